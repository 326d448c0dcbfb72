// Allocation-free variant of the fused LM head (b2s_lm_head_bf16_static).
//
// Who this is for.  b2s_lm_head_bf16 needs a scratch buffer for its per-vocab-tile partial records, and the
// caller has to get that buffer from somewhere.  Some callers cannot: a plug-in that runs inside a host
// application whose device memory is entirely owned by that application's own pool (a raw cudaMalloc from the
// plug-in either fails because the pool already holds the whole GPU, or fragments what the pool thinks it owns);
// a serving loop that must not take the device-wide lock cudaMalloc/cudaFree take (they synchronise the
// device and stall every other stream); and code that is only handed input and output pointers by a foreign
// runtime and has no allocator handle at all.  For them the scratch lives in a fixed array that is part of
// this library's image: it is reserved on every GPU that loads the library, for as long as the process lives,
// so the entry point itself never allocates, frees or resizes anything.
#include "b2s_common.cuh"
#include <mutex>

namespace b2s {

// Capacity: what b2s_lm_head_scratch_bytes(M, V) reports must fit.  4 MiB + 4 KiB covers e.g. 256 rows at a
// 131072-entry vocabulary, or 512 rows at Falcon's 65024 (254 tiles * 8 floats * 4 B = 8128 B per row).
constexpr size_t NOALLOC_SCRATCH_BYTES = (4u << 20) + 4096;

// File-scope device array.  In host code this name is only a handle for the runtime's symbol table, not
// something a kernel can dereference, so the device address is looked up once per device below.
static __device__ __align__(256) unsigned char g_lmhead_static_scratch[NOALLOC_SCRATCH_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
static void* g_scratch_addr[NOALLOC_MAX_DEVICES] = {nullptr};
static std::mutex g_scratch_mutex;

// First use on a device: resolve the array's address on THAT device (the device must be current, each GPU has
// its own copy of the array).  Plain symbol lookup -- nothing is allocated.
static int scratch_for_current_device(void** out) {
  int dev = -1;
  B2S_CHECK_CUDA(cudaGetDevice(&dev));
  B2S_REQUIRE(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "lm_head_static: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (g_scratch_addr[dev] == nullptr) {
    void* p = nullptr;
    B2S_CHECK_CUDA(cudaGetSymbolAddress(&p, g_lmhead_static_scratch));
    B2S_REQUIRE(p != nullptr, "lm_head_static: cudaGetSymbolAddress returned a null address on device %d", dev);
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return B2S_OK;
}

int launch_lmhead_merge(const float* partials, const float* cand_logits, int M, int tiles_n, float* stats, cudaStream_t st);

}  // namespace b2s

using namespace b2s;

extern "C" {

size_t b2s_lm_head_static_capacity(void) { return NOALLOC_SCRATCH_BYTES; }

int b2s_lm_head_bf16_static(const void* hidden, const void* W, const int32_t* cand_ids, int32_t M, int32_t V, int32_t d,
                            float* stats, void* stream) {
  B2S_REQUIRE(hidden && W && cand_ids && stats, "lm_head_static: null argument");
  B2S_REQUIRE(M > 0 && V > 0 && d > 0, "lm_head_static: empty problem");
  const size_t tiles_n = static_cast<size_t>(ceil_div(V, LMH_TILE_N));
  const size_t part_bytes = align_up(4 * static_cast<size_t>(M) * tiles_n * LMH_PART, 256);
  const size_t need = part_bytes + align_up(8 * static_cast<size_t>(M), 256);
  B2S_REQUIRE(need <= NOALLOC_SCRATCH_BYTES,
              "lm_head_static: M=%d rows x V=%d needs %zu bytes of scratch, the built-in buffer holds %zu; "
              "use b2s_lm_head_bf16 with caller-provided scratch, or split the rows",
              M, V, need, NOALLOC_SCRATCH_BYTES);
  void* base = nullptr;
  B2S_TRY(scratch_for_current_device(&base));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  float* partials = reinterpret_cast<float*>(base);
  float* cand_logits = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(base) + part_bytes);
  // The array keeps whatever the previous call left in it.  Partials are fully rewritten by the GEMM epilogue;
  // the candidate-logit slots are only written by the tile that owns the candidate column, so they are cleared.
  B2S_CHECK_CUDA(cudaMemsetAsync(cand_logits, 0, 8 * static_cast<size_t>(M), st));
  LmHeadArgs a{cand_ids, partials, cand_logits};
  B2S_TRY(gemm_tc_launch(hidden, W, nullptr, nullptr, M, V, d, EPI_LMHEAD, &a, st));
  return launch_lmhead_merge(partials, cand_logits, M, static_cast<int>(tiles_n), stats, st);
}

}  // extern "C"
