"""Sustained (seconds-long, power-capped) GEMM throughput on one box: our tcgen05 kernels vs cuBLAS on the four
Falcon-7B layer shapes at the batch size of the real workload, with SM clock and power sampled during each run."""
import json
import subprocess
import sys
import time

import torch

sys.path.insert(0, ".")
from llm_interpretation_replication_b200 import _lib

SHAPES = ((4672, 4544, 0), (4544, 4544, 2), (18176, 4544, 1), (4544, 18176, 2))


def sample_start():
    return subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-lms", "100"],
                            stdout=subprocess.PIPE, text=True)


def sample_stop(p):
    p.terminate()
    out = p.communicate()[0]
    clk, pw = [], []
    for line in out.splitlines():
        a = line.split(",")
        try:
            clk.append(float(a[0])); pw.append(float(a[1]))
        except Exception:
            pass
    k = len(clk) // 3            # ignore the ramp
    return (sorted(clk[k:])[len(clk[k:]) // 2] if clk[k:] else None, sorted(pw[k:])[len(pw[k:]) // 2] if pw[k:] else None)


def main():
    M = int(sys.argv[1]) if len(sys.argv) > 1 else 111104
    secs = float(sys.argv[2]) if len(sys.argv) > 2 else 3.0
    lib = _lib.load()
    st = _lib.current_stream()
    bufs = []
    for (N, K, epi) in SHAPES:
        for dt in (torch.bfloat16, torch.float16):
            A = torch.randn(M, K, device="cuda").to(dt)
            W = (torch.randn(N, K, device="cuda") * 0.02).to(dt)
            out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if epi == 2 else dt)
            bufs.append((N, K, epi, dt, A, W, out))
    fl_layer = sum(2.0 * M * N * K for (N, K, _) in SHAPES)

    def run(name, fn):
        fn(); torch.cuda.synchronize()
        p = sample_start()
        t0 = time.time(); n = 0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        while time.time() - t0 < secs:
            fn(); n += 1
            if n % 4 == 0:
                torch.cuda.synchronize()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        clk, pw = sample_stop(p)
        print(json.dumps({"impl": name, "tflops": round(n * fl_layer / ms / 1e9, 1), "layers": n, "sm_mhz": clk, "power_w": pw}), flush=True)

    def ours(dt, cg):
        fnname = lib.b2s_gemm_bf16 if dt == torch.bfloat16 else lib.b2s_gemm_f16
        def f():
            _lib.check(lib.b2s_set_option(b"gemm_cta_group", cg), "opt")
            for (N, K, epi, d, A, W, out) in bufs:
                if d == dt:
                    _lib.check(fnname(_lib.ptr(A), _lib.ptr(W), None, _lib.ptr(out), M, N, K, epi, st), "gemm")
        return f

    def cublas(dt):
        def f():
            for (N, K, epi, d, A, W, out) in bufs:
                if d == dt:
                    A @ W.T
        return f
    for rnd in range(2):
        run("cublas_bf16", cublas(torch.bfloat16))
        run("ours_pair_bf16", ours(torch.bfloat16, 2))
        run("ours_cta1_bf16", ours(torch.bfloat16, 1))
        run("ours_pair_fp16", ours(torch.float16, 2))
        run("cublas_fp16", cublas(torch.float16))
    _lib.check(lib.b2s_set_option(b"gemm_cta_group", 0), "opt")


if __name__ == "__main__":
    main()
