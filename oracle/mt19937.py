"""MT19937 + NumPy legacy bounded integers, restated (test infrastructure only).

Follows numpy's legacy ``RandomState`` as used by every seeded hot-path bootstrap of the
reference (``np.random.seed(42)`` + ``np.random.choice(n, size=n, replace=True)``:
reference analysis/run_base_vs_instruct_100q.py:616,657;
survey_analysis/analyze_base_vs_instruct_mae_100q.py:296,312,357,382;
analysis/analyze_results_base_versus_instruct.py:304,312,349).  Spec: SURVEY.md Appendix B.
Pinned against numpy itself in tests/test_oracle_bootstrap.py.
"""
from __future__ import annotations

import numpy as np

N, M = 624, 397
MATRIX_A = 0x9908B0DF
UPPER, LOWER = 0x80000000, 0x7FFFFFFF


class MT19937:
    def __init__(self, seed: int):
        mt = np.empty(N, dtype=np.uint64)
        mt[0] = seed & 0xFFFFFFFF
        for i in range(1, N):
            mt[i] = (1812433253 * (int(mt[i - 1]) ^ (int(mt[i - 1]) >> 30)) + i) & 0xFFFFFFFF
        self.mt = mt.astype(np.uint32)
        self.pos = N

    def _twist(self):
        mt = self.mt.astype(np.uint64)
        for kk in range(N):
            y = (int(mt[kk]) & UPPER) | (int(mt[(kk + 1) % N]) & LOWER)
            v = int(mt[(kk + M) % N]) ^ (y >> 1) ^ (MATRIX_A if (y & 1) else 0)
            mt[kk] = v
        self.mt = mt.astype(np.uint32)
        self.pos = 0

    def next_u32(self) -> int:
        if self.pos >= N:
            self._twist()
        y = int(self.mt[self.pos])
        self.pos += 1
        y ^= y >> 11
        y ^= (y << 7) & 0x9D2C5680
        y ^= (y << 15) & 0xEFC60000
        y ^= y >> 18
        return y & 0xFFFFFFFF

    def words(self, count: int) -> np.ndarray:
        return np.array([self.next_u32() for _ in range(count)], dtype=np.uint32)

    def bounded(self, rng: int) -> int:
        """legacy rk_interval: masked rejection, one 32-bit word per attempt (rng < 2**32)."""
        if rng == 0:
            return 0
        mask = rng
        mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16
        while True:
            v = self.next_u32() & mask
            if v <= rng:
                return v

    def choice(self, n: int, size: int) -> np.ndarray:
        """np.random.choice(n, size=size, replace=True) == randint(0, n, size)."""
        return np.array([self.bounded(n - 1) for _ in range(size)], dtype=np.int64)

    def state_words(self) -> np.ndarray:
        """624 state words + position, the layout b2s_mt19937_choice reads/writes."""
        return np.concatenate([self.mt, np.array([self.pos], dtype=np.uint32)])


def choice_table_numpy(seed: int, n: int, B: int) -> np.ndarray:
    """The same [B, n] index table via numpy's own legacy stream (fast path for big cases)."""
    rs = np.random.RandomState(seed)
    return np.stack([rs.choice(n, size=n, replace=True) for _ in range(B)]).astype(np.int32)
