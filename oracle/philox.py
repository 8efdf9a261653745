"""Philox4x32-10 (Salmon et al., SC'11 "Parallel random numbers: as easy as 1, 2, 3"), numpy.

TEST INFRASTRUCTURE ONLY.  The product kernels draw the jitter of `provided/scene.py:63-65`
from this counter-based generator; the oracle tiers read the SAME stream so that both
sample identically (BASELINE.json north_star):

    counter = (idx_lo, idx_hi, 0, 0),  idx = ((col*H + row)*D + dof)*S + aa
    key     = (seed_lo, seed_hi)
    r1, r2, r3 = first three output words * 2**-32   (doubles in [0, 1))

Known-answer vectors (Random123 kat_vectors) are checked in tests/test_philox.py.
"""
import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)
DEFAULT_SEED = 0x5EED


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised over numpy uint32 arrays (counters); k0/k1 scalar ints."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) & MASK for c in (c0, c1, c2, c3))
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)), lo1, (hi0 ^ c3 ^ np.uint64(k1)), lo0
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return tuple(c.astype(np.uint32) for c in (c0, c1, c2, c3))


def jitter_uniforms(idx, seed=DEFAULT_SEED):
    """(n,3) float64 uniforms in [0,1) for sample indices `idx` (uint64 array)."""
    idx = np.asarray(idx, dtype=np.uint64)
    w = philox4x32_10(idx & MASK, idx >> np.uint64(32), 0, 0, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    return np.stack([w[0], w[1], w[2]], axis=-1).astype(np.float64) * 2.0 ** -32


class RandStream:
    """Replacement for `numpy.random.rand` in the reference: the n-th scalar draw of a strip
    render that starts at column `col0` is component n%3 of sample index n//3 + first_idx,
    because `scene.py:47-65` draws three numbers per (col,row,dof,aa) in exactly that order."""

    def __init__(self, first_idx, seed=DEFAULT_SEED, chunk=1 << 16):
        self.next_idx = int(first_idx)
        self.seed = seed
        self.chunk = chunk
        self.buf = []
        self.pos = 0

    def __call__(self, *shape):
        if shape:
            raise NotImplementedError("reference only calls np.random.rand() without arguments")
        if self.pos >= len(self.buf):
            idx = np.arange(self.next_idx, self.next_idx + self.chunk, dtype=np.uint64)
            self.buf = jitter_uniforms(idx, self.seed).reshape(-1).tolist()
            self.pos = 0
            self.next_idx += self.chunk
        v = self.buf[self.pos]
        self.pos += 1
        return v
