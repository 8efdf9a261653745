"""Philox4x32-10 known-answer vectors (Random123 kat_vectors) for the three CPU-side generators;
the CUDA generator is checked against these through the jittered renders in the gpu tests."""
import numpy as np

import oracle
import philox

KAT = [
    ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff,) * 4, (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]


def test_numpy_philox_kat():
    for ctr, key, want in KAT:
        got = tuple(int(x) for x in philox.philox4x32_10(*ctr, *key))
        assert got == want


def test_c_oracle_philox_matches_numpy():
    # the oracle's entry point fixes counter words 2,3 to zero (the jitter stream layout)
    for idx, seed in [(0, 0), (12345, 0x5EED), (2**40 + 17, 0xDEADBEEF12345678), (2**63 + 5, 1)]:
        got = oracle.philox(idx, seed)
        want = philox.philox4x32_10(idx & 0xffffffff, idx >> 32, 0, 0, seed & 0xffffffff, seed >> 32)
        assert got == tuple(int(x) for x in want)


def test_rand_stream_order():
    rs = philox.RandStream(first_idx=7, seed=0x5EED, chunk=8)
    draws = [rs() for _ in range(30)]
    want = philox.jitter_uniforms(np.arange(7, 17, dtype=np.uint64), 0x5EED).reshape(-1)
    assert draws == list(want)
    assert all(0.0 <= x < 1.0 for x in draws)
