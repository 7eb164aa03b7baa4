"""ORACLE (test infrastructure, not product code) -- numpy restatement of the counter-based random numbers of the
device-resident sampler variant (csrc/ensemble.cuh: stretch_randoms_philox_kernel).

Only ``tests/`` may import this module.

Philox4x32-10 is the published generator of Salmon, Moraes, Dror and Shaw, "Parallel random numbers: as easy as 1, 2, 3"
(SC'11): ten rounds of  (c0, c1, c2, c3) <- (hi(M1 c2) ^ c1 ^ k0, lo(M1 c2), hi(M0 c0) ^ c3 ^ k1, lo(M0 c0))  with the key
bumped by the Weyl constants after every round.  Pinned by the known-answer vectors of the Random123 distribution
(kat_vectors: zero counter/key, all-ones, and the digits-of-pi vector) in tests/test_ensemble.py.

The arrays returned by :func:`stretch_randoms` are what ``bi_stretch_randoms_philox`` writes: counter = (proposal k,
half | draw << 1, step lo, step hi), key = (seed lo, seed hi); u = ((w_hi >> 6) 2^26 + (w_lo >> 6) + 1/2) 2^-52.
"""
from __future__ import annotations

import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = np.uint64(0x9E3779B9), np.uint64(0xBB67AE85)
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised over equal-shaped uint64 arrays holding 32-bit words; returns the four output words."""
    c0, c1, c2, c3 = (np.asarray(v, np.uint64) & MASK for v in (c0, c1, c2, c3))
    k0 = np.asarray(k0, np.uint64) & MASK
    k1 = np.asarray(k1, np.uint64) & MASK
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & MASK, lo1, (hi0 ^ c3 ^ k1) & MASK, lo0
        k0 = (k0 + W0) & MASK
        k1 = (k1 + W1) & MASK
    return c0, c1, c2, c3


def u52(hi, lo):
    return ((hi >> np.uint64(6)).astype(np.float64) * 67108864.0 + (lo >> np.uint64(6)).astype(np.float64) + 0.5) \
        * (1.0 / 4503599627370496.0)


def stretch_randoms(seed: int, step0: int, nsteps: int, nwalkers: int, ndim: int, a: float = 2.0):
    """dict of (nsteps, 2, nwalkers // 2) arrays: s_idx, c_idx (int32), zz, factors, logu."""
    ns = nwalkers // 2
    t, h, k = np.meshgrid(np.arange(nsteps, dtype=np.uint64), np.arange(2, dtype=np.uint64),
                          np.arange(ns, dtype=np.uint64), indexing="ij")
    step = t + np.uint64(step0)
    slo, shi = step & MASK, step >> np.uint64(32)
    k0 = np.full_like(k, np.uint64(seed) & MASK)
    k1 = np.full_like(k, (np.uint64(seed) >> np.uint64(32)) & MASK)
    r0 = philox4x32_10(k, h, slo, shi, k0, k1)
    r1 = philox4x32_10(k, h | np.uint64(2), slo, shi, k0, k1)
    u1, u2, u3 = u52(r0[0], r0[1]), u52(r0[2], r0[3]), u52(r1[0], r1[1])
    tt = (a - 1.0) * u1 + 1.0
    zz = tt * tt / a
    j = np.minimum((u2 * float(ns)).astype(np.int64), ns - 1)
    hi_ = h.astype(np.int64)
    return {"s_idx": (2 * k.astype(np.int64) + hi_).astype(np.int32), "c_idx": (2 * j + (1 - hi_)).astype(np.int32),
            "zz": zz, "factors": (ndim - 1.0) * np.log(zz), "logu": np.log(u3)}
