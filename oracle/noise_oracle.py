"""ORACLE (test infrastructure only) -- quantum noise and virtual monoenergetic images, SURVEY.md section 8 row f4.

Parity status: the reference's noise sampler sits in the absent ``xtomosim`` submodule (call sites:
/root/reference/main.py:68 ``spec.rescale_counts(A_iso * dose / N_proj)`` and the noise study at plots.py:631-717), so the
*sampler* is project-defined and its parity is UNPINNED against the reference.  What is pinned:

* ``philox4x32_10`` against the published known-answer vectors of the Random123 library (Salmon, Moraes, Dror, Shaw,
  "Parallel random numbers: as easy as 1, 2, 3", SC'11; kat_vectors file, philox4x32 10 rounds) -- see
  tests/test_oracle_noise.py;
* ``poisson`` statistically (exact pmf for small means, mean/variance for large ones);
* ``make_vmi`` against the reference function itself (/root/reference/plots.py:136-144), restated below.

The CUDA kernels (dex-ct-sim_b200/csrc/noise.cu) use the same counter layout and the same operation order, so the test
compares them element for element.
"""
from __future__ import annotations

import numpy as np
from scipy.special import gammaln

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)
S32 = np.uint64(32)


def philox4x32_10(ctr, key):
    """ctr: four uint32 arrays (broadcastable), key: two uint32 scalars -> four uint32 arrays."""
    c = [np.asarray(x, dtype=np.uint64) & MASK for x in ctr]
    c = list(np.broadcast_arrays(*c))
    k0, k1 = int(key[0]) & 0xFFFFFFFF, int(key[1]) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c[0]
        p1 = M1 * c[2]
        hi0, lo0 = p0 >> S32, p0 & MASK
        hi1, lo1 = p1 >> S32, p1 & MASK
        c = [hi1 ^ c[1] ^ np.uint64(k0), lo1, hi0 ^ c[3] ^ np.uint64(k1), lo0]
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return [x.astype(np.uint32) for x in c]


def u01(x):
    """uint32 -> double in (0, 1): (x + 1/2) 2^-32 (exact)."""
    return (x.astype(np.float64) + 0.5) * 2.3283064365386963e-10


def poisson(lam, seed, stream_id=0, offset=0):
    """Poisson variates of means ``lam`` (any shape; element i of the flattened array uses counter ``offset + i``).

    lam < 10: CDF inversion by sequential search from one uniform; otherwise PTRS (Hoermann 1993), two uniforms per trial.
    Counter = (index low, index high, block, stream_id), key = (seed low, seed high)."""
    lam = np.asarray(lam, dtype=np.float64)
    flat = lam.reshape(-1)
    n = flat.size
    idx = np.uint64(offset) + np.arange(n, dtype=np.uint64)
    c0, c1 = idx & MASK, idx >> S32
    key = (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    out = np.zeros(n, dtype=np.float64)
    out[np.isnan(flat)] = np.nan

    small = np.flatnonzero((flat > 0) & (flat < 10.0))
    if small.size:
        l = flat[small]
        u = u01(philox4x32_10((c0[small], c1[small], 0, stream_id), key)[0])
        p = np.exp(-l)
        F = p.copy()
        k = np.zeros(small.size)
        live = u > F
        it = 0
        while live.any() and it < 200:
            it += 1
            k[live] += 1.0
            p[live] = p[live] * (l[live] / k[live])
            F[live] = F[live] + p[live]
            live &= u > F
        out[small] = k

    big = np.flatnonzero(flat >= 10.0)
    if big.size:
        l = flat[big]
        slam, loglam = np.sqrt(l), np.log(l)
        b = 0.931 + 2.53 * slam
        a = -0.059 + 0.02483 * b
        invalpha = 1.1239 + 1.1328 / (b - 3.4)
        vr = 0.9277 - 3.6224 / (b - 2.0)
        res = np.floor(l + 0.5)
        todo = np.ones(big.size, dtype=bool)
        blk = None
        for t in range(64):
            if not todo.any():
                break
            if t % 2 == 0:
                blk = philox4x32_10((c0[big], c1[big], t >> 1, stream_id), key)
            U = u01(blk[2 * (t & 1)]) - 0.5
            V = u01(blk[2 * (t & 1) + 1])
            us = 0.5 - np.abs(U)
            k = np.floor((2.0 * a / us + b) * U + l + 0.43)
            fast = (us >= 0.07) & (V <= vr)
            skip = ~fast & ((k < 0.0) | ((us < 0.013) & (V > us)))
            with np.errstate(invalid='ignore', divide='ignore'):
                lhs = np.log(V) + np.log(invalpha) - np.log(a / (us * us) + b)
                rhs = -l + k * loglam - gammaln(k + 1.0)
            acc = todo & (fast | (~skip & (lhs <= rhs)))
            res[acc] = k[acc]
            todo &= ~acc
        out[big] = res
    return out.reshape(lam.shape)


def noisy_log(counts, air_counts, floor_counts=0.5):
    return -np.log(np.maximum(counts, floor_counts) / air_counts)


def make_vmi(E0, M1_img, M2_img, u_p_1, u_p_2, u_w, HU=True):
    """/root/reference/plots.py:136-144 with the three xcompy lookups passed in (u_p_1, u_p_2, u_w: mass attenuation of
    the two bases and of water at E0)."""
    vmi = u_p_1 * M1_img + u_p_2 * M2_img
    if HU:
        vmi = 1000 * (vmi - u_w) / u_w
    return vmi.astype(np.float32)
