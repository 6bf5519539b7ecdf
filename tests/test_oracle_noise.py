"""The noise / VMI oracle (oracle/noise_oracle.py) against published known answers and closed forms (CPU only)."""
import numpy as np
from scipy import stats

from oracle import noise_oracle as NO

# Random123 known-answer vectors, philox4x32 with 10 rounds (counter words, key words -> output words)
KAT = [
    ((0x00000000, 0x00000000, 0x00000000, 0x00000000), (0x00000000, 0x00000000),
     (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff), (0xffffffff, 0xffffffff),
     (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
     (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]


def test_philox_known_answers():
    for ctr, key, want in KAT:
        got = tuple(int(x) for x in NO.philox4x32_10(ctr, key))
        assert got == want
    # vectorised call = element-wise calls
    c0 = np.array([0, 0xffffffff, 0x243f6a88], dtype=np.uint32)
    out = NO.philox4x32_10((c0, 0, 0, 0), (0, 0))
    assert int(out[0][0]) == 0x6627e8d5 and out[0].shape == (3,)


def test_uniforms_are_strictly_inside_the_unit_interval():
    u = NO.u01(np.array([0, 1, 0xffffffff], dtype=np.uint32))
    assert u[0] == 2.0 ** -33 and u[2] == 1.0 - 2.0 ** -33 and np.all((u > 0) & (u < 1))


def test_small_means_follow_the_poisson_pmf():
    n = 200000
    for lam in (0.05, 0.7, 3.0, 9.99):
        x = NO.poisson(np.full(n, lam), seed=99, stream_id=3)
        kmax = int(x.max())
        obs = np.bincount(x.astype(int), minlength=kmax + 1)
        exp = stats.poisson.pmf(np.arange(kmax + 1), lam) * n
        keep = exp >= 5
        chi2 = np.sum((obs[keep] - exp[keep]) ** 2 / exp[keep])
        assert chi2 < stats.chi2.ppf(1 - 1e-6, keep.sum() - 1), (lam, chi2)


def test_large_means_follow_the_poisson_pmf():
    n = 200000
    for lam in (10.0, 37.5, 400.0):
        x = NO.poisson(np.full(n, lam), seed=7)
        lo, hi = int(stats.poisson.ppf(1e-4, lam)), int(stats.poisson.ppf(1 - 1e-4, lam))
        obs = np.bincount(np.clip(x.astype(int), lo, hi) - lo, minlength=hi - lo + 1)
        exp = stats.poisson.pmf(np.arange(lo, hi + 1), lam) * n
        exp[0] = stats.poisson.cdf(lo, lam) * n
        exp[-1] = stats.poisson.sf(hi - 1, lam) * n
        keep = exp >= 5
        chi2 = np.sum((obs[keep] - exp[keep]) ** 2 / exp[keep])
        assert chi2 < stats.chi2.ppf(1 - 1e-6, keep.sum() - 1), (lam, chi2)
    x = NO.poisson(np.full(n, 1.0e5), seed=8)
    assert abs(x.mean() - 1e5) < 5 * np.sqrt(1e5 / n) and abs(x.var() / 1e5 - 1) < 0.03


def test_streams_offsets_and_edge_cases():
    lam = np.array([0.0, -3.0, np.nan, 5.0, 5.0e4])
    x = NO.poisson(lam, seed=1)
    assert x[0] == 0 and x[1] == 0 and np.isnan(x[2]) and x[3] >= 0 and x[4] > 4.0e4
    assert NO.poisson(np.zeros((0,)), seed=1).shape == (0,)
    big = np.full(1000, 20.0)
    whole = NO.poisson(big, seed=5, stream_id=2)
    # a shard starting at element 400 draws what the whole array drew there
    np.testing.assert_array_equal(NO.poisson(big[400:], seed=5, stream_id=2, offset=400), whole[400:])
    assert np.any(NO.poisson(big, seed=5, stream_id=3) != whole) and np.any(NO.poisson(big, seed=6, stream_id=2) != whole)
    # counters beyond 2^32 use the high word
    assert np.any(NO.poisson(big[:8], seed=5, offset=1 << 32) != NO.poisson(big[:8], seed=5, offset=0))


def test_make_vmi_matches_the_reference_expression():
    r = np.random.default_rng(3)
    m1, m2 = r.uniform(0, 1.2, (16, 16)).astype(np.float32), r.uniform(0, 0.4, (16, 16)).astype(np.float32)
    u1, u2, uw = np.array([0.2048]), np.array([0.2229]), np.array([0.2059])
    hu = NO.make_vmi(70, m1, m2, u1, u2, uw)
    assert hu.dtype == np.float32
    np.testing.assert_allclose(hu, 1000 * ((0.2048 * m1.astype(float) + 0.2229 * m2.astype(float)) - 0.2059) / 0.2059, rtol=1e-6, atol=1e-3)
    raw = NO.make_vmi(70, m1, m2, u1, u2, uw, HU=False)
    np.testing.assert_allclose(raw, 0.2048 * m1 + 0.2229 * m2, rtol=1e-6)
