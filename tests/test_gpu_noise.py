"""Poisson noise and VMI kernels (csrc/noise.cu, through the C ABI) against oracle/noise_oracle.py.

The generator is integer arithmetic and the sampler's accept/reject decisions are written operation for operation like
the oracle's, so variates are compared ELEMENT FOR ELEMENT.  Only the slow acceptance branch evaluates log/lgamma, whose
last bit may differ between CUDA's and the host's libm; a decision flips only if the two sides of the test agree to
~1e-15, so the test allows 1 differing element per 10^5 (none has been observed).
"""
import numpy as np
import pytest
import torch

import xcompy as xc
import matdecomp
from dexct import config, dist as ddist
from dexct import synthetic as S
from dexct.ops import F32, F64, poisson_noise, vmi
from oracle import noise_oracle as NO
from xtomosim import forward_project as FP

pytestmark = pytest.mark.gpu
TD = {F32: torch.float32, F64: torch.float64}


def _means(n, seed):
    r = np.random.default_rng(seed)
    lam = np.concatenate([r.uniform(0, 10, n // 4), r.uniform(10, 200, n // 4), 10 ** r.uniform(2, 6.5, n // 4),
                          r.uniform(0, 1e5, n - 3 * (n // 4))])
    lam[:6] = [0.0, -1.0, 9.999999, 10.0, 1e-30, 3.0e6]
    return lam


@pytest.mark.parametrize('code', [F32, F64])
def test_variates_equal_the_oracle_element_for_element(code):
    n = 300000
    lam = _means(n, 1)
    if code == F32:
        lam = lam.astype(np.float32).astype(np.float64)
    t = torch.as_tensor(lam, dtype=TD[code], device='cuda')
    lg = torch.empty_like(t)
    poisson_noise(t, seed=0x1234567890ABCDEF, stream_id=5, offset=(1 << 33) + 17, logs=lg, air_counts=1.0e5, floor_counts=0.5,
                  dtype=code)
    want = NO.poisson(lam, seed=0x1234567890ABCDEF, stream_id=5, offset=(1 << 33) + 17)
    got = t.cpu().numpy().astype(np.float64)
    if code == F32:
        want = want.astype(np.float32).astype(np.float64)       # counts above 2^24 round to float
    diff = np.flatnonzero(got != want)
    assert diff.size <= n // 100000, (diff[:10], got[diff[:10]], want[diff[:10]])
    want_log = NO.noisy_log(got, 1.0e5, 0.5)
    np.testing.assert_allclose(lg.cpu().numpy(), want_log, rtol=2e-6 if code == F32 else 1e-14, atol=1e-6 if code == F32 else 1e-14)


def test_nan_and_empty():
    t = torch.tensor([float('nan'), 0.0, 4.0], device='cuda')
    poisson_noise(t, seed=1)
    h = t.cpu().numpy()
    assert np.isnan(h[0]) and h[1] == 0 and h[2] >= 0
    poisson_noise(torch.empty(0, device='cuda'), seed=1)
    with pytest.raises(ValueError):
        poisson_noise(torch.ones(4, device='cuda', dtype=torch.float64), seed=1, dtype=F32)


def test_view_shards_draw_what_the_whole_scan_draws():
    lam = torch.full((40, 1, 50), 300.0, device='cuda')
    whole = poisson_noise(lam.clone(), seed=9, stream_id=1)
    parts = []
    for rank in range(3):
        v0, nv = ddist.shard_range(40, rank, 3)
        parts.append(poisson_noise(lam[v0:v0 + nv].clone(), seed=9, stream_id=1, offset=v0 * 50))
    assert torch.equal(torch.cat(parts), whole)


def test_get_sino_with_noise_is_poisson_around_the_noiseless_scan(monkeypatch):
    # every call re-traces (no stored lengths), so the noiseless counts under the noise are bit-identical to clean_raw
    monkeypatch.setattr(config, 'cache_lengths', False)
    ct, ph, lo, hi = S.make_config('C2', n_views=64, n=128)
    FP.clear_cache()
    clean_raw, clean_log = FP.get_sino(ct, ph, lo)
    ct.noise, ct.noise_seed = True, 77
    raw1, log1 = FP.get_sino(ct, ph, lo)
    raw2, _ = FP.get_sino(ct, ph, lo)                       # the next acquisition takes the next random stream
    assert ct.noise_stream == 2
    assert np.all(raw1 == np.round(raw1)) and np.any(raw1 != raw2)
    ok = clean_raw > 20.0                                   # (rays through metal detect next to nothing)
    z = ((raw1 - clean_raw) / np.sqrt(np.maximum(clean_raw, 1.0)))[ok]
    assert abs(z.mean()) < 0.02 and abs(z.std() - 1.0) < 0.02
    # element for element: the oracle's variates of the (float32) noiseless counts
    want = NO.poisson(clean_raw.astype(np.float32).astype(np.float64), seed=77, stream_id=0)
    assert np.mean(want != raw1) < 1e-5
    air = float(np.sum(FP.spectral_weights(ct, lo)))
    np.testing.assert_allclose(log1, NO.noisy_log(raw1, air, config.noise_floor), rtol=1e-5, atol=1e-5)
    # dual-spectrum call: two streams, both noisy; same seed and stream numbering -> reproducible
    ct.noise_stream = 0
    (r_lo, _), (r_hi, _) = FP.get_sinos(ct, ph, [lo, hi])
    assert np.array_equal(r_lo, raw1) and np.all(r_hi == np.round(r_hi)) and ct.noise_stream == 2
    ct.noise = False
    again, _ = FP.get_sino(ct, ph, lo)
    assert np.array_equal(again, clean_raw)


@pytest.mark.parametrize('code', [F32, F64])
def test_vmi_matches_the_reference_expression(code):
    r = np.random.default_rng(5)
    m1 = r.uniform(-0.1, 1.2, (64, 64)).astype(np.float32)
    m2 = r.uniform(-0.1, 0.5, (64, 64)).astype(np.float32)
    E = np.array([70.0])
    u1, u2 = xc.mixatten(matdecomp.matcomp1, E), xc.mixatten(matdecomp.matcomp2, E)
    uw = 1.0 * xc.mixatten('H(11.2)O(88.8)', E)
    for hu in (True, False):
        want = NO.make_vmi(70.0, m1, m2, u1, u2, uw, HU=hu)
        got = vmi(torch.as_tensor(m1, dtype=TD[code], device='cuda'), torch.as_tensor(m2, dtype=TD[code], device='cuda'),
                  float(u1[0]), float(u2[0]), float(uw[0]), hu=hu, dtype=code).cpu().numpy()
        assert got.dtype == np.float32
        if code == F64:
            np.testing.assert_allclose(got, want, rtol=2e-7, atol=1e-30)        # one float32 rounding of the same double
        else:
            np.testing.assert_allclose(got, want, rtol=1e-5, atol=2e-3 if hu else 1e-6)
    config_dtype = config.dtype
    try:
        config.dtype = 'f64' if code == F64 else 'f32'
        got = matdecomp.make_vmi(70.0, m1, m2)
        np.testing.assert_allclose(got, NO.make_vmi(70.0, m1, m2, u1, u2, uw), rtol=1e-5, atol=2e-3)
    finally:
        config.dtype = config_dtype
