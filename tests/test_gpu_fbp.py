"""Parity of the CUDA fan-beam FBP (csrc/fbp.cu, through the C ABI) with the float64 oracle (oracle/fbp_oracle.py).

Tolerances (scale-relative, max|a - ref| / max|ref|): 1e-10 in f64 mode; 2e-5 in f32 mode -- the back-projection sums
~10^3 interpolated samples of a filtered line whose neighbouring samples differ by O(1), and the single-precision
channel coordinate (atan, 2^-24 relative) moves each sample by ~1e-5 of that difference.
"""
import os

import numpy as np
import pytest
import torch

from conftest import scale_rel_err
from dexct import _native as N
from dexct import synthetic as S
from dexct.ops import F32, F64, fbp_fan, hounsfield, make_scan
from oracle import fbp_oracle as FO

pytestmark = pytest.mark.gpu

FAN, SID, SDD = 0.8230337, 60.0, 100.0
TOL = {F32: 2e-5, F64: 1e-10}
TD = {F32: torch.float32, F64: torch.float64}


def _case(n_ch=200, n_views=120, n_rows=1, seed=11, beta0=0.0):
    scan = make_scan(n_views, n_ch, FAN, SID, SDD, n_rows=n_rows, h_iso=0.3, beta0=beta0)
    beta = scan.beta0 + np.arange(n_views) * scan.dbeta
    gam = scan.gamma0 + np.arange(n_ch) * scan.dgamma
    r = np.random.default_rng(seed)
    rows = []
    for k in range(n_rows):
        m = k % 5                                         # five distinct slices, repeated
        s = FO.disc_sinogram(beta, gam, SID, (3.0 - m, -2.0 + 0.5 * m), 8.0 - m, 0.2)
        s += FO.disc_sinogram(beta, gam, SID, (-5.0, 4.0), 2.0, 0.35)
        s += 0.01 * r.standard_normal(s.shape)            # noise: exercises the filter at every frequency
        rows.append(s)
    sino = np.stack(rows, axis=1)                          # [views, rows, ch]
    return scan, beta, sino


def _oracle(scan, beta, sino, n, fov, ramp):
    return FO.fbp_fan(sino, beta, scan.gamma0, scan.dgamma, scan.dbeta, SID, n, fov, ramp)


@pytest.mark.parametrize('dtype', [F32, F64])
@pytest.mark.parametrize('ramp', [1.0, 0.8, 0.5])
def test_image_matches_oracle(dtype, ramp):
    scan, beta, sino = _case()
    ref = _oracle(scan, beta, sino, 96, 40.0, ramp)
    got = fbp_fan(scan, torch.as_tensor(sino, device='cuda'), 96, 40.0, ramp, dtype=dtype)
    assert got.shape == (1, 96, 96) and got.dtype == TD[dtype]
    assert scale_rel_err(got.cpu().numpy(), ref) < TOL[dtype]


@pytest.mark.parametrize('dtype', [F32, F64])
def test_multi_row_odd_sizes_and_start_angle(dtype):
    scan, beta, sino = _case(n_ch=173, n_views=77, n_rows=3, beta0=0.3)
    ref = _oracle(scan, beta, sino, 61, 33.0, 0.9)
    got = fbp_fan(scan, torch.as_tensor(sino, device='cuda'), 61, 33.0, 0.9, dtype=dtype)
    assert got.shape == (3, 61, 61)
    assert scale_rel_err(got.cpu().numpy(), ref) < TOL[dtype]


@pytest.mark.parametrize('dtype', [F32, F64])
@pytest.mark.parametrize('n_rows', [8, 9, 16, 21, 24, 40, 50, 100])
def test_many_rows_share_the_in_plane_geometry(dtype, n_rows):
    """n_rows >= 8 takes the multi-row kernels (row-fastest filtered data): 8 or 16 rows per thread up to 23 rows, from 24
    rows on lanes along the rows with 1, 2 or 4 rows per lane; 9, 21, 40, 50 and 100 leave a partly filled last row block.  Every slice must still match the oracle's slice-by-slice reconstruction, and view
    chunks must accumulate."""
    scan, beta, sino = _case(n_ch=131, n_views=45, n_rows=n_rows, beta0=-0.2)
    ref = _oracle(scan, beta, sino, 50, 30.0, 0.8)
    d = torch.as_tensor(sino, device='cuda')
    got = fbp_fan(scan, d, 50, 30.0, 0.8, dtype=dtype)
    assert got.shape == (n_rows, 50, 50)
    err = [scale_rel_err(got[k].cpu().numpy(), ref[k]) for k in range(n_rows)]
    assert max(err) < TOL[dtype], err
    img = torch.zeros_like(got)
    for v0, v1 in ((0, 20), (20, 45)):
        fbp_fan(scan, d[v0:v1], 50, 30.0, 0.8, view_begin=v0, dtype=dtype, image=img, accumulate=True)
    assert scale_rel_err(img.cpu().numpy(), got.cpu().numpy()) < (1e-6 if dtype == F32 else 1e-13)


@pytest.mark.parametrize('dtype', [F32, F64])
@pytest.mark.parametrize('rotation', [2.0 * np.pi, -2.0 * np.pi])
@pytest.mark.parametrize('n_rows', [1, 3])
def test_quarter_turn_symmetry_path(dtype, rotation, n_rows, monkeypatch):
    """Whole full-rotation scans with 4 | n_views and an even image take the kernel that shares each (pixel, view) geometry
    between the four quarter-turn images of the pixel: it must match the oracle and the general kernel, for either sense of
    rotation and any start angle, also when accumulating."""
    n_views, n_ch, n = 80, 151, 64
    scan = make_scan(n_views, n_ch, FAN, SID, SDD, n_rows=n_rows, h_iso=0.3, rotation=rotation, beta0=0.37)
    beta = scan.beta0 + np.arange(n_views) * scan.dbeta
    gam = scan.gamma0 + np.arange(n_ch) * scan.dgamma
    r = np.random.default_rng(4)
    sino = np.stack([FO.disc_sinogram(beta, gam, SID, (4.0 - k, -3.0 + 2 * k), 6.0, 0.2) + 0.01 * r.standard_normal((n_views, n_ch))
                     for k in range(n_rows)], axis=1)
    ref = FO.fbp_fan(sino, beta, scan.gamma0, scan.dgamma, scan.dbeta, SID, n, 38.0, 0.8)
    d = torch.as_tensor(sino, device='cuda')
    got = fbp_fan(scan, d, n, 38.0, 0.8, dtype=dtype)
    assert scale_rel_err(got.cpu().numpy(), ref) < TOL[dtype]
    monkeypatch.setenv('DEXCT_FBP_SYM', '0')
    plain = fbp_fan(scan, d, n, 38.0, 0.8, dtype=dtype)
    monkeypatch.delenv('DEXCT_FBP_SYM')
    assert not torch.equal(plain, got)                       # (two kernels: the sums run in different orders)
    assert scale_rel_err(got.cpu().numpy(), plain.cpu().numpy()) < (2e-6 if dtype == F32 else 1e-13)
    img = got.clone()
    fbp_fan(scan, d, n, 38.0, 0.8, dtype=dtype, image=img, accumulate=True)
    assert scale_rel_err(img.cpu().numpy(), 2.0 * got.cpu().numpy()) < 1e-6


def test_view_chunks_accumulate_to_the_whole_scan():
    """What view-sharded ranks do: partial images of disjoint view ranges sum to the image of the whole scan."""
    scan, beta, sino = _case(n_views=90)
    d = torch.as_tensor(sino, device='cuda')
    whole = fbp_fan(scan, d, 64, 40.0, 0.8, dtype=F64)
    img = torch.zeros_like(whole)
    for v0, v1 in ((0, 31), (31, 31), (31, 90)):            # includes an empty range
        fbp_fan(scan, d[v0:v1], 64, 40.0, 0.8, view_begin=v0, dtype=F64, image=img, accumulate=True)
    assert scale_rel_err(img.cpu().numpy(), whole.cpu().numpy()) < 1e-13
    parts = [fbp_fan(scan, d[a:b], 64, 40.0, 0.8, view_begin=a, dtype=F64) for a, b in ((0, 45), (45, 90))]
    assert scale_rel_err((parts[0] + parts[1]).cpu().numpy(), whole.cpu().numpy()) < 1e-13


def test_hounsfield_and_linearity():
    scan, beta, sino = _case()
    d = torch.as_tensor(sino, device='cuda')
    a = fbp_fan(scan, d, 48, 40.0, 1.0, dtype=F64)
    b = fbp_fan(scan, 2.5 * d, 48, 40.0, 1.0, dtype=F64)
    assert scale_rel_err(b.cpu().numpy(), 2.5 * a.cpu().numpy()) < 1e-14
    hu = hounsfield(a, 0.19, dtype=F64)
    assert np.allclose(hu.cpu().numpy(), FO.hounsfield(a.cpu().numpy(), 0.19), rtol=1e-13, atol=1e-10)


def test_argument_errors_are_reported():
    scan, beta, sino = _case(n_ch=64, n_views=16)
    d = torch.as_tensor(sino, device='cuda')
    with pytest.raises(N.DexctError, match='ramp'):
        fbp_fan(scan, d, 32, 40.0, 0.0)
    with pytest.raises(N.DexctError, match='source circle'):
        fbp_fan(scan, d, 32, 90.0, 1.0)
    with pytest.raises(N.DexctError, match='view range'):
        fbp_fan(scan, d, 32, 40.0, 1.0, view_begin=8)
    with pytest.raises(ValueError):
        fbp_fan(scan, d[:, :, :60], 32, 40.0, 1.0)


def test_c2_pipeline_get_sino_then_get_recon_recovers_the_phantom():
    """Full size (C2: 900 ch x 1000 views -> 512 x 512): project the synthetic phantom with a narrow spectrum and
    reconstruct sino_log through the drop-in entry points; the soft-tissue body and the air around it come back at their
    attenuation, and the f32 image agrees with the oracle on a strip of pixels (the oracle is too slow for all of them)."""
    from xtomosim.back_project import effective_mu_water, get_recon
    from xtomosim.forward_project import get_sino
    ct, ph, lo, hi = S.make_config('C2')
    mono = S.gaussian_spectrum('mono', 70.0, 0.4, lo.E)
    raw, log = get_sino(ct, ph, mono)
    rec, hu = get_recon(log, ct, mono, 512, 51.2, 0.8)
    assert rec.shape == (512, 512) and rec.dtype == np.float64
    mu70 = ph.mu_table(np.array([70.0]))[:, 0]
    lab = ph.M[0]
    from scipy.ndimage import binary_erosion
    body = binary_erosion(lab == 1, iterations=6)
    yy, xx = np.mgrid[0:512, 0:512] - 255.5
    # the fan covers a disc of radius SID sin(fan/2) = 24 cm; the corners of the square are outside the scanned field
    air = binary_erosion(lab == 0, iterations=6) & (np.hypot(xx, yy) * 0.1 < 22.0)
    assert abs(rec[body].mean() - mu70[1]) < 0.01 * mu70[1]
    assert abs(rec[air].mean()) < 0.01 * mu70[1]
    mw = effective_mu_water(ct, mono)
    assert np.allclose(hu, 1000.0 * (rec - mw) / mw, rtol=1e-5, atol=1e-2)
    # strip parity against the oracle: 8 image rows through the middle
    scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, rotation=ct.rotation)
    beta = scan.beta0 + np.arange(ct.N_proj) * scan.dbeta
    n, fov = 512, 51.2
    q = FO.filter_lines(log, scan.gamma0, scan.dgamma, SID, 0.8)
    qp = np.pad(q, ((0, 0), (1, 1)))
    c = (np.arange(n) - 0.5 * (n - 1)) * (fov / n)
    X, Y = np.meshgrid(c, c[252:260])
    ref = np.zeros_like(X)
    for v in range(ct.N_proj):
        cb, sb = np.cos(beta[v]), np.sin(beta[v])
        A = SID - (X * cb + Y * sb)
        B = Y * cb - X * sb
        u = np.clip((np.arctan2(-B, A) - scan.gamma0) / scan.dgamma, -1.0, float(ct.N_channels))
        fl = np.minimum(np.floor(u), ct.N_channels - 1.0)
        i0 = fl.astype(np.int64)
        ref += (qp[v][i0 + 1] + (u - fl) * (qp[v][i0 + 2] - qp[v][i0 + 1])) / (A * A + B * B)
    ref *= abs(scan.dbeta)
    assert np.max(np.abs(rec[252:260] - ref)) / np.max(np.abs(rec)) < 5e-5


@pytest.mark.parametrize('dtype', [F32, F64])
@pytest.mark.parametrize('n_rows', [1, 3, 9, 17, 30])
def test_wide_fan_and_forced_kernel_variants(dtype, n_rows, monkeypatch):
    """A 1.6 rad fan lies outside the range of the single-precision kernels' arctangent polynomial (|tan gamma| <= 0.85), so
    every back-projection kernel runs its general variant; the one-line filter kernel and the smaller row blocks (tuning
    knobs DEXCT_FBP_FILTER / DEXCT_FBP_RB) must reproduce the same image."""
    fan, n_ch, n_views = 1.6, 181, 64
    scan = make_scan(n_views, n_ch, fan, SID, SDD, n_rows=n_rows, h_iso=0.3, beta0=0.1)
    beta = scan.beta0 + np.arange(n_views) * scan.dbeta
    gam = scan.gamma0 + np.arange(n_ch) * scan.dgamma
    r = np.random.default_rng(n_rows)
    sino = np.stack([FO.disc_sinogram(beta, gam, SID, (2.0 - 0.3 * (k % 4), 1.0), 9.0 - (k % 3), 0.2)
                     + 0.01 * r.standard_normal((n_views, n_ch)) for k in range(n_rows)], axis=1)
    ref = FO.fbp_fan(sino, beta, scan.gamma0, scan.dgamma, scan.dbeta, SID, 44, 36.0, 0.8)
    d = torch.as_tensor(sino, device='cuda')
    got = fbp_fan(scan, d, 44, 36.0, 0.8, dtype=dtype)
    assert scale_rel_err(got.cpu().numpy(), ref) < TOL[dtype]
    for knob, values in (('DEXCT_FBP_FILTER', ['1']), ('DEXCT_FBP_RB', ['0', '8', '16'])):
        for v in values:
            monkeypatch.setenv(knob, v)
            alt = fbp_fan(scan, d, 44, 36.0, 0.8, dtype=dtype)
            monkeypatch.delenv(knob)
            assert scale_rel_err(alt.cpu().numpy(), ref) < TOL[dtype], (knob, v)
            assert scale_rel_err(alt.cpu().numpy(), got.cpu().numpy()) < (2e-6 if dtype == F32 else 1e-13), (knob, v)


@pytest.mark.parametrize('dtype', [F32, F64])
@pytest.mark.parametrize('rotation,n_rows', [(np.pi + FAN + 0.05, 1), (-4.4, 1), (5.6, 3), (4.0, 9), (-3.99, 30)])
def test_short_scans_take_parker_weights(dtype, rotation, n_rows):
    """rotation_angle_total is a parameter of the reference's scanner (input/params.txt:24): anything from pi + fan angle up to
    a full turn reconstructs, with Parker's short-scan weights applied where the filter reads the sinogram -- every kernel
    variant (1, 3, 9 and 30 rows), both directions of rotation, view chunks accumulating; against the oracle, whose weights
    are checked on the CPU (conjugate rays add up to 1; a disc reconstructs to its attenuation)."""
    n_ch, n_views = 151, 96
    scan = make_scan(n_views, n_ch, FAN, SID, SDD, n_rows=n_rows, h_iso=0.3, rotation=rotation, beta0=0.4)
    beta = scan.beta0 + np.arange(n_views) * scan.dbeta
    gam = scan.gamma0 + np.arange(n_ch) * scan.dgamma
    rows = [FO.disc_sinogram(beta, gam, SID, (3.0 - (k % 4), -2.0), 7.0 - (k % 3), 0.2) +
            FO.disc_sinogram(beta, gam, SID, (-5.0, 4.0), 2.0, 0.35) for k in range(n_rows)]
    sino = np.stack(rows, axis=1)
    ref = FO.fbp_fan(sino, beta, scan.gamma0, scan.dgamma, scan.dbeta, SID, 64, 36.0, 0.8, short_scan=True)
    d = torch.as_tensor(sino, device='cuda')
    got = fbp_fan(scan, d, 64, 36.0, 0.8, dtype=dtype)
    assert scale_rel_err(got.cpu().numpy(), ref) < TOL[dtype]
    # the level is right (the weights are normalised): centre of the big disc of the first slice
    c = (np.arange(64) - 31.5) * (36.0 / 64)
    X, Y = np.meshgrid(c, c)
    inside = np.hypot(X - 3.0, Y + 2.0) < 4.0
    assert abs(float(got[0].cpu().numpy()[inside].mean()) - 0.2) < 0.01
    img = torch.zeros_like(got)
    for v0, v1 in ((0, 41), (41, 96)):
        fbp_fan(scan, d[v0:v1], 64, 36.0, 0.8, view_begin=v0, dtype=dtype, image=img, accumulate=True)
    assert scale_rel_err(img.cpu().numpy(), got.cpu().numpy()) < (1e-6 if dtype == F32 else 1e-13)


def test_unsupported_rotations_are_refused_and_get_recon_takes_short_scans():
    scan = make_scan(60, 100, FAN, SID, SDD, rotation=np.pi + 0.5 * FAN)           # less than pi + fan: not enough data
    with pytest.raises(N.DexctError, match='pi \\+ fan'):
        fbp_fan(scan, torch.zeros((60, 100), device='cuda'), 32, 20.0, 1.0)
    scan = make_scan(60, 100, FAN, SID, SDD, rotation=4.0 * np.pi)                  # two turns
    with pytest.raises(N.DexctError, match='full turn'):
        fbp_fan(scan, torch.zeros((60, 100), device='cuda'), 32, 20.0, 1.0)
    # the drop-in entry point: a 4.2 rad scan of a disc through get_recon
    from xtomosim.back_project import get_recon
    from xtomosim.system import FanBeamGeometry
    ct = FanBeamGeometry(160, 120, FAN, SID, SDD, rotation=4.2)
    mono = S.gaussian_spectrum('m', 70.0, 2.0, np.arange(60.0, 81.0))
    sino = FO.disc_sinogram(ct.betas, ct.gammas, SID, (2.0, 1.0), 6.0, 0.2)
    rec, hu = get_recon(sino, ct, mono, 64, 30.0, 0.9)
    c = (np.arange(64) - 31.5) * (30.0 / 64)
    X, Y = np.meshgrid(c, c)
    assert abs(rec[np.hypot(X - 2.0, Y - 1.0) < 4.0].mean() - 0.2) < 0.005


@pytest.mark.parametrize('n_ch,n_rows', [(700, 32), (513, 24), (256, 8)])
def test_tensor_core_filter_equals_the_fp32_pipe_filter(n_ch, n_rows):
    """Several slabs of 256 output channels, a partly filled last slab, line counts that end inside a 128-line tile: images
    reconstructed through fbp_filter_tc_kernel (3xTF32 on tcgen05, the default for row-fastest single-precision data) against
    the same through the FP32-pipe filter (DEXCT_FBP_TC=0), which the oracle tests above pin."""
    n_views, n = 30, 48
    scan = make_scan(n_views, n_ch, FAN, SID, SDD, n_rows=n_rows, h_iso=0.3, beta0=0.2)
    r = np.random.default_rng(n_ch)
    sino = torch.as_tensor(r.random((n_views, n_rows, n_ch)).astype(np.float32) + 0.5, device='cuda')
    imgs = {}
    for flag in ('1', '0'):
        os.environ['DEXCT_FBP_TC'] = flag
        try:
            imgs[flag] = fbp_fan(scan, sino, n, 30.0, 0.8, dtype=F32).cpu().numpy()
        finally:
            os.environ.pop('DEXCT_FBP_TC', None)
    ref = fbp_fan(scan, sino.double(), n, 30.0, 0.8, dtype=F64).cpu().numpy()
    e_tc, e_fp = scale_rel_err(imgs['1'], ref), scale_rel_err(imgs['0'], ref)
    print(f'\nn_ch {n_ch}, {n_rows} rows: tensor-core filter vs float64 {e_tc:.2e}, FP32-pipe filter vs float64 {e_fp:.2e}')
    assert np.isfinite(imgs['1']).all()
    # white noise on an offset is the hard case for single precision (zero DC gain: the image is what is left after the
    # partial sums cancel; measured 1.2e-5 .. 3.7e-5 for BOTH filters, the back-projection's rounding included): the
    # tensor-core filter must not be worse than the FP32-pipe one, and the two images must agree far better than that
    assert e_tc < 1.2 * e_fp + 2e-6 and e_fp < 1e-4
    assert scale_rel_err(imgs['1'], imgs['0']) < 1e-5
