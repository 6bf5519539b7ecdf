"""Parity of the CUDA Gauss-Newton decomposition (csrc/gn.cu, through the C ABI and the drop-in matdecomp module) with
the reference: golden vectors produced by /root/reference/matdecomp.py itself (tools/make_golden.py) and the pinned numpy
restatement (oracle/gn_oracle.py).

Tolerances (BASELINE.json north_star): f64 mode 1e-10, f32 mode 1e-5, both scale-relative per array
(max|a - ref| / max|ref|): the small basis component of a pixel is ill-determined, so an element-wise relative bound is
not attainable by ANY float32-input implementation (hard part 2).  In f32 mode the oracle is fed the
same float32-rounded counts the kernel sees.
"""
import os
import types

import numpy as np
import pytest
import torch

from conftest import scale_rel_err
from dexct import _native as N
from dexct import synthetic as S
from dexct.ops import F32, F64, Projector, gn_decompose, make_scan
from oracle import gn_oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=['tensor-core', 'fp32-pipe'])
def f32_kernel(request):
    """Every test of this module runs twice: single precision through the tensor-core kernel (gn_tc_kernel, the default up
    to 448 energy bins) and through the FP32-pipe kernel (gn_kernel: DEXCT_GN_TC=0, and the fallback beyond 448 bins)."""
    old = os.environ.get('DEXCT_GN_TC')
    os.environ['DEXCT_GN_TC'] = '1' if request.param == 'tensor-core' else '0'
    yield request.param
    if old is None:
        os.environ.pop('DEXCT_GN_TC', None)
    else:
        os.environ['DEXCT_GN_TC'] = old


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + '.npz'))


def _oracle(g, i0, mus, n_iters):
    nB = g.shape[2]
    i0_t = np.tile(i0, (1, nB)).reshape(2, nB, i0.shape[1])
    return gn_oracle.optimize_sino_cpu(g, None, i0_t, mus, n_iters)


@pytest.mark.parametrize('name', ['gn_synth_10it', 'gn_synth_25it', 'gn_c1_tables'])
def test_f64_matches_reference_golden(golden_dir, name):
    d = _load(golden_dir, name)
    out = gn_decompose(d['g'], d['i0'], d['mus'], int(d['n_iters']), dtype=F64).cpu().numpy()
    assert out.shape == d['a_ref'].shape
    assert scale_rel_err(out, d['a_ref']) < 1e-10


@pytest.mark.parametrize('name', ['gn_synth_25it', 'gn_c1_tables'])
def test_f32_matches_reference_on_rounded_inputs(golden_dir, name):
    d = _load(golden_dir, name)
    g32 = d['g'].astype(np.float32).astype(np.float64)
    i032 = d['i0'].astype(np.float32).astype(np.float64)
    mus32 = d['mus'].astype(np.float32).astype(np.float64)
    ref = _oracle(g32, i032, mus32, int(d['n_iters']))
    out = gn_decompose(d['g'], d['i0'], d['mus'], int(d['n_iters']), dtype=F32).cpu().numpy()
    assert scale_rel_err(out, ref) < 1e-5
    # and it still lands on the truth the golden file was synthesised from, to float32-conditioning accuracy
    # (gn_c1_tables has four planted air pixels at the start of view 0 whose a_true is not what was measured)
    skip = 4 if name == 'gn_c1_tables' else 0
    assert scale_rel_err(out.reshape(-1, 2)[skip:], d['a_true'].reshape(-1, 2)[skip:]) < 2e-3


def test_air_mask_and_dropin_entry_point(golden_dir):
    """get_basismat_sinos through the drop-in module == the reference's get_basismat_sinos (golden mat1/mat2)."""
    import matdecomp
    from dexct import config
    d = _load(golden_dir, 'gn_c1_tables')
    ct = types.SimpleNamespace(det_E=d['det_E'], det_eta_E=d['det_eta_E'], eid=bool(d['eid']))
    s1 = types.SimpleNamespace(E=d['spec1_E'], I0=d['spec1_I0'])
    s2 = types.SimpleNamespace(E=d['spec2_E'], I0=d['spec2_I0'])
    old = config.dtype
    try:
        config.dtype = 'f64'
        m1, m2 = matdecomp.get_basismat_sinos(ct, d['g'][0], d['g'][1], s1, s2, n_iters=int(d['n_iters']),
                                              mask_thresh=float(d['mask_thresh']))
        assert m1.dtype == np.float64 and m1.shape == d['mat1'].shape
        assert scale_rel_err(m1, d['mat1']) < 1e-10 and scale_rel_err(m2, d['mat2']) < 1e-10
        assert np.array_equal(m1 == 0, d['mat1'] == 0) and (m1[0, :4] == 0).all() and (m2[0, :4] == 0).all()
        config.dtype = 'f32'
        m1, m2 = matdecomp.get_basismat_sinos(ct, d['g'][0], d['g'][1], s1, s2, n_iters=int(d['n_iters']),
                                              mask_thresh=float(d['mask_thresh']))
        assert np.array_equal(m1 == 0, d['mat1'] == 0)
        assert scale_rel_err(m1, d['mat1']) < 2e-3 and scale_rel_err(m2, d['mat2']) < 2e-3
        # the reference's own function names resolve to the kernel too
        a = matdecomp.optimize_sino_cpu(d['g'], d['ee'], np.tile(d['i0'][:, None, :], (1, 48, 1)), d['mus'], 30)
        assert a.shape == (2, 48, 2)
    finally:
        config.dtype = old


@pytest.mark.parametrize('dtype', [F32, F64])
def test_edge_sizes(dtype, golden_dir):
    d = _load(golden_dir, 'gn_synth_25it')
    g, i0, mus = d['g'], d['i0'], d['mus']
    # zero iterations: the initial value (matdecomp.py:98)
    out = gn_decompose(g, i0, mus, 0, dtype=dtype).cpu().numpy()
    assert np.allclose(out, 1e-6, rtol=1e-6)
    # a single pixel, and a ragged pixel count that is no multiple of anything
    flat = g.reshape(2, -1)
    ref = d['a_ref'].reshape(-1, 2)
    for P in (1, 37, 119):
        out = gn_decompose(flat[:, :P], i0, mus, 25, dtype=dtype).cpu().numpy()
        assert out.shape == (P, 2)
        assert scale_rel_err(out, ref[:P]) < (1e-10 if dtype == F64 else 2e-3)
    # no pixels at all
    out = gn_decompose(np.zeros((2, 0)), i0, mus, 5, dtype=dtype)
    assert out.shape == (0, 2)
    with pytest.raises(ValueError):
        gn_decompose(flat, i0[:1], mus, 3, dtype=dtype)
    L = N.lib()
    assert L.dexct_gn_decompose(None, None, None, 1, 1, 1, 0.0, None, None, 0, 0, None) == -1


def test_c3_full_size_self_consistency():
    """Config C3 (2048 x 2048 pixel pairs, 120 bins): at 20 iterations the recovered line integrals reproduce the measured
    counts through the forward model (a well-conditioned, size-independent property), in both precisions; a subset is
    compared with the reference restatement."""
    g, i0, mus, a_true = S.gn_problem(256, 2048)              # 512 Ki pixels on the host; tiled to 2048 x 2048 on the device
    gd = torch.as_tensor(g).cuda()
    gd = gd.repeat(1, 8, 1).contiguous()                      # [2, 2048, 2048]
    i0d, musd = torch.as_tensor(i0).cuda(), torch.as_tensor(mus).cuda()
    for dtype, tol in ((F32, 2e-5), (F64, 1e-12)):
        a = gn_decompose(gd, i0d, musd, 20, dtype=dtype).double()
        assert a.shape == (2048, 2048, 2)
        att = torch.exp(-(a[..., 0:1] * musd[0] + a[..., 1:2] * musd[1]))
        for m in range(2):
            nu = (att * i0d[m]).sum(-1)
            assert float(((nu - gd[m]).abs() / gd[m]).max()) < tol
        # tiles are identical problems -> identical answers (no cross-pixel state)
        assert torch.equal(a[:256], a[1792:])
    ref = _oracle(g[:, :2], i0, mus, 20)
    out = gn_decompose(g[:, :2], i0, mus, 20, dtype=F64).cpu().numpy()
    assert scale_rel_err(out, ref) < 1e-10
    assert scale_rel_err(out, a_true[:2]) < 1e-8


def test_starved_pixels_stay_finite_in_single_precision():
    """Photon-starved / metal-like pixels: the Newton iterates go where 2^x leaves the float range, and only the
    reference's exponent clip (matdecomp.py:116) keeps its float64 sums finite.  The single-precision kernel hands such
    pixels to its float64 pass: the result must be finite wherever the oracle's is, and equal to it to float accuracy
    (same float-rounded inputs), while ordinary pixels are untouched."""
    E = np.arange(1.0, 141.0)
    lo, hi = S.gaussian_spectrum('lo', 55.0, 18.0, E, 1e8), S.gaussian_spectrum('hi', 95.0, 25.0, E, 1e8)
    i0 = np.stack([lo.I0, hi.I0])
    mus = np.stack([0.2 * (60.0 / E) ** 2.2 + 0.15, 0.9 * (60.0 / E) ** 2.9 + 0.12])      # steep at 1 keV like real mu/rho
    r = np.random.default_rng(9)
    a = np.stack([r.uniform(0, 25, 400), r.uniform(0, 4, 400)], axis=-1)
    att = np.exp(-(a[:, 0:1] * mus[0] + a[:, 1:2] * mus[1]))
    g = np.stack([att @ i0[0], att @ i0[1]])                   # [2, 400] consistent counts
    g[1, :40] *= r.uniform(1e-9, 1e-3, 40)                     # starve the second measurement of 40 pixels
    g[0, 40:60] *= r.uniform(5.0, 50.0, 20)                    # and make 20 more inconsistent the other way
    r32 = lambda x: np.asarray(x).astype(np.float32).astype(np.float64)                  # noqa: E731
    gq, i0q, musq = r32(g), r32(i0), r32(mus)
    ref = np.full((400, 2), np.nan)
    with np.errstate(all='ignore'):
        for p in range(400):                                   # pixel by pixel: one singular Hessian must not abort the rest
            try:
                ref[p] = _oracle(gq[:, None, p:p + 1], i0q, musq, 30)[0, 0]
            except np.linalg.LinAlgError:
                pass
    out32 = gn_decompose(gq, i0q, musq, 30, dtype=F32).cpu().numpy()
    out64 = gn_decompose(gq, i0q, musq, 30, dtype=F64).cpu().numpy()
    fin = np.isfinite(ref).all(axis=1)
    assert fin[60:].all() and fin[:60].sum() >= 10            # the oracle itself loses some of the starved pixels
    assert np.isfinite(out32[fin]).all() and np.isfinite(out64[fin]).all()
    # ordinary pixels: the usual tolerances
    assert scale_rel_err(out64[60:], ref[60:]) < 1e-9
    assert scale_rel_err(out32[60:], ref[60:]) < 1e-4
    # starved pixels: the Newton map is chaotic there (float64 results of two summation orders already differ in the
    # 4th digit), so only the bulk is compared: the float64 pass of the single-precision kernel follows the float64
    # kernel (it sums the bins in another order) on most of them
    st = np.flatnonzero(fin[:60])
    rel = np.abs(out32[st] - out64[st]).max(axis=1) / np.maximum(np.abs(out64[st]).max(axis=1), 1.0)
    assert np.median(rel) < 1e-4


@pytest.mark.parametrize('nE', [2, 3, 509, 510, 511, 640])
def test_bin_counts_around_the_constant_bank_limit(nE):
    """The single-precision kernel keeps its tables in the constant bank up to 510 bins (two padding rows of the 512) and in
    shared memory beyond; odd bin counts run the two-bin pipelined loop into a zero row.  Every size must give the
    reference's numbers (oracle/gn_oracle.py) -- synthetic spectra, counts from the forward model of matdecomp.py:116-118."""
    rng = np.random.default_rng(nE)
    E = np.linspace(40.0, 140.0, nE)
    mus = np.stack([0.2 + 0.6 * (E / 20.0) ** -3 + 0.1 * (E / 60.0) ** -0.3,
                    0.15 + 6.0 * (E / 20.0) ** -3 + 0.12 * (E / 60.0) ** -0.3])
    i0 = (np.stack([np.exp(-0.5 * ((E - 55) / 18) ** 2), np.exp(-0.5 * ((E - 95) / 25) ** 2)]) + 1e-3) * 3e6 / nE
    a_true = np.stack([rng.uniform(0, 12, (3, 41)), rng.uniform(0, 3, (3, 41))], -1)
    att = np.exp(-(a_true[..., None, :] * mus.T[None, None]).sum(-1))             # [3, 41, nE]
    g = np.stack([(att * i0[m]).sum(-1) for m in range(2)])
    want = _oracle(g, i0, mus, 20)
    assert scale_rel_err(want, a_true) < 1e-8                                      # the oracle recovers the truth
    out64 = gn_decompose(g, i0, mus, 20, dtype=F64).cpu().numpy()
    assert scale_rel_err(out64, want) < 1e-10
    g32 = g.astype(np.float32).astype(np.float64)
    i032 = i0.astype(np.float32).astype(np.float64)
    mus32 = mus.astype(np.float32).astype(np.float64)
    want32 = _oracle(g32, i032, mus32, 20)
    out32 = gn_decompose(g, i0, mus, 20, dtype=F32).cpu().numpy()
    assert np.isfinite(out32).all()
    assert scale_rel_err(out32, want32) < 1e-5


def _sampled_f32_parity(g_dev, i0, mus, n_iters, lines, tol=1e-5):
    """Single-precision decomposition of a FULL-size problem against the oracle on a sample of its detector lines: the oracle
    gets exactly the float32 values the kernel read (counts, i0, mus)."""
    a = gn_decompose(g_dev, i0, mus, n_iters, dtype=F32)
    g32 = g_dev.float()[:, lines].double().cpu().numpy()                        # what the kernel read, as float64
    r32 = lambda x: np.asarray(x).astype(np.float32).astype(np.float64)        # noqa: E731
    with np.errstate(all='ignore'):
        ref = _oracle(g32, r32(i0), r32(mus), n_iters)
    got = a[lines].double().cpu().numpy()
    assert got.shape == ref.shape
    fin = np.isfinite(ref).all(-1)
    assert np.isfinite(got[fin]).all()
    return got, ref, g32


def test_c3_full_size_f32_against_the_oracle_on_64_lines():
    """Config C3 (2048 x 2048 pixel pairs, 120 bins, the fixed 10 iterations BASELINE.json names): the single-precision kernel
    on the whole problem, compared with the reference restatement on 64 of its 2048 detector lines (131 072 pixels) at 1e-5."""
    g, i0, mus, a_true = S.gn_problem(2048, 2048)
    gd = torch.as_tensor(g).cuda()
    lines = np.linspace(0, 2048, 64, endpoint=False).astype(int)
    got, ref, _ = _sampled_f32_parity(gd, i0, mus, 10, lines)
    err = scale_rel_err(got, ref)
    assert err < 1e-5, f'C3 f32 vs oracle, 64 lines: {err:.3e}'


def test_c2_full_size_f32_against_the_oracle_on_64_views():
    """Config C2: sinograms of the whole 1000-view scan from the projector, decomposed in single precision (10 iterations, air
    mask off so that every pixel is compared), against the oracle on 64 views.  Pixels with less than one detected photon in
    either measurement (rays through the steel and titanium inserts) are photon-starved: their Newton iterates are still
    travelling after 10 steps and the map is chaotic there, so they are held to finiteness only."""
    from xtomosim.forward_project import spectral_weights
    import matdecomp
    ct, ph, lo, hi = S.make_config('C2')
    scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, rotation=ct.rotation)
    pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz))
    W = np.stack([spectral_weights(ct, lo), spectral_weights(ct, hi)])
    counts, _, _ = pj.project(ph.mu_table(lo.E), W, dtype=F32, want_log=False)
    gd = counts[:, :, 0, :].contiguous()                                       # [2, 1000, 900]
    _, i0, mus = matdecomp.spectral_tables(ct, lo, hi)
    views = np.linspace(0, ct.N_proj, 64, endpoint=False).astype(int)
    got, ref, g32 = _sampled_f32_parity(gd, i0, mus, 10, views)
    lit = g32.min(axis=0) >= 1.0
    # measured 0.63: the low spectrum (mean 55 keV) leaves less than one photon behind the long chords of the 43 cm body
    assert lit.mean() > 0.5
    err = float(np.max(np.abs(got[lit] - ref[lit])) / np.max(np.abs(ref[lit])))
    assert err < 1e-5, f'C2 f32 vs oracle, 64 views, lit pixels: {err:.3e}'


@pytest.mark.parametrize('n_meas,n_mats', [(2, 2), (3, 3), (3, 2), (4, 3), (1, 1), (4, 4)])
def test_general_form_per_bin_spectra_and_other_material_counts(n_meas, n_mats):
    """optimize_sino_cpu in the generality the reference wrote it with (matdecomp.py:95,101-105,118,125): a spectrum per
    detector bin, 1..4 measurements and basis materials -- through the drop-in entry point, against the oracle (which
    tests/test_oracle_gn.py pins to the reference function on the same shapes)."""
    import matdecomp
    from dexct import config
    rng = np.random.default_rng(100 * n_meas + n_mats)
    nV, nB, nE = 3, 29, 64
    E = 30.0 + 1.5 * np.arange(nE)
    base = np.stack([0.2 + 5.0 / E, 0.15 + 60.0 / E ** 1.5, 0.1 + 900.0 / E ** 2.5, 0.05 + 2.0e4 / E ** 3.2])
    mus = base[:n_mats]
    shape = np.exp(-0.5 * ((E[None, :] - np.linspace(45, 110, n_meas)[:, None]) / 16.0) ** 2) + 0.02
    i0 = shape[:, None, :] * rng.uniform(0.5, 1.5, (n_meas, nB, 1)) * rng.uniform(0.9, 1.1, (n_meas, nB, nE)) * 1e5
    a = np.stack([rng.uniform(0, [12, 2.5, 0.3, 0.05][q], (nV, nB)) for q in range(n_mats)], axis=-1)
    att = np.exp(-(a[..., None, :] * mus.T[None, None]).sum(-1))
    g = np.stack([(att * i0[m][None]).sum(-1) for m in range(n_meas)])
    n_iters = 12
    want = gn_oracle.optimize_sino_cpu(g, E, i0, mus, n_iters)
    old = config.dtype
    try:
        config.dtype = 'f64'
        got = matdecomp.optimize_sino_cpu(g, E, i0, mus, n_iters, verbose=False)
        assert got.shape == (nV, nB, n_mats) and got.dtype == np.float64
        assert scale_rel_err(got, want) < 1e-10
        if n_meas == n_mats:
            # a square, consistent system: the truth comes back (four materials are so ill-conditioned that 12 Newton steps
            # of the reference's own iteration are still 5e-5 away -- kernel and oracle agree on that to 1e-10 above)
            assert scale_rel_err(got, a) < (1e-6 if n_mats <= 3 else 1e-3)
        # single-precision I/O: the oracle is fed the rounded inputs
        config.dtype = 'f32'
        r32 = lambda x: np.asarray(x).astype(np.float32).astype(np.float64)            # noqa: E731
        want32 = gn_oracle.optimize_sino_cpu(r32(g), E, r32(i0), r32(mus), n_iters)
        got32 = matdecomp.optimize_sino(g, E, i0, mus, n_iters)
        assert scale_rel_err(got32, want32) < 1e-5
        # one spectrum tiled over the bins with three materials takes the same kernel ([nMeas, nE] form inside)
        if n_mats != 2 or n_meas != 2:
            i0_t = np.tile(i0[:, :1, :], (1, nB, 1))
            g_t = np.stack([(att * i0_t[m][None]).sum(-1) for m in range(n_meas)])
            config.dtype = 'f64'
            assert scale_rel_err(matdecomp.optimize_sino_cpu(g_t, E, i0_t, mus, n_iters),
                                 gn_oracle.optimize_sino_cpu(g_t, E, i0_t, mus, n_iters)) < 1e-10
    finally:
        config.dtype = old
    with pytest.raises(N.DexctError, match='1 to 4'):
        from dexct.ops import gn_decompose_general
        gn_decompose_general(np.ones((5, 1, 4)), np.ones((5, 8)), np.ones((2, 8)), 1, dtype=F64)


@pytest.mark.parametrize('P', [1, 127, 128, 129, 1000, 5000, 70001])
def test_air_pixels_left_out_of_the_list_equals_the_full_decomposition(P, f32_kernel):
    """With an air mask the tensor-core kernel iterates only a LIST of the non-air pixels (gn_compact_kernel zeroes the others
    up front).  Same results, bit for bit, as decomposing every tile and zeroing afterwards (DEXCT_GN_COMPACT=0) -- including
    sinograms that are all air, have no air, or end in the middle of a tile."""
    g, i0, mus, _ = S.gn_problem(71, 1000)
    gg = np.ascontiguousarray(g.reshape(2, -1)[:, :P]).copy()
    r = np.random.default_rng(P)
    air = r.random(P) < 0.3
    air[: P // 7] = True                                      # a run of air at the start, like the end of a detector line
    gg[0, air] = gg[0].max() * 1.2
    gg[1, air] = gg[1].max()
    for mask in (0.8, 0.0):
        outs = {}
        for flag in ('1', '0'):
            os.environ['DEXCT_GN_COMPACT'] = flag
            try:
                outs[flag] = gn_decompose(gg, i0, mus, 7, mask_thresh=mask, dtype=F32).cpu().numpy()
            finally:
                os.environ.pop('DEXCT_GN_COMPACT', None)
        assert np.array_equal(outs['1'], outs['0'])
        if mask > 0:
            masked = gg[0].astype(np.float32) >= np.float32(mask) * gg[0].astype(np.float32).max()     # matdecomp.py:196
            assert (outs['1'][masked] == 0).all() and (outs['1'][~masked] != 0).all()
            assert masked[air].all()
    # all air
    os.environ['DEXCT_GN_COMPACT'] = '1'
    try:
        z = gn_decompose(np.full((2, P), 5.0), i0, mus, 7, mask_thresh=0.5, dtype=F32).cpu().numpy()
    finally:
        os.environ.pop('DEXCT_GN_COMPACT', None)
    assert (z == 0).all()
