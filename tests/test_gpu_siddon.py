"""Parity of the CUDA forward projector (csrc/siddon.cu, through the C ABI) with the oracle.

* voxel-index / material-id SEQUENCES: bit-exact against oracle section (B) (the operation-for-operation CPU twin), which
  tests/test_oracle_siddon.py ties to Siddon 1985 (section A);
* per-material lengths, counts, log sinograms: against Siddon 1985 in float64 -- 1e-10 scale-relative in f64 mode,
  1e-5 in f32 mode (the tolerances of BASELINE.json's north_star);
* full-size properties at config C2 and a slab of C4: chord-length conservation, linearity, closed forms.
"""
import ctypes as C

import numpy as np
import pytest
import torch

import oracle
from conftest import elem_rel_err, per_ray_len_err, scale_rel_err
from dexct import _native as N
from dexct import synthetic as S
from dexct.ops import F32, F64, Projector, make_scan, poly_detect

pytestmark = pytest.mark.gpu

FAN = 0.8230337
TOL = {F32: 1e-5, F64: 1e-10}
# element-wise bounds (conftest.elem_rel_err / per_ray_len_err): counts on every pixel within six decades of the air count,
# lengths of every ray against that ray's own chord.  A short chord is a sum of few voxel pieces, each placed with the
# absolute accuracy of the column coordinate (2^-24 of the column count in single precision), so the per-chord bound of
# short rays is looser than the scale-relative one by the ratio of the chords.
TOL_ELEM = {F32: 1e-5, F64: 1e-10}
TOL_RAY = {F32: 2e-5, F64: 1e-10}


def _labels(n, nz, K, seed=3):
    r = np.random.default_rng(seed)
    lab = r.integers(0, K, (nz, n, n)).astype(np.uint8)
    lab[:, n // 4:n // 2, n // 3:2 * n // 3] = min(2, K - 1)
    return lab


def _pair(n=96, n_ch=120, n_views=36, n_rows=1, nz=1, K=4, fov=51.2, h_iso=0.4, beta0=0.0, seed=3, fan=FAN):
    """(Projector, oracle.Scan built from the DEVICE's own angle tables)."""
    lab = _labels(n, nz, K, seed)
    vox = fov / n
    voxel = (vox, vox, vox if nz > 1 else 0.1)
    scan = make_scan(n_views, n_ch, fan, 60.0, 100.0, n_rows=n_rows, h_iso=h_iso, beta0=beta0)
    pj = Projector(scan, lab, voxel)
    sc = oracle.Scan(lab, voxel, 60.0, 100.0, pj.tables_host())
    return pj, sc


def _tables(K, nE=24, nS=2, seed=5):
    r = np.random.default_rng(seed)
    E = np.linspace(30, 120, nE)
    mu = np.stack([(0.02 + 0.08 * m) * (0.6 + (60.0 / E) ** (1.0 + 0.2 * m)) for m in range(K)])
    mu[0] = 0.0002                                    # air
    W = r.uniform(0.2, 1.0, (nS, nE)) * 1e4
    return mu, W


def test_device_tables_are_the_stated_conventions():
    pj, _ = _pair(n_rows=3, nz=4, beta0=0.3)
    t = pj.tables_host()
    want = oracle.fan_tables(120, 36, FAN, beta0=0.3, N_rows=3, h_iso=0.4)
    for k in want:
        assert np.allclose(t[k], want[k], rtol=0, atol=2e-15), k


@pytest.mark.parametrize('dtype', [F32, F64])
@pytest.mark.parametrize('nz,n_rows', [(1, 1), (16, 6)])
def test_voxel_sequences_bit_exact(dtype, nz, n_rows):
    pj, sc = _pair(nz=nz, n_rows=n_rows)
    r = np.random.default_rng(11)
    n_all = sc.n_views * sc.n_rows * sc.n_ch
    ids = np.unique(np.concatenate([r.integers(0, n_all, 1500),
                                    # central and edge channels of the axis-parallel / diagonal views
                                    [((v * sc.n_rows + row) * sc.n_ch + c) for v in (0, 9, 18, 27, 4, 5)
                                     for row in (0, sc.n_rows - 1) for c in (0, 59, 60, 119)]]))
    got = pj.trace(ids, max_seg=1024, dtype=dtype)
    n_seg = 0
    for rid, (vox, lab, ln) in zip(ids, got):
        c = int(rid % sc.n_ch)
        row = int((rid // sc.n_ch) % sc.n_rows)
        v = int(rid // (sc.n_ch * sc.n_rows))
        w_vox, w_lab, w_len = sc.slab_trace(v, row, c, dtype == F64)
        assert np.array_equal(vox, w_vox), (rid, v, row, c)
        assert np.array_equal(lab, w_lab), (rid, v, row, c)
        assert np.array_equal(ln.astype(np.float64), w_len), (rid, v, row, c)
        n_seg += len(vox)
    assert n_seg > 50 * len(ids)


def _on_grid_plane(sc, v, row, c, shape, xmajor):
    """The in-plane ray runs along a plane of the minor axis (both end points within 1e-6 voxel of the same plane): which of
    the two neighbouring voxel rows it belongs to is a tie that rounding decides."""
    nz, ny, nx = shape
    S, T = sc.endpoints(v, row, c)
    ax, n, d = (1, ny, sc.voxel[1]) if xmajor else (0, nx, sc.voxel[0])
    gs, gt = S[ax] / d + 0.5 * n, T[ax] / d + 0.5 * n
    return abs(gs - round(gs)) < 1e-6 and abs(gt - round(gt)) < 1e-6 and round(gs) == round(gt)


def _check_production_walk(pj, sc, ids, dtype, K, shape, ties=None):
    """The steps the PRODUCTION walk (walk_skip / walk_skip3 over the skip layouts, recorded by a sink inside the routine the
    production kernels inline) takes along each ray, against the oracle's voxel sequence of the same ray:
    * the steps tile the ray's column range without gap or overlap;
    * an ordinary column names exactly the oracle's voxels of that column, in order, with their labels (bit-exact indices
      and material ids; lengths to rounding -- the walk computes them from differently grouped operations);
    * a homogeneous stretch of S columns covers oracle voxels of ONE label only -- the one the walk books it to -- and its
      length is the sum of theirs;
    * hence the run-length-compressed material-id sequence of the ray is identical.
    Returns (steps, voxels, homogeneous steps)."""
    nz, ny, nx = shape
    f64 = dtype == F64
    recs, L = pj.walk_trace(ids, K, max_steps=4096, dtype=dtype)
    L = L.cpu().numpy().astype(np.float64)
    n_steps = n_vox = n_hom = 0
    for t, (rid, rec) in enumerate(zip(ids, recs)):
        c = int(rid % sc.n_ch)
        row = int((rid // sc.n_ch) % sc.n_rows)
        v = int(rid // (sc.n_ch * sc.n_rows))
        w_vox, w_lab, w_len = sc.slab_trace(v, row, c, f64)
        tag = (int(rid), v, row, c)
        if not rec['hit']:
            assert len(w_vox) == 0 and rec['n_steps'] == 0, tag
            continue
        if ties is not None and _on_grid_plane(sc, v, row, c, shape, rec['xmajor']):
            # The production walk evaluates such a ray in the mirrored layout when its slope is -1e-16 (1 - 0 rounds to 1: the
            # row below), the oracle in the unmirrored one (the row above): both are Siddon's answer for a ray ON the plane.
            # Held to: the steps tile the column range and the lengths add up to the chord.
            assert abs(float(sum(x.astype(np.float64).sum() for x in rec['len'])) - float(w_len.sum())) <= 1e-5 * max(float(w_len.sum()), 1e-3), tag
            nxt = int(rec['col'][0])
            for s_col, s_run in zip(rec['col'], rec['run']):
                assert int(s_col) == nxt, tag
                nxt += max(int(s_run), 1)
            ties.append(tag)
            continue
        if not rec['fwd']:                                    # the walk always runs towards +major
            w_vox, w_lab, w_len = w_vox[::-1], w_lab[::-1], w_len[::-1]
        o_col = (w_vox % nx) if rec['xmajor'] else ((w_vox // nx) % ny)
        total = float(w_len.sum())
        tiny = (2e-6 if not f64 else 1e-12) * max(total, 1e-3)
        keep = w_len > tiny
        o_vox, o_lab, o_len, o_col = w_vox[keep], w_lab[keep], w_len[keep], o_col[keep]
        assert (np.diff(o_col) >= 0).all(), tag
        pos = 0
        nxt = int(rec['col'][0])
        got_labels, want_labels = [], []
        for s_col, s_run, s_vox, s_lab, s_len in zip(rec['col'], rec['run'], rec['vox'], rec['lab'], rec['len']):
            assert int(s_col) == nxt, tag
            s_len = s_len.astype(np.float64)
            if s_run > 0:
                n_hom += 1
                end = pos
                while end < len(o_col) and o_col[end] < s_col + s_run:
                    end += 1
                lab0 = int(s_lab[0])
                used = s_len > 0                               # a 2-D step has two pieces, the third slot is padding
                used[0] = True
                assert (s_lab[used] == lab0).all(), tag
                assert (o_lab[pos:end] == lab0).all(), (tag, int(s_col), int(s_run))
                assert abs(float(s_len.sum()) - float(o_len[pos:end].sum())) <= 4 * tiny + 1e-6 * (not f64) * float(s_len.sum()), tag
                if s_len.sum() > tiny:
                    got_labels.append(lab0)
                want_labels.extend(int(x) for x in o_lab[pos:end])
                pos = end
                nxt = int(s_col) + int(s_run)
            else:
                end = pos
                while end < len(o_col) and o_col[end] == s_col:
                    end += 1
                live = (s_len > tiny) & (s_vox >= 0)
                assert np.array_equal(s_vox[live], o_vox[pos:end]), (tag, int(s_col), s_vox, s_len, o_vox[pos:end], o_len[pos:end])
                assert np.array_equal(s_lab[live], o_lab[pos:end]), (tag, int(s_col))
                assert np.max(np.abs(s_len[live] - o_len[pos:end]), initial=0.0) <= 4 * tiny, (tag, int(s_col))
                got_labels.extend(int(x) for x in s_lab[live])
                want_labels.extend(int(x) for x in o_lab[pos:end])
                n_vox += end - pos
                pos = end
                nxt = int(s_col) + 1
        assert pos == len(o_col), tag
        rle = lambda a: [x for i, x in enumerate(a) if i == 0 or x != a[i - 1]]          # noqa: E731
        assert rle(got_labels) == rle(want_labels), tag
        # and the walk's per-material sums are the oracle's, label by label
        want_L = np.bincount(w_lab, weights=w_len, minlength=K)[:K]
        assert np.max(np.abs(L[t] - want_L)) <= (3e-6 if not f64 else 1e-11) * max(total, 1e-3) * 4, tag
        n_steps += rec['n_steps']
    return n_steps, n_vox, n_hom


@pytest.mark.parametrize('dtype', [F32, F64])
@pytest.mark.parametrize('nz,n_rows', [(1, 1), (16, 6)])
def test_production_walk_sequences_random_labels(dtype, nz, n_rows):
    """Same ray set as test_voxel_sequences_bit_exact, through the production walk (few homogeneous stretches: random labels)."""
    pj, sc = _pair(nz=nz, n_rows=n_rows)
    r = np.random.default_rng(11)
    n_all = sc.n_views * sc.n_rows * sc.n_ch
    ids = np.unique(np.concatenate([r.integers(0, n_all, 1500),
                                    [((v * sc.n_rows + row) * sc.n_ch + c) for v in (0, 9, 18, 27, 4, 5)
                                     for row in (0, sc.n_rows - 1) for c in (0, 59, 60, 119)]]))
    n_steps, n_vox, n_hom = _check_production_walk(pj, sc, ids, dtype, 4, (nz, 96, 96))
    assert n_vox > 40 * len(ids) and n_hom > 0


@pytest.mark.parametrize('dtype', [F32, F64])
@pytest.mark.parametrize('nz,n_rows,h_iso,voxz', [(1, 1, 0.4, 0.1), (48, 16, 1.2, 51.2 / 96), (48, 16, 0.5, 0.2)])
def test_production_walk_sequences_anatomy_like(dtype, nz, n_rows, h_iso, voxz):
    """Body ellipsoid + inserts: most of every ray is covered by homogeneous stretches, at all z-homogeneity levels."""
    n, n_ch, n_views = 96, 110, 24
    lab = S.phantom_labels(n, nz)
    vox = 51.2 / n
    voxel = (vox, vox, voxz)
    scan = make_scan(n_views, n_ch, FAN, 60.0, 100.0, n_rows=n_rows, h_iso=h_iso, beta0=0.1)
    pj = Projector(scan, lab, voxel)
    sc = oracle.Scan(lab, voxel, 60.0, 100.0, pj.tables_host())
    r = np.random.default_rng(5)
    ids = np.unique(r.integers(0, n_views * n_rows * n_ch, 1200))
    n_steps, n_vox, n_hom = _check_production_walk(pj, sc, ids, dtype, S.K, (nz, n, n))
    assert n_hom > 2 * len(ids)
    # the point of the layouts: far fewer steps than voxels crossed
    segs = sum(len(sc.slab_trace(int(i // (n_ch * n_rows)), int((i // n_ch) % n_rows), int(i % n_ch), dtype == F64)[0]) for i in ids[:200])
    steps = sum(rec['n_steps'] for rec in pj.walk_trace(ids[:200], S.K, dtype=dtype, steps=False)[0])
    assert steps < 0.6 * segs


def test_production_walk_degenerate_rays():
    """Rays along grid planes, through voxel corners, axis-parallel (the set of test_degenerate_rays_bit_exact)."""
    for n, n_ch, fov, beta0 in ((64, 9, 64.0, 0.0), (65, 9, 65.0, 0.0), (64, 8, 64.0, np.pi / 4), (64, 9, 200.0, 0.0)):
        lab = _labels(n, 1, 4)
        vox = fov / n
        scan = make_scan(8, n_ch, 0.4, 60.0, 100.0, beta0=beta0)
        pj = Projector(scan, lab, (vox, vox, 0.1))
        sc = oracle.Scan(lab, (vox, vox, 0.1), 60.0, 100.0, pj.tables_host())
        for dtype in (F32, F64):
            ties = []
            _check_production_walk(pj, sc, np.arange(8 * n_ch), dtype, 4, (1, n, n), ties)
            assert len(ties) <= 8, ties                        # at most the central ray of each view


def test_degenerate_rays_bit_exact():
    """Rays along grid planes, through voxel corners, axis-parallel; even and odd grids; source inside the volume."""
    for n, n_ch, fov, beta0 in ((64, 9, 64.0, 0.0), (65, 9, 65.0, 0.0), (64, 8, 64.0, np.pi / 4), (64, 9, 200.0, 0.0)):
        lab = _labels(n, 1, 4)
        vox = fov / n
        # 8 views -> beta = k pi/4 (+beta0): the central channel of an odd detector is exactly axis-parallel/diagonal
        scan = make_scan(8, n_ch, 0.4, 60.0, 100.0, beta0=beta0)
        pj = Projector(scan, lab, (vox, vox, 0.1))
        sc = oracle.Scan(lab, (vox, vox, 0.1), 60.0, 100.0, pj.tables_host())
        ids = np.arange(8 * n_ch)
        for dtype in (F32, F64):
            for rid, (vx, lb, ln) in zip(ids, pj.trace(ids, max_seg=512, dtype=dtype)):
                w = sc.slab_trace(int(rid // n_ch), 0, int(rid % n_ch), dtype == F64)
                assert np.array_equal(vx, w[0]) and np.array_equal(lb, w[1]) and np.array_equal(ln.astype(np.float64), w[2])
        # and the float64 lengths are Siddon's -- except on rays that run ALONG a grid plane (within rounding of it), where
        # which of the two adjacent voxel rows the ray "is in" is a coin toss for any implementation, the paper's
        # midpoint rule included; there only the total is defined
        L = sc.slab_lengths(4, True)
        Lc, _ = sc.classic_lengths(4)
        assert np.max(np.abs(L.sum(-1) - Lc.sum(-1))) < 1e-10
        for v in range(8):
            for c in range(n_ch):
                S_, T_ = sc.endpoints(v, 0, c)
                grazing = any(abs(S_[a] - T_[a]) < 1e-9 and abs(S_[a] / vox + n / 2.0 - round(S_[a] / vox + n / 2.0)) < 1e-9
                              for a in (0, 1))
                if not grazing:
                    assert np.max(np.abs(L[v, 0, c] - Lc[v, 0, c])) < 1e-10, (n, v, c)


@pytest.mark.parametrize('dtype', [F32, F64])
@pytest.mark.parametrize('nz,n_rows,K', [(1, 1, 4), (16, 6, 4), (1, 1, 1), (1, 1, 7), (12, 3, 13), (1, 1, 20)])
def test_lengths_counts_logs_vs_siddon1985(dtype, nz, n_rows, K):
    pj, sc = _pair(nz=nz, n_rows=n_rows, K=K)
    mu, W = _tables(K)
    counts, logs, L = pj.project(mu, W, dtype=dtype, want_log=True, want_L=True)
    torch.cuda.synchronize()
    L_ref, seg = sc.classic_lengths(K)
    assert seg > 0
    assert scale_rel_err(L.cpu().numpy(), L_ref) < TOL[dtype]
    e_ray = per_ray_len_err(L.cpu().numpy(), L_ref)
    assert e_ray < TOL_RAY[dtype], f'per-ray length error {e_ray:.3e}'
    c_ref, l_ref, _ = sc.project_poly(mu, W)
    assert scale_rel_err(counts.cpu().numpy(), c_ref) < TOL[dtype]
    e_cnt = elem_rel_err(counts.cpu().numpy(), c_ref)
    assert e_cnt < TOL_ELEM[dtype], f'element-wise count error {e_cnt:.3e}'
    # the log sinogram is O(1..10): absolute error of tol x its scale
    assert scale_rel_err(logs.cpu().numpy(), l_ref) < (1e-9 if dtype == F64 else 3e-5)
    # detection from the stored lengths reproduces the fused epilogue
    c2, l2 = poly_detect(L, mu, W, dtype=dtype)
    assert scale_rel_err(c2.cpu().numpy(), counts.cpu().numpy()) < (1e-14 if dtype == F64 else 1e-6)


def test_view_ranges_and_spectrum_counts():
    pj, sc = _pair()
    mu, W4 = _tables(4, nS=4)
    full, _, _ = pj.project(mu, W4, dtype=F64, want_log=False)
    assert full.shape == (4, 36, 1, 120)
    a, _, _ = pj.project(mu, W4, view_begin=0, view_count=17, dtype=F64, want_log=False)
    b, _, _ = pj.project(mu, W4, view_begin=17, view_count=19, dtype=F64, want_log=False)
    assert torch.equal(torch.cat([a, b], dim=1), full)               # view sharding is exact
    one, _, _ = pj.project(mu, W4[2:3], dtype=F64, want_log=False)
    assert torch.equal(one[0], full[2])
    empty, _, _ = pj.project(mu, W4, view_begin=5, view_count=0, dtype=F64, want_log=False)
    assert empty.shape == (4, 0, 1, 120)


def test_rays_that_miss_see_air_exactly():
    # a 6.4 cm object under a 47 degree fan: the outer channels pass beside it
    pj, sc = _pair(n=64, fov=6.4)
    mu, W = _tables(4)
    counts, logs, L = pj.project(mu, W, dtype=F32, want_L=True)
    Lh = L.cpu().numpy()
    miss = Lh.sum(-1) == 0
    assert miss.any() and (~miss).any()
    air = torch.as_tensor(W, dtype=torch.float32).cuda().sum(1)
    c = counts.cpu().numpy()
    assert np.all(np.abs(c[:, miss] / air.cpu().numpy()[:, None] - 1) < 1e-6)
    assert np.all(np.abs(logs.cpu().numpy()[:, miss]) < 1e-6)
    L_ref, _ = sc.classic_lengths(4)
    assert np.array_equal(L_ref.sum(-1) == 0, miss)


def test_argument_errors_are_reported():
    pj, _ = _pair()
    mu, W = _tables(4)
    with pytest.raises(N.DexctError, match='view range'):
        pj.project(mu, W, view_begin=30, view_count=10)
    with pytest.raises(N.DexctError, match='nSpec'):
        pj.project(mu, np.tile(W, (3, 1)))
    with pytest.raises(ValueError):
        pj.project(mu, W[:, :-1])
    big = np.ones((200, 300))
    with pytest.raises(N.DexctError, match='shared memory'):
        pj.project(big, np.ones((1, 300)), dtype=F64)
    L = N.lib()
    assert L.dexct_siddon_poly(None, None, None, 0, 0, None, 1, 1, None, 1, None, None, None, 0, None) == -1
    # a label without a row in the attenuation table is refused (the kernels index their accumulators with the label)
    pj.labels[0, 5, 5] = 9
    pj.prepare()
    with pytest.raises(ValueError, match='label 9'):
        pj.project(mu, W)
    mu10 = np.concatenate([mu, np.tile(mu[-1:], (6, 1))])
    assert pj.project(mu10, W)[0].shape == (2, 36, 1, 120)
    pj.labels[0, 5, 5] = 1
    pj.prepare()
    assert torch.isfinite(pj.project(mu, W)[0]).all()
    # a cone so steep that a ray would cross more than one slice per column is refused, not traced wrongly (the slab
    # traversal assumes at most one z crossing per column): 5 micrometre slices under an 8-row detector
    scan = make_scan(12, 40, FAN, 60.0, 100.0, n_rows=8, h_iso=1.0)
    with pytest.raises(N.DexctError, match='cone too steep'):
        Projector(scan, _labels(32, 6, 4), (1.6, 1.6, 0.005)).project(mu, W)
    # ... while the same thin slices under a single-row detector are an ordinary 2-D scan
    scan1 = make_scan(12, 40, FAN, 60.0, 100.0, n_rows=1)
    assert torch.isfinite(Projector(scan1, _labels(32, 1, 4), (1.6, 1.6, 0.005)).project(mu, W)[0]).all()


# ---- full-size properties ---------------------------------------------------------------------------------------

def _box_chords(ct, half):
    """Analytic chord of every (view, channel) ray of a 2-D fan with the square [-half, half]^2 (float64 numpy)."""
    b = ct.betas[:, None]
    g = ct.gammas[None, :]
    S = np.stack([ct.SID * np.cos(b) + 0 * g, ct.SID * np.sin(b) + 0 * g])
    D = np.stack([-ct.SDD * np.cos(b + g), -ct.SDD * np.sin(b + g)])
    with np.errstate(divide='ignore', invalid='ignore'):
        t1, t2 = (-half - S) / D, (half - S) / D
    lo = np.minimum(t1, t2).max(0).clip(0, 1)
    hi = np.maximum(t1, t2).min(0).clip(0, 1)
    return np.maximum(hi - lo, 0) * ct.SDD


@pytest.mark.parametrize('dtype', [F32, F64])
def test_c2_full_size_chord_conservation_and_linearity(dtype):
    """Config C2 (512x512, 900 ch x 1000 views, 120 bins): sum_m L_m is the chord of the ray with the phantom's bounding
    square; doubling W doubles the counts exactly; a one-material phantom gives the closed-form spectrum integral."""
    ct, ph, lo, hi = S.make_config('C2')
    from xtomosim.forward_project import spectral_weights
    scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, rotation=ct.rotation)
    pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz))
    mu = ph.mu_table(lo.E)
    W = np.stack([spectral_weights(ct, lo), spectral_weights(ct, hi)])
    counts, logs, L = pj.project(mu, W, dtype=dtype, want_L=True)
    chord = _box_chords(ct, 25.6)
    Ls = L.sum(-1)[:, 0].cpu().numpy()
    assert np.max(np.abs(Ls - chord)) < (2e-4 if dtype == F32 else 1e-9)      # chords are up to 72 cm
    c2, _, _ = pj.project(mu, 2.0 * W, dtype=dtype, want_log=False)
    assert torch.equal(c2, 2.0 * counts)
    # checksum against the oracle on a view subset (the whole scan takes the CPU minutes)
    tabs = pj.tables_host()
    sub = {k: (v[::125] if k.endswith('beta') else v) for k, v in tabs.items()}
    sc = oracle.Scan(ph.M, (ph.dx, ph.dy, ph.dz), ct.SID, ct.SDD, sub)
    c_ref, l_ref, seg = sc.project_poly(mu, W)
    assert 500 < seg / (8 * 900) < 700
    assert scale_rel_err(counts[:, ::125].cpu().numpy(), c_ref) < TOL[dtype]
    e_cnt = elem_rel_err(counts[:, ::125].cpu().numpy(), c_ref)
    assert e_cnt < TOL_ELEM[dtype], f'element-wise count error {e_cnt:.3e}'
    L_ref, _ = sc.classic_lengths(S.K)
    e_ray = per_ray_len_err(L[::125].cpu().numpy(), L_ref)
    assert e_ray < TOL_RAY[dtype], f'per-ray length error {e_ray:.3e}'
    # homogeneous phantom: closed form
    uni = np.ones_like(ph.M)
    pj1 = Projector(scan, uni, (ph.dx, ph.dy, ph.dz))
    c1, _, _ = pj1.project(mu, W, dtype=dtype, want_log=False)
    want = np.einsum('se,vce->svc', W, np.exp(-chord[..., None] * mu[1]))
    assert scale_rel_err(c1[:, :, 0].cpu().numpy(), want) < (3e-5 if dtype == F32 else 1e-10)


def test_c4_slab_3d_parity_and_chords():
    """Config C4 geometry (512^3, 64 rows x 900 ch) on 6 of its 1200 views: lengths vs Siddon 1985 on a ray subset,
    chord conservation on all rays."""
    ct, ph, lo, hi = S.make_config('C4', n_views=6, nz=128)
    from xtomosim.forward_project import spectral_weights
    scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, n_rows=ct.N_rows, h_iso=ct.h_iso,
                     rotation=ct.rotation)
    pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz))
    mu = ph.mu_table(lo.E)
    W = np.stack([spectral_weights(ct, lo), spectral_weights(ct, hi)])
    tabs = pj.tables_host()
    sc = oracle.Scan(ph.M, (ph.dx, ph.dy, ph.dz), ct.SID, ct.SDD, tabs)
    for dtype in (F32, F64):
        counts, logs, L = pj.project(mu, W, dtype=dtype, want_L=True)
        L_ref, seg = sc.classic_lengths(S.K, v0=2, nv=1)
        assert scale_rel_err(L[2:3].cpu().numpy(), L_ref) < TOL[dtype]
        c_ref, _ = oracle.poly_detect(L_ref, mu, W, want_log=False)
        assert scale_rel_err(counts[:, 2].reshape(2, -1).cpu().numpy(), c_ref) < TOL[dtype]


@pytest.mark.parametrize('dtype', [F32, F64])
@pytest.mark.parametrize('h_iso,voxz', [(1.2, 51.2 / 96), (0.5, 0.2), (2.0, 1.0)])
def test_steep_cone_on_anatomy_like_volume(dtype, h_iso, voxz):
    """3-D skip layouts under stress: ellipsoidal body and inserts (long homogeneous runs in x, y AND z), wide cone angles
    and anisotropic voxels, so that the run caps of every z-homogeneity level come into play."""
    n, nz, n_rows, n_ch, n_views = 96, 48, 16, 110, 24
    lab = S.phantom_labels(n, nz)
    vox = 51.2 / n
    voxel = (vox, vox, voxz)
    scan = make_scan(n_views, n_ch, FAN, 60.0, 100.0, n_rows=n_rows, h_iso=h_iso, beta0=0.1)
    pj = Projector(scan, lab, voxel)
    sc = oracle.Scan(lab, voxel, 60.0, 100.0, pj.tables_host())
    mu, W = _tables(S.K)
    counts, logs, L = pj.project(mu, W, dtype=dtype, want_log=True, want_L=True)
    L_ref, seg = sc.classic_lengths(S.K)
    assert seg > 0
    assert scale_rel_err(L.cpu().numpy(), L_ref) < TOL[dtype]
    e_ray = per_ray_len_err(L.cpu().numpy(), L_ref)
    assert e_ray < TOL_RAY[dtype], f'per-ray length error {e_ray:.3e}'
    c_ref, l_ref, _ = sc.project_poly(mu, W)
    assert scale_rel_err(counts.cpu().numpy(), c_ref) < TOL[dtype]
    e_cnt = elem_rel_err(counts.cpu().numpy(), c_ref)
    assert e_cnt < TOL_ELEM[dtype], f'element-wise count error {e_cnt:.3e}'


@pytest.mark.parametrize('name', ['C4', 'C5'])
def test_c4_c5_full_size_properties(name):
    """Configs C4 (512^3, 64 rows x 900 ch x 1200 views) and C5 (1024^3, 128 x 1800 x 2400) at FULL volume and detector size,
    on view ranges spread over the rotation (the phantom is rasterised on the device; the CPU oracle traces one detector
    view of it): sum_m L_m is the chord of the ray with the volume's bounding box; the lengths of a view match Siddon 1985;
    doubling W doubles the counts exactly; a view range projected on its own equals the same views inside a larger range
    (what view-sharded ranks rely on)."""
    from xtomosim.forward_project import spectral_weights
    ct, ph, lo, hi = S.make_config(name, device='cuda')
    assert tuple(ph.M.shape) == (S.CONFIGS[name]['nz'], S.CONFIGS[name]['n'], S.CONFIGS[name]['n'])
    scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, n_rows=ct.N_rows, h_iso=ct.h_iso,
                     rotation=ct.rotation)
    pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz))
    mu = ph.mu_table(lo.E)
    W = np.stack([spectral_weights(ct, lo), spectral_weights(ct, hi)])
    tabs = pj.tables_host()
    half = np.array([ph.Nx * ph.dx, ph.Ny * ph.dy, ph.Nz * ph.dz]) / 2
    nv = 3
    for v0 in (0, ct.N_proj // 8 + 1, ct.N_proj // 2 + 5, ct.N_proj - nv):
        counts, logs, L = pj.project(mu, W, view_begin=v0, view_count=nv, want_L=True)
        # chords with the box, from the device's own angle tables
        cb, sb = tabs['cos_beta'][v0:v0 + nv, None, None], tabs['sin_beta'][v0:v0 + nv, None, None]
        cg, sg = tabs['cos_gamma'][None, None, :], tabs['sin_gamma'][None, None, :]
        z = tabs['zdet'][None, :, None]
        src = np.stack(np.broadcast_arrays(ct.SID * cb, ct.SID * sb, 0 * z + 0 * cg))
        d = np.stack(np.broadcast_arrays(-ct.SDD * (cb * cg - sb * sg), -ct.SDD * (sb * cg + cb * sg), z + 0 * cb + 0 * cg))
        with np.errstate(divide='ignore', invalid='ignore'):
            t1 = (-half[:, None, None, None] - src) / d
            t2 = (half[:, None, None, None] - src) / d
        t_lo = np.nanmax(np.minimum(t1, t2), axis=0).clip(0, 1)
        t_hi = np.nanmin(np.maximum(t1, t2), axis=0).clip(0, 1)
        chord = np.maximum(t_hi - t_lo, 0) * np.sqrt((d ** 2).sum(0))
        Ls = L.sum(-1).cpu().numpy()
        assert np.max(np.abs(Ls - chord)) < 3e-4, (name, v0, np.max(np.abs(Ls - chord)))
        c2, _, _ = pj.project(mu, 2.0 * W, view_begin=v0, view_count=nv, want_log=False)
        assert torch.equal(c2, 2.0 * counts)
        lo_v = max(0, v0 - 2)
        c_big, l_big, _ = pj.project(mu, W, view_begin=lo_v, view_count=min(nv + 4, ct.N_proj - lo_v))
        assert torch.equal(c_big[:, v0 - lo_v:v0 - lo_v + nv], counts) and torch.equal(l_big[:, v0 - lo_v:v0 - lo_v + nv], logs)
    # one view against Siddon 1985 on the host copy of the volume
    v = ct.N_proj // 8 + 2
    sub = {k: (a[v:v + 1] if k.endswith('beta') else a) for k, a in tabs.items()}
    sc = oracle.Scan(ph.M.cpu().numpy(), (ph.dx, ph.dy, ph.dz), ct.SID, ct.SDD, sub)
    L_ref, seg = sc.classic_lengths(S.K)
    counts, _, L = pj.project(mu, W, view_begin=v, view_count=1, want_L=True)
    assert scale_rel_err(L.cpu().numpy(), L_ref) < TOL[F32]
    c_ref, _ = oracle.poly_detect(L_ref, mu, W, want_log=False)
    assert scale_rel_err(counts.reshape(2, -1).cpu().numpy(), c_ref) < TOL[F32]
    per_ray = seg / (ct.N_rows * ct.N_channels)
    assert (500 < per_ray < 800) if name == 'C4' else (1000 < per_ray < 1600)


@pytest.mark.parametrize('dtype', [F32, F64])
@pytest.mark.parametrize('nz,n_rows,K', [(1, 1, 128), (1, 1, 129), (10, 4, 129), (1, 1, 256), (10, 4, 256)])
def test_up_to_256_materials(dtype, nz, n_rows, K):
    """Label values up to 255 (the uint8 phantom format's maximum).  The skip layouts keep 7 bits per label, so K <= 128 walks
    them and K > 128 falls back to the pair-layout kernel (one column per step); the lengths stay in shared memory instead of
    registers for K > 16.  Same numbers as Siddon 1985 either way."""
    pj, sc = _pair(nz=nz, n_rows=n_rows, K=K, n_views=12, seed=K)
    r = np.random.default_rng(K)
    E = np.linspace(30, 120, 24)
    mu = (0.01 + r.uniform(0, 0.4, (K, 1))) * (0.6 + (60.0 / E) ** r.uniform(1.0, 2.5, (K, 1)))
    mu[0] = 0.0002
    W = r.uniform(0.2, 1.0, (2, 24)) * 1e4
    counts, logs, L = pj.project(mu, W, dtype=dtype, want_log=True, want_L=True)
    L_ref, seg = sc.classic_lengths(K)
    assert seg > 0 and L_ref[..., K - 1].max() > 0                                 # the highest label is really met
    assert scale_rel_err(L.cpu().numpy(), L_ref) < TOL[dtype]
    e_ray = per_ray_len_err(L.cpu().numpy(), L_ref)
    assert e_ray < TOL_RAY[dtype], f'per-ray length error {e_ray:.3e}'
    c_ref, l_ref, _ = sc.project_poly(mu, W)
    assert scale_rel_err(counts.cpu().numpy(), c_ref) < TOL[dtype]
    e_cnt = elem_rel_err(counts.cpu().numpy(), c_ref)
    assert e_cnt < TOL_ELEM[dtype], f'element-wise count error {e_cnt:.3e}'
    assert scale_rel_err(logs.cpu().numpy(), l_ref) < (1e-9 if dtype == F64 else 3e-5)
    c2, l2 = poly_detect(L, mu, W, dtype=dtype)
    assert scale_rel_err(c2.cpu().numpy(), counts.cpu().numpy()) < (1e-14 if dtype == F64 else 1e-6)


@pytest.mark.parametrize('dtype', [F32, F64])
@pytest.mark.parametrize('shape,voxel,n_rows', [((1, 70, 120), (0.40, 0.55, 0.1), 1), ((1, 120, 70), (0.60, 0.35, 0.1), 1),
                                               ((9, 120, 70), (0.55, 0.30, 0.7), 5), ((14, 50, 131), (0.25, 0.62, 0.4), 6)])
def test_rectangular_volumes_and_anisotropic_voxels(dtype, shape, voxel, n_rows):
    """nx != ny and dx != dy != dz: the two pair / skip layouts have different pitches and column counts, and which in-plane
    axis is the major one depends on the slope in INDEX units.  Blobby labels (long homogeneous runs for the skip layouts)
    with a noisy band across them; lengths, counts and logs against Siddon 1985."""
    nz, ny, nx = shape
    r = np.random.default_rng(nx * 1000 + ny)
    coarse = r.integers(0, 6, (nz // 4 + 1, ny // 9 + 1, nx // 9 + 1))
    lab = np.repeat(np.repeat(np.repeat(coarse, 4, 0), 9, 1), 9, 2)[:nz, :ny, :nx].astype(np.uint8)
    lab[:, ny // 3:ny // 3 + 5, :] = r.integers(0, 6, (nz, 5, nx))
    scan = make_scan(30, 140, FAN, 60.0, 100.0, n_rows=n_rows, h_iso=0.35, beta0=0.05)
    pj = Projector(scan, lab, voxel)
    sc = oracle.Scan(lab, voxel, 60.0, 100.0, pj.tables_host())
    mu, W = _tables(6)
    counts, logs, L = pj.project(mu, W, dtype=dtype, want_log=True, want_L=True)
    L_ref, seg = sc.classic_lengths(6)
    assert seg > 0
    assert scale_rel_err(L.cpu().numpy(), L_ref) < TOL[dtype]
    e_ray = per_ray_len_err(L.cpu().numpy(), L_ref)
    assert e_ray < TOL_RAY[dtype], f'per-ray length error {e_ray:.3e}'
    c_ref, l_ref, _ = sc.project_poly(mu, W)
    assert scale_rel_err(counts.cpu().numpy(), c_ref) < TOL[dtype]
    e_cnt = elem_rel_err(counts.cpu().numpy(), c_ref)
    assert e_cnt < TOL_ELEM[dtype], f'element-wise count error {e_cnt:.3e}'
    assert scale_rel_err(logs.cpu().numpy(), l_ref) < (1e-9 if dtype == F64 else 3e-5)
    # a view range on its own equals the same views inside the whole scan
    c_part, _, _ = pj.project(mu, W, view_begin=7, view_count=9, dtype=dtype)
    assert torch.equal(c_part, counts[:, 7:16])


@pytest.mark.parametrize('dtype', [F32, F64])
@pytest.mark.parametrize('n,nz,n_ch,n_views,n_rows,h_iso', [(1, 1, 1, 1, 1, 0.4), (2, 1, 3, 5, 1, 0.4), (33, 1, 7, 3, 4, 0.09),
                                                             (17, 3, 33, 2, 1, 0.4), (40, 5, 1, 9, 2, 0.3), (64, 2, 40, 6, 3, 0.2)])
def test_tiny_and_lopsided_sizes(dtype, n, nz, n_ch, n_views, n_rows, h_iso):
    """One voxel, one ray, one view; a single slice under a multi-row detector (rays leave the slab through its faces);
    odd plane counts under a single row in the mid-plane; fewer channels than a warp."""
    pj, sc = _pair(n=n, n_ch=n_ch, n_views=n_views, n_rows=n_rows, nz=nz, K=5, fov=25.6 if n > 2 else 4.0, h_iso=h_iso,
                   beta0=0.3, seed=n)
    mu, W = _tables(5)
    counts, logs, L = pj.project(mu, W, dtype=dtype, want_log=True, want_L=True)
    assert counts.shape == (2, n_views, n_rows, n_ch) and L.shape == (n_views, n_rows, n_ch, 5)
    L_ref, seg = sc.classic_lengths(5)
    assert scale_rel_err(L.cpu().numpy(), L_ref) < TOL[dtype]
    e_ray = per_ray_len_err(L.cpu().numpy(), L_ref)
    assert e_ray < TOL_RAY[dtype], f'per-ray length error {e_ray:.3e}'
    c_ref, l_ref, _ = sc.project_poly(mu, W)
    assert scale_rel_err(counts.cpu().numpy(), c_ref) < TOL[dtype]
    e_cnt = elem_rel_err(counts.cpu().numpy(), c_ref)
    assert e_cnt < TOL_ELEM[dtype], f'element-wise count error {e_cnt:.3e}'
    # (absolute where the log sinogram is ~0: the outer rows of case 33-1-7-3-4 miss the 1 mm slab, the inner ones leave it half way)
    assert np.max(np.abs(logs.cpu().numpy() - l_ref)) < (1e-9 if dtype == F64 else 3e-5) * max(1.0, np.abs(l_ref).max())
