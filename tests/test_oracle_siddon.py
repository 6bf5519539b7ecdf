"""Known-answer tests of the Siddon oracle (oracle/siddon_oracle.c).  The reference holds no vectors for this path
(parity unpinned, see the oracle header), so the anchors are analytic: exact chord lengths through uniform blocks,
conservation of total path length, the monoenergetic closed form, and agreement of the two independent traversals."""
import numpy as np
import pytest

import oracle

VOX = (0.1, 0.1, 0.1)


def test_axis_aligned_and_diagonal_exact_lengths():
    lab = np.ones((1, 64, 64), np.uint8)
    L, n = oracle.classic_lengths_ray([-10, 0.05, 0], [10, 0.05, 0], lab, VOX, 2)
    assert n == 64 and L[0] == 0 and abs(L[1] - 6.4) < 1e-12
    L, n = oracle.classic_lengths_ray([0.05, 10, 0], [0.05, -10, 0], lab, VOX, 2)
    assert n == 64 and abs(L[1] - 6.4) < 1e-12
    # exact 45 degree diagonal through voxel corners: 64 voxels, zero-length corner touches are dropped
    L, n = oracle.classic_lengths_ray([-10, -10, 0], [10, 10, 0], lab, VOX, 2)
    assert n == 64 and abs(L[1] - 6.4 * np.sqrt(2)) < 1e-12


def test_two_material_split_and_ray_inside_volume():
    lab = np.zeros((1, 32, 32), np.uint8)
    lab[:, :, 16:] = 1                       # right half material 1
    L, _ = oracle.classic_lengths_ray([-5, 0.03, 0], [5, 0.03, 0], lab, VOX, 2)
    assert abs(L[0] - 1.6) < 1e-12 and abs(L[1] - 1.6) < 1e-12
    # source inside the volume: alpha clamps to [0, 1]
    L, _ = oracle.classic_lengths_ray([-0.5, 0.03, 0], [5, 0.03, 0], lab, VOX, 2)
    assert abs(L[0] - 0.5) < 1e-12 and abs(L[1] - 1.6) < 1e-12
    # ray that misses
    L, n = oracle.classic_lengths_ray([-5, 3.0, 0], [5, 3.0, 0], lab, VOX, 2)
    assert n == 0 and L.sum() == 0
    # ray exactly on the top face (half-open volume): misses; exactly on the bottom face: inside
    L, n = oracle.classic_lengths_ray([-5, 1.6, 0], [5, 1.6, 0], lab, VOX, 2)
    assert n == 0
    L, n = oracle.classic_lengths_ray([-5, -1.6, 0], [5, -1.6, 0], lab, VOX, 2)
    assert n == 32 and abs(L.sum() - 3.2) < 1e-12


def test_3d_oblique_total_length_equals_box_chord(rng):
    lab = rng.integers(0, 5, (24, 40, 56)).astype(np.uint8)
    vox = (0.1, 0.07, 0.13)
    half = 0.5 * np.array([56 * vox[0], 40 * vox[1], 24 * vox[2]])
    for _ in range(200):
        S = rng.uniform(-1, 1, 3) * half * 3
        T = rng.uniform(-1, 1, 3) * half * 3
        L, n = oracle.classic_lengths_ray(S, T, lab, vox, 5)
        # slab-method chord of the segment S->T with the box
        d = T - S
        with np.errstate(divide='ignore', invalid='ignore'):
            t1, t2 = (-half - S) / d, (half - S) / d
        lo, hi = np.minimum(t1, t2).max(), np.maximum(t1, t2).min()
        lo, hi = max(lo, 0.0), min(hi, 1.0)
        chord = max(hi - lo, 0.0) * np.linalg.norm(d)
        assert abs(L.sum() - chord) < 1e-10


def _scan(n=96, n_ch=120, n_views=36, n_rows=1, nz=1, seed=3):
    r = np.random.default_rng(seed)
    lab = r.integers(0, 4, (nz, n, n)).astype(np.uint8)
    lab[:, n // 4:n // 2, n // 3:2 * n // 3] = 2
    t = oracle.fan_tables(n_ch, n_views, 0.8230337, N_rows=n_rows, h_iso=0.4, SID=60.0, SDD=100.0)
    vox = 51.2 / n
    return oracle.Scan(lab, (vox, vox, vox if nz > 1 else 0.1), 60.0, 100.0, t)


@pytest.mark.parametrize('nz,n_rows', [(1, 1), (16, 6)])
def test_slab_twin_agrees_with_siddon_1985(nz, n_rows):
    sc = _scan(nz=nz, n_rows=n_rows)
    L, seg = sc.classic_lengths(4)
    assert seg > 0
    L64 = sc.slab_lengths(4, f64=True)
    L32 = sc.slab_lengths(4, f64=False)
    assert np.max(np.abs(L64 - L)) < 1e-11
    assert np.max(np.abs(L32 - L)) < 2e-5 * L.sum(-1).max()
    # ordered traces: identical voxel sequences in float64 on generic rays
    r = np.random.default_rng(0)
    for _ in range(300):
        v, row, c = r.integers(sc.n_views), r.integers(sc.n_rows), r.integers(sc.n_ch)
        a = sc.classic_trace(v, row, c)
        b = sc.slab_trace(v, row, c, True)
        assert len(a[0]) == len(b[0]) and np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
        if len(a[0]):
            assert np.max(np.abs(a[2] - b[2])) < 1e-10


def test_monoenergetic_closed_form_and_air_normalisation():
    sc = _scan()
    L, _ = sc.classic_lengths(4)
    mu = np.array([[0.0], [0.2], [0.5], [0.3]])
    W = np.array([[1000.0]])
    counts, logs = oracle.poly_detect(L, mu, W)
    line = (L.reshape(-1, 4) * mu[:, 0]).sum(-1)
    assert np.allclose(counts[0], 1000.0 * np.exp(-line), rtol=1e-14)
    assert np.allclose(logs[0], line, atol=1e-12)
    c2, l2, seg = sc.project_poly(mu, W)
    assert np.array_equal(c2.reshape(-1), counts[0])


def test_segment_count_matches_survey_probe():
    """about 620 segments per ray on the 512 grid with the params.txt geometry."""
    lab = np.zeros((1, 512, 512), np.uint8)
    t = oracle.fan_tables(900, 8, 0.8230337)
    sc = oracle.Scan(lab, VOX, 60.0, 100.0, t)
    _, seg = sc.classic_lengths(1)
    assert 600 < seg / (8 * 900) < 660


@pytest.mark.parametrize('shape,voxel,n_rows,K', [((1, 70, 120), (0.40, 0.55, 0.1), 1, 6), ((9, 120, 70), (0.55, 0.30, 0.7), 5, 6),
                                                 ((14, 50, 131), (0.25, 0.62, 0.4), 6, 256)])
def test_two_traversals_agree_on_rectangular_anisotropic_volumes(shape, voxel, n_rows, K):
    """The checker's own consistency on the shapes tests/test_gpu_siddon.py uses for the device: nx != ny, dx != dy != dz and the
    full uint8 label range.  Siddon 1985 (A) and the slab traversal (B) give the same lengths, total = chord of the box."""
    nz, ny, nx = shape
    r = np.random.default_rng(nx * 1000 + ny)
    lab = r.integers(0, K, shape).astype(np.uint8)
    t = oracle.fan_tables(140, 30, 0.8230337, N_rows=n_rows, h_iso=0.35, SID=60.0, SDD=100.0)
    sc = oracle.Scan(lab, voxel, 60.0, 100.0, t)
    L, seg = sc.classic_lengths(K)
    assert seg > 0 and L.shape[-1] == K
    L64 = sc.slab_lengths(K, f64=True)
    assert np.max(np.abs(L64 - L)) < 1e-10
    assert np.max(np.abs(sc.slab_lengths(K, f64=False) - L)) < 2e-5 * L.sum(-1).max()
    half = np.array([nx * voxel[0], ny * voxel[1], nz * voxel[2]]) / 2
    for v, row, c in [(0, 0, 0), (7, n_rows - 1, 70), (29, n_rows // 2, 139), (13, 0, 33)]:
        S_, T_ = sc.endpoints(v, row, c)
        d = T_ - S_
        with np.errstate(divide='ignore', invalid='ignore'):
            t1, t2 = (-half - S_) / d, (half - S_) / d
        lo, hi = np.nanmax(np.minimum(t1, t2)).clip(0, 1), np.nanmin(np.maximum(t1, t2)).clip(0, 1)
        assert abs(L[v, row, c].sum() - max(hi - lo, 0.0) * np.linalg.norm(d)) < 1e-9
