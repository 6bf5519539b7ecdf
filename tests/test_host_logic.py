"""Host-side logic of the drop-in modules (no GPU): data objects, parameter file, material tables, the spectral tables of
the decomposition against the oracle restatement, view sharding with a 2-rank gloo group."""
import json
import os
import types

import numpy as np
import pytest
import torch

from dexct import dist as ddist
from dexct import synthetic as S
from oracle import gn_oracle
from xtomosim import xcompy as xc
from xtomosim.system import (FanBeamGeometry, ScannerGeometry, VoxelPhantom, read_parameter_file, write_materials,
                             xRaySpectrum)


def test_spectrum_and_detector_files(golden_dir):
    s = xRaySpectrum(os.path.join(golden_dir, 'input/spectrum/80kV_1mGy_float32.bin'), '80kV')
    assert s.E.shape == (140,) and s.E[0] == 1 and s.E[-1] == 140 and s.I0.max() > 1e10
    before = s.I0.copy()
    s.rescale_counts(0.5)
    assert np.array_equal(s.I0, before * 0.5)
    mv = xRaySpectrum(os.path.join(golden_dir, 'input/spectrum/detunedMV_1mGy_float32.bin'), 'detunedMV')
    assert mv.E.shape == (100,) and abs(mv.E[-1] - 6000) < 1e-3
    ct = FanBeamGeometry(N_channels=800, N_proj=1200, gamma_fan=0.8230337, SID=60.0, SDD=100.0, h_iso=1.0, eid=True,
                         detector_file=os.path.join(golden_dir, 'input/detector/eta_eid_mv.bin'))
    assert ScannerGeometry is FanBeamGeometry
    assert ct.det_E.shape == (6000,) and ct.det_eta_E.max() == 1.0
    assert abs(ct.A_iso - 60.0 * 0.8230337 / 800 * 1.0) < 1e-15          # main.py:68 uses ct.A_iso, ct.N_proj
    assert ct.N_proj == 1200 and ct.eid


def test_xcompy_standin():
    assert xc.parse_matcomp('H(11.2)O(88.8)') == [('H', pytest.approx(0.112)), ('O', pytest.approx(0.888))]
    with pytest.raises(ValueError):
        xc.parse_matcomp('water')
    E = np.array([30.0, 60.0, 100.0])
    w = xc.mixatten('H(11.2)O(88.8)', E)
    assert np.allclose(w, [0.3756, 0.2059, 0.1707], rtol=0.02)          # published NIST values
    assert (np.diff(w) < 0).all()
    import xcompy as top                                                 # matdecomp.py:7 imports it top-level
    assert np.array_equal(top.mixatten('Ca(100)', E), xc.mixatten('Ca(100)', E))
    xc.register_element_table('Ca', [10, 100], [90.0, 0.25])
    try:
        assert abs(xc.mixatten('Ca(100)', np.array([100.0]))[0] - 0.25) < 1e-12
    finally:
        xc._tables.pop('Ca')


def test_phantom_and_parameter_file(tmp_path, golden_dir):
    lab = S.phantom_labels(64)
    (tmp_path / 'ph.bin').write_bytes(lab.tobytes())
    write_materials(tmp_path / 'mats.csv', S.MATERIALS)
    ph = VoxelPhantom('p', str(tmp_path / 'ph.bin'), str(tmp_path / 'mats.csv'), 64, 64, 1, z_index=0)
    assert ph.M.shape == (1, 64, 64) and ph.K == 8 and np.array_equal(ph.M, lab)
    mu = ph.mu_table(np.array([60.0, 80.0]))
    assert mu.shape == (8, 2) and mu[2, 0] > mu[1, 0] > mu[0, 0] >= 0
    hu = ph.M_mono(80).get()                                             # plots.py:251
    assert hu.shape == (64, 64) and hu.min() < -900                      # air
    with pytest.raises(ValueError):
        VoxelPhantom('p', str(tmp_path / 'ph.bin'), str(tmp_path / 'mats.csv'), 32, 32, 1)
    params = {"RUN_ID": "t", "forward_project": True, "back_project": True, "phantom_type": "voxel", "phantom_id": "p",
              "phantom_filename": str(tmp_path / 'ph.bin'), "matcomp_filename": str(tmp_path / 'mats.csv'),
              "Nx": 64, "Ny": 64, "Nz": 1, "dx": 0.8, "dy": 0.8, "dz": 0.8, "z_index": 0,
              "scanner_geometry": "fan_beam", "SID": 60.0, "SDD": 100.0, "N_channels": 100, "N_projections": 90,
              "fan_angle_total": 0.8230337, "rotation_angle_total": 6.283185, "detector_px_height": 1.0,
              "detector_mode": "eid", "detector_filename": os.path.join(golden_dir, 'input/detector/eta_eid_mv.bin'),
              "spectrum_id": "NA", "spectrum_filename": "NA", "N_photons_per_cm2_per_scan": "NA",
              "N_recon_matrix": 64, "FOV_recon": 50.0, "ramp_filter_percent_Nyquist": 0.8}
    (tmp_path / 'params.txt').write_text(json.dumps(params))
    (run_id, do_fp, do_bp, ct, phantom, spec, N_matrix, FOV, ramp), = read_parameter_file(str(tmp_path / 'params.txt'))
    assert (run_id, do_fp, do_bp, spec, N_matrix, FOV, ramp) == ('t', True, True, None, 64, 50.0, 0.8)
    assert ct.N_channels == 100 and ct.N_proj == 90 and ct.eid and phantom.dx == 0.8


def test_spectral_tables_match_oracle_restatement(golden_dir):
    """matdecomp.spectral_tables (product, host side) == oracle gn_model_tables == what the reference built (golden)."""
    import matdecomp
    d = np.load(os.path.join(golden_dir, 'gn_c1_tables.npz'))
    ct = types.SimpleNamespace(det_E=d['det_E'], det_eta_E=d['det_eta_E'], eid=bool(d['eid']))
    s1 = types.SimpleNamespace(E=d['spec1_E'], I0=d['spec1_I0'])
    s2 = types.SimpleNamespace(E=d['spec2_E'], I0=d['spec2_I0'])
    ee, i0, mus = matdecomp.spectral_tables(ct, s1, s2)
    assert np.array_equal(ee, d['ee']) and np.array_equal(i0, d['i0']) and np.array_equal(mus, d['mus'])
    assert (matdecomp.mat1, matdecomp.matcomp1, matdecomp.density1) == (gn_oracle.mat1, gn_oracle.matcomp1, gn_oracle.density1)
    assert (matdecomp.mat2, matdecomp.matcomp2, matdecomp.density2) == (gn_oracle.mat2, gn_oracle.matcomp2, gn_oracle.density2)


def test_scan_tables_match_oracle_conventions():
    import oracle
    from dexct.ops import make_scan
    s = make_scan(1000, 900, 0.8230337, 60.0, 100.0, n_rows=4, h_iso=0.1, rotation=6.283185, beta0=0.25)
    t = oracle.fan_tables(900, 1000, 0.8230337, rotation_angle_total=6.283185, beta0=0.25, N_rows=4, h_iso=0.1)
    v, c, r = np.arange(1000), np.arange(900), np.arange(4)
    assert np.allclose(np.cos(s.beta0 + v * s.dbeta), t['cos_beta'], atol=1e-15)
    assert np.allclose(np.sin(s.gamma0 + c * s.dgamma), t['sin_gamma'], atol=1e-15)
    assert np.allclose(s.z0 + r * s.dzdet, t['zdet'], atol=1e-15)


def test_shard_range_partitions():
    for n in (0, 1, 7, 1000, 1201):
        for ws in (1, 2, 3, 8):
            parts = [ddist.shard_range(n, r, ws) for r in range(ws)]
            assert parts[0][0] == 0 and sum(c for _, c in parts) == n
            assert all(parts[i][0] + parts[i][1] == parts[i + 1][0] for i in range(ws - 1))
            assert max(c for _, c in parts) - min(c for _, c in parts) <= 1
    with pytest.raises(ValueError):
        ddist.shard_range(10, 2, 2)


def _gloo_worker(rank, ws, port, n_views, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=ws)
    try:
        full = torch.arange(n_views * 3 * 5, dtype=torch.float32).reshape(n_views, 3, 5)
        v0, nv = ddist.shard_range(n_views, rank, ws)
        got = ddist.all_gather_views(full[v0:v0 + nv].clone(), n_views, dim=0)
        ok = torch.equal(got, full)
        # gather along a non-leading axis too ([nSpec, views, ch] layout)
        full2 = full.permute(1, 0, 2).contiguous()
        got2 = ddist.all_gather_views(full2[:, v0:v0 + nv].clone(), n_views, dim=1)
        # a set-up step that failed on one rank only is seen by all of them
        ok = ok and ddist.all_agree(True) and not ddist.all_agree(rank != 1) and not ddist.all_agree(False)
        q.put((rank, ok and torch.equal(got2, full2)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('n_views', [8, 7])
def test_all_gather_views_gloo_world2(n_views):
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + (os.getpid() + n_views) % 400
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, n_views, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]


def test_all_agree_without_a_process_group():
    assert ddist.all_agree(True) and not ddist.all_agree(False) and not ddist.all_agree(0)


def test_spectral_tables_memo_follows_its_inputs():
    """The decomposition tables are memoised on the CONTENT of their inputs: the same content returns equal (separate)
    arrays, a changed spectrum or a changed attenuation provider returns new ones."""
    import matdecomp
    import xcompy
    from dexct import synthetic as S
    ct, ph, lo, hi = S.make_config('C2', n_views=4, n=32)
    a = matdecomp.spectral_tables(ct, lo, hi)
    b = matdecomp.spectral_tables(ct, lo, hi)
    assert all(np.array_equal(x, y) and x is not y for x, y in zip(a, b))
    b[1][...] = 0                                              # the caller's copy is its own
    assert np.array_equal(matdecomp.spectral_tables(ct, lo, hi)[1], a[1])
    lo.I0 = lo.I0 * 2.0
    c = matdecomp.spectral_tables(ct, lo, hi)
    assert np.allclose(c[1][0], 2.0 * a[1][0]) and np.array_equal(c[1][1], a[1][1])
    try:
        xcompy.register_element_table('Ca', [1.0, 1000.0], [1.0, 1.0])
        d = matdecomp.spectral_tables(ct, lo, hi)
        assert not np.array_equal(d[2], c[2])                  # bone contains calcium
    finally:
        import xtomosim.xcompy as impl
        impl._tables.pop('Ca', None)
        impl.generation += 1
    assert np.array_equal(matdecomp.spectral_tables(ct, lo, hi)[2], c[2])


def test_cone_recurrence_equals_runs_then_diagonal_minimum():
    """The skip layouts' run length is defined by the cone recurrence S(i,j) = 1 + min(S'(i+1,j), S'(i+1,j+1)); the device
    builds it as R = run of homogeneous words of one label along the columns, then S(i,j) = min(R(i,j), 1 + S'(i+1,j+1))
    along diagonals (csrc/siddon.cu, skip_runs_kernel / skip_cone_kernel).  Same numbers, on random and on blobby images."""
    rng = np.random.default_rng(1)

    def cone(lo, hi):
        npc, ne = lo.shape
        S = np.zeros((npc, ne), int)
        pS, pL = np.zeros(ne + 1, int), np.full(ne + 1, 255)
        for i in range(npc - 1, -1, -1):
            cS, cL = np.zeros(ne + 1, int), np.full(ne + 1, 255)
            for j in range(ne):
                if lo[i, j] == hi[i, j]:
                    sa = pS[j] if pL[j] == lo[i, j] else 0
                    sb = pS[j + 1] if pL[j + 1] == lo[i, j] else 0
                    S[i, j] = cS[j] = min(1 + min(sa, sb), 127)
                    cL[j] = lo[i, j]
            pS, pL = cS, cL
        return S

    def two_pass(lo, hi):
        npc, ne = lo.shape
        hom = lo == hi
        R = np.zeros((npc + 1, ne + 1), int)
        S = np.zeros((npc + 1, ne + 1), int)
        for i in range(npc - 1, -1, -1):
            for j in range(ne):
                if hom[i, j]:
                    ok = i + 1 < npc and hom[i + 1, j] and lo[i + 1, j] == lo[i, j]
                    R[i, j] = min(1 + (R[i + 1, j] if ok else 0), 127)
        for i in range(npc - 1, -1, -1):
            for j in range(ne):
                if hom[i, j]:
                    ok = i + 1 < npc and j + 1 < ne and hom[i + 1, j + 1] and lo[i + 1, j + 1] == lo[i, j]
                    S[i, j] = min(R[i, j], 1 + (S[i + 1, j + 1] if ok else 0))
        return S[:npc, :ne]

    for trial in range(12):
        npc, ne = int(rng.integers(3, 36)), int(rng.integers(3, 36))
        img = rng.integers(0, 3, (npc, ne + 1))
        if trial % 2:
            img = np.repeat(np.repeat(rng.integers(0, 3, (npc // 4 + 1, ne // 4 + 2)), 4, 0), 4, 1)[:npc, :ne + 1]
        assert np.array_equal(cone(img[:, :-1], img[:, 1:]), two_pass(img[:, :-1], img[:, 1:]))


def test_phantom_digest_sees_every_voxel():
    """The workspace cache of get_sino is keyed on the WHOLE label volume (host arrays of any size and device tensors):
    a one-voxel in-place edit, or moving a block of voxels by a word, changes the key."""
    import torch
    from xtomosim.forward_project import _content_digest
    rng = np.random.default_rng(3)
    a = rng.integers(0, 8, size=(72, 1024, 1024), dtype=np.uint8)          # 72 MiB: the multi-piece path
    d = _content_digest(a)
    assert _content_digest(a.copy()) == d
    b = a.copy()
    b[40, 517, 4099 % 1024] ^= 1
    assert _content_digest(b) != d
    c = a.copy()
    c[0, 0, :8], c[0, 0, 8:16] = a[0, 0, 8:16], a[0, 0, :8]
    assert (c != a).any() and _content_digest(c) != d
    small = a[:1, :64, :64]
    assert _content_digest(small) == _content_digest(np.ascontiguousarray(small))
    t = torch.from_numpy(a[:2].copy())
    dt = _content_digest(t)
    assert _content_digest(t.clone()) == dt
    u = t.clone()
    u[1, 1000, 1000] ^= 1
    assert _content_digest(u) != dt
    w = t.clone()
    w[0, 0, :8], w[0, 0, 8:16] = t[0, 0, 8:16], t[0, 0, :8]
    assert _content_digest(w) != dt
