"""The reference's own driver script, UNCHANGED, against the drop-in modules -- and with it configuration C1 at its real
shape (BASELINE.json configs[0]; /root/reference/input/params.txt:18-35: 800 channels x 1200 views, fan 0.8230337 rad,
energy-integrating detector file eta_eid_mv.bin, recon 512^2 / FOV 50 / ramp 0.8; main.py:101,153: detunedMV + 80 kV at
doses 9 and 1, 50 Newton iterations, four reconstructions).

``tests/golden/reference_main.py`` is a byte-for-byte copy of /root/reference/main.py (test fixture, like the .bin inputs
beside it; the copy is checked against the mounted reference where that exists).  The test builds the working directory
main.py expects (``./input/params.txt`` = the reference's own parameter file, its spectrum and detector files, and -- because
the XCAT phantom is licensed and not shipped -- a synthetic 512 x 512 uint8 phantom + material table at the paths params.txt
names), puts ``dex-ct-sim_b200/`` and the headless matplotlib stand-in on ``sys.path`` and ``runpy``s the script with
``show_imgs`` left at True.  It then checks the .bin outputs of main.py:121-122,135-136,154-155,169 and compares a 32-view
sample of the decomposition with the oracle (oracle/gn_oracle.py, the pinned restatement of matdecomp.py:87-127).
"""
import filecmp
import json
import os
import runpy
import shutil
import sys
import time

import numpy as np
import pytest

from conftest import PKG, REFERENCE, ROOT, scale_rel_err
from dexct import synthetic as S

pytestmark = pytest.mark.gpu

MAIN = os.path.join(ROOT, 'tests', 'golden', 'reference_main.py')
SHIMS = os.path.join(ROOT, 'tests', 'shims')


def _workdir(tmp_path, golden_dir):
    inp = tmp_path / 'input'
    shutil.copytree(os.path.join(golden_dir, 'input'), inp)                  # params.txt, spectrum/, detector/ (reference's own)
    params = json.loads((inp / 'params.txt').read_text())
    ph_path = tmp_path / params['phantom_filename']
    os.makedirs(ph_path.parent, exist_ok=True)
    n = params['Nx']
    lab = S.phantom_labels(n)
    ph_path.write_bytes(lab.tobytes())
    from xtomosim.system import write_materials
    write_materials(tmp_path / params['matcomp_filename'], [dict(m) for m in S.MATERIALS])
    return params


@pytest.mark.skipif(not os.path.exists(os.path.join(REFERENCE, 'main.py')), reason='reference tree not mounted')
def test_fixture_is_the_reference_file():
    assert filecmp.cmp(MAIN, os.path.join(REFERENCE, 'main.py'), shallow=False)
    for rel in ('params.txt', 'detector/eta_eid_mv.bin', 'spectrum/detunedMV_1mGy_float32.bin', 'spectrum/80kV_1mGy_float32.bin'):
        assert filecmp.cmp(os.path.join(ROOT, 'tests/golden/input', rel), os.path.join(REFERENCE, 'input', rel), shallow=False)


def test_reference_main_py_runs_unchanged(tmp_path, golden_dir, capsys):
    params = _workdir(tmp_path, golden_dir)
    from dexct import config
    from xtomosim import forward_project as FP
    old_cwd, old_path, old_dtype = os.getcwd(), list(sys.path), config.dtype
    for name in [m for m in sys.modules if m == 'matplotlib' or m.startswith('matplotlib.')]:
        del sys.modules[name]                                                 # (none on the GPU box; the shim must win)
    sys.path[:0] = [SHIMS, PKG]
    os.chdir(tmp_path)
    try:
        config.dtype = 'f32'
        FP.clear_cache()
        t0 = time.perf_counter()
        runpy.run_path(MAIN, run_name='__main__')
        wall = time.perf_counter() - t0
    finally:
        os.chdir(old_cwd)
        sys.path[:] = old_path
        config.dtype = old_dtype
    out = capsys.readouterr().out
    assert 'matdecomp finished for detunedMV-80kV' in out                    # main.py:178

    nv, nc, nm = params['N_projections'], params['N_channels'], params['N_recon_matrix']
    run = tmp_path / 'output' / params['RUN_ID']
    assert (run / 'params.txt').read_text() == (tmp_path / 'input' / 'params.txt').read_text()       # main.py:98

    def rd(rel, shape):
        a = np.fromfile(run / rel, dtype=np.float32)
        assert a.size == int(np.prod(shape)), rel
        return a.reshape(shape).astype(np.float64)

    sinos = {}
    for sub in ('detunedMV_9000uGy', '80kV_1000uGy'):                         # main.py:110
        raw, log = rd(f'{sub}/sino_raw_float32.bin', (nv, nc)), rd(f'{sub}/sino_log_float32.bin', (nv, nc))    # :121-122
        assert np.isfinite(raw).all() and (raw > 0).all() and (log > -1e-5).all()
        sinos[sub] = raw
        rec, hu = rd(f'{sub}/recon_raw_float32.bin', (nm, nm)), rd(f'{sub}/recon_HU_float32.bin', (nm, nm))    # :135-136
        assert np.isfinite(rec).all() and np.isfinite(hu).all()
        tissue = rec[nm // 2 - 8:nm // 2 + 8, 120:136].mean()                 # inside the body ellipse, 50 cm FOV over 512 px
        assert tissue > 0.01
    md = 'matdecomp_detunedMV_80kV_9000uGy_1000uGy'
    m1, m2 = rd(f'{md}/mat1_sino_float32.bin', (nv, nc)), rd(f'{md}/mat2_sino_float32.bin', (nv, nc))          # :154-155
    assert np.isfinite(m1).all() and np.isfinite(m2).all()
    for i in (1, 2):
        assert np.isfinite(rd(f'{md}/mat{i}_recon_float32.bin', (nm, nm))).all()                              # :169
    s1, s2 = sinos['detunedMV_9000uGy'], sinos['80kV_1000uGy']
    air = s1 >= 0.95 * s1.max()                                               # matdecomp.py:195-196 (float32 files: ties aside)
    assert (m1[air & (s1 >= 0.951 * s1.max())] == 0).all() and m1[~air].max() > 1.0

    # ---- parity sample: 32 views of the full-shape decomposition against the oracle, fed what the kernel was fed
    from oracle import gn_oracle
    import matdecomp
    from xtomosim.system import read_parameter_file, xRaySpectrum
    os.chdir(tmp_path)
    try:
        ct = read_parameter_file('./input/params.txt')[0][3]
        specs = []
        for sid, dose in (('detunedMV', 9), ('80kV', 1)):                     # main.py:64-69,101
            sp = xRaySpectrum(f'./input/spectrum/{sid}_1mGy_float32.bin', sid)
            sp.rescale_counts(ct.A_iso * dose / ct.N_proj)
            specs.append(sp)
    finally:
        os.chdir(old_cwd)
    ee, i0, mus = gn_oracle.gn_model_tables(ct, specs[0], specs[1], matdecomp.xc.mixatten)
    assert len(ee) > 200                                                       # the union grid of the MV and kV spectra
    views = np.linspace(0, nv, 32, endpoint=False).astype(int)
    # The .bin files hold float32 roundings of the float64 sinograms main.py passed on, so decompose THESE values on both
    # sides: kernel (float32 and float64 mode) against the oracle.
    g = np.stack([s1[views], s2[views]])
    r32 = lambda a: np.asarray(a).astype(np.float32).astype(np.float64)       # noqa: E731
    i0_t = np.tile(r32(i0), (1, nc)).reshape(2, nc, len(ee))
    ref = gn_oracle.optimize_sino_cpu(g, ee, i0_t, r32(mus), 50)
    from dexct.ops import F32, F64, gn_decompose
    a64 = gn_decompose(g, r32(i0), r32(mus), 50, dtype=F64).cpu().numpy()
    a32 = gn_decompose(g, i0, mus, 50, dtype=F32).cpu().numpy()
    # MV + kV is an ill-conditioned pair and 50 Newton steps amplify rounding: the reference restatement itself, summed over
    # the energies in the opposite order (mathematically the same function), moves by `spread`.  Pixels where it does not move
    # are held to the stated tolerances; everywhere the kernel must stay within the restatement's own spread.
    ref_rev = gn_oracle.optimize_sino_cpu(g, ee[::-1].copy(), i0_t[..., ::-1].copy(), r32(mus)[:, ::-1].copy(), 50)
    ok = np.isfinite(ref).all(-1) & np.isfinite(ref_rev).all(-1)
    scale = float(np.max(np.abs(ref[ok])))
    spread = np.max(np.abs(ref - ref_rev), axis=-1) / scale
    e64 = np.max(np.abs(a64 - ref), axis=-1) / scale
    e32 = np.max(np.abs(a32 - ref), axis=-1) / scale
    stable = ok & (spread < 1e-12)
    from dexct.ops import gn_fixup_count
    frac = gn_fixup_count() / float(g.shape[1] * g.shape[2])
    msg = (f'[C1] reference main.py unchanged: wall {wall:.2f} s ({nv} x {nc}, {len(ee)} energies, 50 iterations, 4 x FBP {nm}^2); '
           f'oracle spread under reversed summation max {spread[ok].max():.3e}, stable pixels {stable.mean():.4f}; '
           f'f64 vs oracle: stable {e64[stable].max():.3e}, all {e64[ok].max():.3e}; '
           f'f32 vs oracle: stable {e32[stable].max():.3e}, all {e32[ok].max():.3e}; float64 fix-up fraction {frac:.4f}')
    with capsys.disabled():
        print('\n' + msg)
    assert stable.mean() > 0.5, msg
    # measured: stable pixels f64 1.3e-10 / f32 2.0e-7; all pixels f64 2.1e-6 / f32 2.2e-6 with the oracle's own spread 2.2e-6.
    # (One reversed summation is a single sample of a pixel's sensitivity, so "stable" pixels get 1e-9 rather than 1e-10.)
    assert e64[stable].max() < 1e-9 and e64[ok].max() < 10 * spread[ok].max(), msg
    assert e32[stable].max() < 1e-5 and e32[ok].max() < max(1e-5, 10 * spread[ok].max()), msg
    assert np.isfinite(a32[ok]).all(), msg
