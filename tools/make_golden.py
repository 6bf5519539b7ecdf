#!/usr/bin/env python3
"""Generate tests/golden/*.npz by RUNNING THE REFERENCE ITSELF (only possible where /root/reference is mounted).

The reference's Gauss-Newton code (/root/reference/matdecomp.py) is imported unmodified through
``oracle.gn_oracle.load_reference_matdecomp`` (cupy / xcompy are stubbed for the import only) and its outputs are
stored next to the inputs that produced them.  The GPU box has no /root/reference, so the `-m gpu` tests compare the
CUDA path with these files.  Also copies the reference's small binary INPUT DATA (spectra, detector efficiencies --
data, not source code) to tests/golden/input/ because config C1 is defined on them.

    python tools/make_golden.py
"""
import os
import shutil
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'dex-ct-sim_b200'))
GOLD = os.path.join(ROOT, 'tests', 'golden')
REF = '/root/reference'

from oracle import gn_oracle  # noqa: E402
from xtomosim import xcompy as standin_xc  # noqa: E402
from xtomosim.system import FanBeamGeometry, xRaySpectrum  # noqa: E402


def main():
    ref = gn_oracle.load_reference_matdecomp(REF)
    if ref is None:
        sys.exit('reference tree not mounted; golden vectors can only be regenerated in the build container')
    os.makedirs(os.path.join(GOLD, 'input', 'spectrum'), exist_ok=True)
    os.makedirs(os.path.join(GOLD, 'input', 'detector'), exist_ok=True)
    for sub in ('spectrum', 'detector'):
        for f in sorted(os.listdir(os.path.join(REF, 'input', sub))):
            shutil.copyfile(os.path.join(REF, 'input', sub, f), os.path.join(GOLD, 'input', sub, f))

    rng = np.random.default_rng(20240)

    # --- G1: optimize_sino_cpu on a synthetic 120-bin problem (C2/C3 shape, shrunk) -----------------------------
    nV, nB, nE = 3, 40, 120
    E = 20.0 + np.arange(nE)
    i0 = np.stack([np.exp(-0.5 * ((E - 55) / 18) ** 2), np.exp(-0.5 * ((E - 95) / 25) ** 2)])
    i0 *= 1e5 / i0.sum(axis=1, keepdims=True)
    mus = np.stack([standin_xc.mixatten(gn_oracle.matcomp1, E), standin_xc.mixatten(gn_oracle.matcomp2, E)])
    a_true = np.stack([rng.uniform(0, 30, (nV, nB)), rng.uniform(0, 5, (nV, nB))], axis=-1)
    g = gn_oracle.forward_counts(a_true, i0, mus)
    i0_t = np.tile(i0, (1, nB)).reshape(2, nB, nE)
    for n_iters in (10, 25):
        out = ref.optimize_sino_cpu(g, E, i0_t, mus, n_iters, verbose=False)
        np.savez_compressed(os.path.join(GOLD, f'gn_synth_{n_iters}it.npz'), g=g, ee=E, i0=i0, mus=mus,
                            n_iters=n_iters, a_true=a_true, a_ref=out)

    # --- G2: the shipped spectra: detunedMV + 80 kV on the EID detector (config C1 model, 239 energies) ----------
    ct = FanBeamGeometry(800, 1200, 0.8230337, 60.0, 100.0, 1.0, True, os.path.join(REF, 'input/detector/eta_eid_mv.bin'))
    specs = []
    for sid, dose in (('detunedMV', 9), ('80kV', 1)):
        s = xRaySpectrum(os.path.join(REF, f'input/spectrum/{sid}_1mGy_float32.bin'), sid)
        s.rescale_counts(ct.A_iso * dose / ct.N_proj)            # main.py:68
        specs.append(s)
    # run the reference's own table assembly + solver with the stand-in mixatten injected as its `xc`
    ref.xc = types.SimpleNamespace(mixatten=standin_xc.mixatten)
    nV, nB = 2, 48
    a_true = np.stack([rng.uniform(0, 35, (nV, nB)), rng.uniform(0, 6, (nV, nB))], axis=-1)
    ee, i0u, musu = gn_oracle.gn_model_tables(ct, specs[0], specs[1], standin_xc.mixatten)
    gg = gn_oracle.forward_counts(a_true, i0u, musu)
    gg[:, 0, :4] = i0u.sum(axis=1)[:, None]                       # a few air pixels -> exercised by the mask
    ct_small = types.SimpleNamespace(det_E=ct.det_E, det_eta_E=ct.det_eta_E, eid=ct.eid)
    out = ref.do_matdecomp_gn(ct_small, gg[0], gg[1], specs[0], specs[1], 30)
    m1, m2 = ref.get_basismat_sinos(ct_small, gg[0].copy(), gg[1].copy(), specs[0], specs[1], n_iters=30, mask_thresh=0.95)
    np.savez_compressed(os.path.join(GOLD, 'gn_c1_tables.npz'), g=gg, ee=ee, i0=i0u, mus=musu, n_iters=30,
                        a_true=a_true, a_ref=out, mat1=m1, mat2=m2, mask_thresh=0.95,
                        spec1_E=specs[0].E, spec1_I0=specs[0].I0, spec2_E=specs[1].E, spec2_I0=specs[1].I0,
                        det_E=ct.det_E, det_eta_E=ct.det_eta_E, eid=ct.eid)
    print('golden vectors written to', GOLD)


if __name__ == '__main__':
    main()
