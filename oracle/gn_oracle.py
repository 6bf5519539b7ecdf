"""ORACLE (test infrastructure, never the product path): numpy float64 restatement of the
reference's Gauss-Newton two-basis decomposition.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference`` legs may
import this module.  The shipped path (``dex-ct-sim_b200/matdecomp.py`` -> ``libdexct.so``) never does.

Parity status: PINNED.  ``tests/test_oracle_gn.py`` checks this restatement (a) against the reference function
itself, loaded unmodified from ``/root/reference/matdecomp.py`` when that tree is present (this container), and
(b) against ``tests/golden/gn_*.npz`` which were produced by the reference function (``tools/make_golden.py``).

What it follows, line by line (reference = ``/root/reference/matdecomp.py``):

* ``optimize_sino_cpu``  matdecomp.py:87-127  -> :func:`optimize_sino_cpu`
* ``do_matdecomp_gn``    matdecomp.py:130-164 -> :func:`gn_model_tables` + :func:`do_matdecomp_gn`
* ``get_basismat_sinos`` matdecomp.py:167-207 -> :func:`get_basismat_sinos`
"""
from __future__ import annotations

import numpy as np

# matdecomp.py:11-17 -- hard-coded basis materials of the reference.
mat1 = 'ICRU tissue'
matcomp1 = 'H(10.2)C(14.3)N(3.4)O(70.8)Na(0.2)P(0.3)S(0.3)Cl(0.2)K(0.3)'
density1 = 1.06
mat2 = 'ICRU bone'
matcomp2 = 'H(3.4)C(15.5)N(4.2)O(43.5)Na(0.1)Mg(0.2)P(10.3)S(0.3)Ca(22.5)'
density2 = 1.92

EPS_INIT = 1e-6          # matdecomp.py:98  (initial value of every density line integral)
EXP_CLIP = 700.0         # matdecomp.py:116 (.clip(-700, 700) ahead of exp)


def optimize_sino_cpu(Sino_gg, ee, i0, mus, n_iters, verbose=False):
    """Full-Newton iterations on the Poisson negative log-likelihood, one independent 2x2 problem per
    (view, channel) pixel.  Follows matdecomp.py:87-127 statement by statement (same broadcasting shapes, same
    reduction axes, hence the same numpy pairwise-summation order over the energy axis); the residual
    ``g/nu - 1`` is named ``r`` instead of being spelled out three times.

    Sino_gg [nMeas, nViews, nBins] counts; ee [nE] (unused by the loop, as in the reference);
    i0 [nMeas, nBins, nE]; mus [nMats, nE]; returns [nViews, nBins, nMats] float64.
    """
    Sino_gg = np.asarray(Sino_gg, dtype=np.float64)
    i0 = np.asarray(i0, dtype=np.float64)
    mus = np.asarray(mus, dtype=np.float64)
    n_meas, nViews, nBins = Sino_gg.shape
    nMats = mus.shape[0]
    out = np.full([nViews, nBins, nMats], EPS_INIT)                         # :98-99

    ssff = i0[:, np.newaxis, :, :] * mus[np.newaxis, :, np.newaxis, :]        # :102 [nMeas,nMats,nBins,nE]
    ssff2 = i0[:, np.newaxis, np.newaxis, :, :] * \
        (mus[np.newaxis, :, :] * mus[:, np.newaxis, :])[np.newaxis, :, :, np.newaxis, :]   # :105

    for j in range(nViews):                                                  # :109
        a = out[j]                                                           # [nBins, nMats]
        g = Sino_gg[:, j, :]                                                 # [nMeas, nBins]
        for _ in range(n_iters):                                             # :114
            atten = np.exp(np.sum(-1.0 * a[:, :, np.newaxis] * mus, axis=1).clip(-EXP_CLIP, EXP_CLIP))  # :116
            nu = np.sum(i0 * atten, axis=2)                                  # :118 [nMeas,nBins]
            nu_grad = -1.0 * np.sum(ssff * atten, axis=3)                    # :119 [nMeas,nMats,nBins]
            nu_hess = np.sum(ssff2 * atten, axis=4)                          # :120 [nMeas,nMats,nMats,nBins]
            r = g / nu - 1.0
            dF = -1.0 * np.sum(r[:, np.newaxis, :] * nu_grad, axis=0)        # :122 [nMats,nBins]
            H = -1.0 * np.sum(r[:, np.newaxis, np.newaxis, :] * nu_hess
                              - (g / (nu * nu))[:, np.newaxis, np.newaxis, :]
                              * (nu_grad[:, np.newaxis, :, :] * nu_grad[:, :, np.newaxis, :]), axis=0)  # :123
            a -= (np.linalg.inv(H.transpose([2, 0, 1])) * (dF.T)[:, np.newaxis, :]).sum(axis=2)        # :125
    return out


def gn_model_tables(ct, spec1, spec2, mixatten):
    """Spectral model of the decomposition: union energy grid, bin widths, detector response, per-measurement
    weights and the two basis mass-attenuation rows.  Follows matdecomp.py:138-160.

    Returns (ee [nE], i0_2xE [2,nE], mus [2,nE]).  The reference tiles ``i0`` to [2, N_channels, nE] at :151;
    every channel row is identical, so the tiling is left to the caller.
    """
    ee = np.array(sorted(list(set(np.append(spec1.E, spec2.E)))))            # :140
    dE = np.append([ee[0]], ee[1:] - ee[:-1])                                # :142 first bin is 0..E[0]
    detresponse = np.array(np.interp(ee, ct.det_E, ct.det_eta_E))            # :146
    if ct.eid:
        detresponse = detresponse * ee                                       # :147-148
    i01 = np.interp(ee, spec1.E, spec1.I0) * detresponse * dE                # :149
    i02 = np.interp(ee, spec2.E, spec2.I0) * detresponse * dE                # :150
    mus = np.array([mixatten(matcomp1, ee), mixatten(matcomp2, ee)])         # :155-160 (density multiply is
    return ee, np.array([i01, i02]), mus                                     #  commented out at :157)


def do_matdecomp_gn(ct, sino1, sino2, spec1, spec2, n_iters, mixatten):
    """matdecomp.py:130-164."""
    N_proj, N_channels = sino1.shape
    Sino_gg = np.array([sino1, sino2])                                       # :136
    ee, i0_2xE, mus = gn_model_tables(ct, spec1, spec2, mixatten)
    i0 = np.tile(i0_2xE, (1, N_channels)).reshape([2, N_channels, len(ee)])  # :151
    return optimize_sino_cpu(Sino_gg, ee, i0, mus, n_iters)                  # :163


def get_basismat_sinos(ct, sino_raw_1, sino_raw_2, spec1, spec2, mixatten, n_iters=30, mask_thresh=0.95):
    """matdecomp.py:167-207: air mask from measurement 1, decomposition, zero the masked pixels."""
    sino_mask = np.zeros(sino_raw_1.shape, dtype=bool)                       # :195
    sino_mask[sino_raw_1 >= mask_thresh * np.max(sino_raw_1)] = 1            # :196
    dli = do_matdecomp_gn(ct, sino_raw_1, sino_raw_2, spec1, spec2, n_iters, mixatten)   # :199
    sino_mat1 = dli[:, :, 0]                                                 # :200
    sino_mat2 = dli[:, :, 1]
    sino_mat1[sino_mask] = 0                                                 # :204-205
    sino_mat2[sino_mask] = 0
    return sino_mat1, sino_mat2


def forward_counts(a, i0_2xE, mus):
    """The decomposition's own forward model (matdecomp.py:116-118): nu[m] = sum_e i0[m,e] exp(-sum_k a_k mu_k(e)).
    a [...,2] -> counts [2, ...].  Used by tests to synthesise consistent, noiseless measurements."""
    a = np.asarray(a, dtype=np.float64)
    t = -(a[..., 0, None] * mus[0] + a[..., 1, None] * mus[1])
    att = np.exp(np.clip(t, -EXP_CLIP, EXP_CLIP))
    return np.stack([np.sum(i0_2xE[0] * att, axis=-1), np.sum(i0_2xE[1] * att, axis=-1)])


def load_reference_matdecomp(ref_root='/root/reference'):
    """Import the reference's own ``matdecomp`` module UNMODIFIED from ``ref_root`` (only possible where that
    tree exists, i.e. in the build container).  ``cupy`` and ``xcompy`` are absent there; both are only touched
    at import time (matdecomp.py:6-7,20), so inert stand-ins are placed in ``sys.modules`` for the import."""
    import importlib.util
    import os
    import sys
    import types
    path = os.path.join(ref_root, 'matdecomp.py')
    if not os.path.exists(path):
        return None
    saved = {k: sys.modules.get(k) for k in ('cupy', 'xcompy')}
    cp = types.ModuleType('cupy')
    cp.float32 = np.float32                        # default argument at matdecomp.py:20
    xc = types.ModuleType('xcompy')
    sys.modules['cupy'], sys.modules['xcompy'] = cp, xc
    try:
        spec = importlib.util.spec_from_file_location('_reference_matdecomp', path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mod
