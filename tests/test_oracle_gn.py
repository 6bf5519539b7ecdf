"""Pin the GN oracle (oracle/gn_oracle.py): against the reference function itself where /root/reference is mounted, and
against the golden vectors the reference produced (tools/make_golden.py) everywhere."""
import os
import types

import numpy as np
import pytest

from oracle import gn_oracle


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + '.npz'))


@pytest.mark.parametrize('name', ['gn_synth_10it', 'gn_synth_25it', 'gn_c1_tables'])
def test_restatement_matches_golden(golden_dir, name):
    d = _load(golden_dir, name)
    g, i0, mus = d['g'], d['i0'], d['mus']
    nB = g.shape[2]
    i0_t = np.tile(i0, (1, nB)).reshape(2, nB, i0.shape[1])
    out = gn_oracle.optimize_sino_cpu(g, d['ee'], i0_t, mus, int(d['n_iters']))
    # same statements, same numpy reductions: identical to the last bit
    assert np.array_equal(out, d['a_ref'])


def test_golden_converged_recovers_truth(golden_dir):
    d = _load(golden_dir, 'gn_synth_25it')
    assert np.max(np.abs(d['a_ref'] - d['a_true'])) < 1e-10     # round-trip KAT (SURVEY 8c ii)


def test_restatement_matches_reference_function(have_reference, rng):
    if not have_reference:
        pytest.skip('reference tree not mounted on this machine')
    ref = gn_oracle.load_reference_matdecomp()
    nV, nB, nE = 2, 33, 57
    E = 30.0 + 1.5 * np.arange(nE)
    i0 = rng.uniform(0.1, 1.0, (2, nE)) * 1e4
    mus = np.stack([0.2 + 5.0 / E, 0.15 + 60.0 / E ** 1.5])
    a = np.stack([rng.uniform(0, 20, (nV, nB)), rng.uniform(0, 4, (nV, nB))], axis=-1)
    g = gn_oracle.forward_counts(a, i0, mus)
    i0_t = np.tile(i0, (1, nB)).reshape(2, nB, nE)
    for n_iters in (0, 1, 7, 20):
        want = ref.optimize_sino_cpu(g, E, i0_t, mus, n_iters, verbose=False)
        got = gn_oracle.optimize_sino_cpu(g, E, i0_t, mus, n_iters)
        assert np.array_equal(got, want), n_iters


def test_table_assembly_matches_reference(have_reference, golden_dir):
    """do_matdecomp_gn / get_basismat_sinos restatements vs the golden run of the reference's own functions."""
    d = _load(golden_dir, 'gn_c1_tables')
    from xtomosim import xcompy as xc
    ct = types.SimpleNamespace(det_E=d['det_E'], det_eta_E=d['det_eta_E'], eid=bool(d['eid']))
    s1 = types.SimpleNamespace(E=d['spec1_E'], I0=d['spec1_I0'])
    s2 = types.SimpleNamespace(E=d['spec2_E'], I0=d['spec2_I0'])
    ee, i0, mus = gn_oracle.gn_model_tables(ct, s1, s2, xc.mixatten)
    assert np.array_equal(ee, d['ee']) and np.array_equal(i0, d['i0']) and np.array_equal(mus, d['mus'])
    assert len(ee) == 239                                        # SURVEY: detunedMV U 80kV
    m1, m2 = gn_oracle.get_basismat_sinos(ct, d['g'][0].copy(), d['g'][1].copy(), s1, s2, xc.mixatten,
                                          n_iters=int(d['n_iters']), mask_thresh=float(d['mask_thresh']))
    assert np.array_equal(m1, d['mat1']) and np.array_equal(m2, d['mat2'])
    assert (m1[0, :4] == 0).all() and (m2[0, :4] == 0).all()       # the air pixels planted by make_golden.py


def test_restatement_matches_reference_function_in_its_general_form(have_reference, rng):
    """matdecomp.py:87-127 as written takes a spectrum per detector bin and any number of measurements / basis materials
    (:95,101-105,118,125); the restatement must follow it there too (the general CUDA kernel is checked against it)."""
    if not have_reference:
        pytest.skip('reference tree not mounted on this machine')
    ref = gn_oracle.load_reference_matdecomp()
    nV, nB, nE = 2, 17, 41
    E = 30.0 + 2.0 * np.arange(nE)
    base = np.stack([0.2 + 5.0 / E, 0.15 + 60.0 / E ** 1.5, 0.1 + 900.0 / E ** 2.5])
    for n_meas, n_mats in ((2, 2), (3, 3), (3, 2), (4, 3), (1, 1)):
        mus = base[:n_mats]
        shape = np.exp(-0.5 * ((E[None, :] - np.linspace(50, 95, n_meas)[:, None]) / 18.0) ** 2) + 0.02
        i0 = shape[:, None, :] * rng.uniform(0.5, 1.5, (n_meas, nB, 1)) * rng.uniform(0.9, 1.1, (n_meas, nB, nE)) * 1e4
        a = np.stack([rng.uniform(0, [12, 2.5, 0.3][q], (nV, nB)) for q in range(n_mats)], axis=-1)
        att = np.exp(-(a[..., None, :] * mus.T[None, None]).sum(-1))               # [nV, nB, nE]
        g = np.stack([(att * i0[m][None]).sum(-1) for m in range(n_meas)])
        for n_iters in (1, 6):
            want = ref.optimize_sino_cpu(g, E, i0, mus, n_iters, verbose=False)
            got = gn_oracle.optimize_sino_cpu(g, E, i0, mus, n_iters)
            assert np.array_equal(got, want), (n_meas, n_mats, n_iters)
