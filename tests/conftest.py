"""Test configuration.

`-m "not gpu"`: oracle vs golden vectors / the reference itself, host logic, C-ABI symbol checks (CPU only).
`-m gpu`      : parity tests proper -- the CUDA path, called through the C ABI, against the oracle.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, 'dex-ct-sim_b200')
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, 'tests', 'golden')
REFERENCE = '/root/reference'


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a B200 (run with -m gpu on the GPU box)')


def pytest_collection_modifyitems(config, items):
    """GPU tests fail loudly (not skip) when selected on a machine without a device: a silent skip would hide a
    missing CUDA path."""
    return


@pytest.fixture(scope='session')
def golden_dir():
    return GOLDEN


@pytest.fixture(scope='session')
def have_reference():
    return os.path.exists(os.path.join(REFERENCE, 'matdecomp.py'))


@pytest.fixture(scope='session')
def rng():
    return np.random.default_rng(20240)


def scale_rel_err(a, ref, axis=None):
    """max |a - ref| / max |ref| (scale-relative error; see DESIGN.md section 8 for why not element-wise)."""
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    den = np.max(np.abs(ref), axis=axis, keepdims=axis is not None)
    return np.max(np.abs(a - ref) / np.maximum(den, 1e-300))


COUNT_FLOOR = 1e-6          # element-wise checks of count sinograms cover every pixel above this fraction of the air count


def elem_rel_err(a, ref, floor=COUNT_FLOOR):
    """max over elements with |ref| >= floor * max|ref| of |a - ref| / |ref| -- the element-wise relative error the
    north_star's "within 1e-5 relative" names, on every pixel down to six decades below the brightest one (count sinograms
    span ~4 decades: a scale-relative bound alone would let percent-level errors through on strongly attenuated rays)."""
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    big = np.abs(ref) >= floor * np.max(np.abs(ref))
    if not big.any():
        return 0.0
    return float(np.max(np.abs(a[big] - ref[big]) / np.abs(ref[big])))


def per_ray_len_err(L, L_ref):
    """max over rays of max_m |L_m - Lref_m| / chord(ray), chord = sum_m Lref_m (rays that meet the volume): the error of
    every per-material length against THAT ray's own path length, not against the longest chord of the scan."""
    L = np.asarray(L, dtype=np.float64)
    L_ref = np.asarray(L_ref, dtype=np.float64)
    chord = L_ref.sum(-1)
    hit = chord > 0
    if not hit.any():
        return 0.0
    return float(np.max(np.max(np.abs(L - L_ref), axis=-1)[hit] / chord[hit]))
