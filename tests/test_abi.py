"""The C-ABI library loads on a CPU-only machine and exports every symbol include/dexct.h declares; without a B200
every compute entry point fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from dexct import _native as N

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, 'include', 'dexct.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(dexct_[a-z0-9_]+)\s*\(', text)))


def test_library_exports_every_declared_symbol():
    names = _declared_symbols()
    assert len(names) >= 11
    L = C.CDLL(N.LIB_PATH)
    for n in names:
        assert hasattr(L, n), f'{n} declared in include/dexct.h but not exported by libdexct.so'
    # and the ctypes binding covers exactly the declared set
    assert sorted(N.SIGNATURES) == names


def test_struct_layouts_match_header():
    assert C.sizeof(N.DexctScan) == 4 * 4 + 8 * 8
    assert C.sizeof(N.DexctVolume) == 4 * 4 + 3 * 8


def test_version_and_workspace_size_need_no_gpu():
    L = N.lib()
    assert L.dexct_version() == 100
    from dexct.ops import make_scan
    scan = make_scan(1000, 900, 0.8230337, 60.0, 100.0)
    vol = N.DexctVolume(512, 512, 1, 0, 0.1, 0.1, 0.1)
    nbytes = L.dexct_siddon_workspace_bytes(C.byref(scan), C.byref(vol))
    # two padded uint16 pair layouts of 3 planes each + four one-plane skip layouts (2-D scan) + small tables
    assert (2 * 3 + 4) * 512 * 514 * 2 <= nbytes < (2 * 3 + 4) * 512 * 530 * 2 + 64 * 1024
    vol3 = N.DexctVolume(512, 512, 4, 0, 0.1, 0.1, 0.1)
    n3 = L.dexct_siddon_workspace_bytes(C.byref(scan), C.byref(vol3))
    # 3-D: two pair layouts and four skip layouts of nz + 2 planes each, plus two z-homogeneity radius volumes (1 B / voxel)
    rz = 2 * 512 * 512 * 4
    assert 6 * 6 * 512 * 514 * 2 + rz <= n3 < 6 * 6 * 512 * 530 * 2 + rz + 64 * 1024
    bad = N.DexctVolume(0, 512, 1, 0, 0.1, 0.1, 0.1)
    assert L.dexct_siddon_workspace_bytes(C.byref(scan), C.byref(bad)) == 0


def test_no_device_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip('this check is for the CPU-only machine')
    L = N.lib()
    assert L.dexct_device_count() == 0
    rc = L.dexct_gn_decompose(None, None, None, 1, 1, 1, 0.0, None, None, 0, 0, None)
    assert rc == -3 and b'no CPU fallback' in L.dexct_last_error()
    with pytest.raises(N.DexctError):
        N.require_device()
    # the drop-in entry points refuse too, instead of quietly computing on the host
    from dexct import synthetic as S
    from xtomosim.forward_project import get_sino
    ct, ph, lo, hi = S.make_config('C2', n_views=4)
    with pytest.raises(N.DexctError):
        get_sino(ct, ph, lo)
    import matdecomp
    import numpy as np
    with pytest.raises(N.DexctError):
        matdecomp.get_basismat_sinos(ct, np.ones((4, 900)), np.ones((4, 900)), lo, hi, n_iters=2)


def test_product_never_imports_the_oracle():
    """The shipped path must not route through oracle/ (or any CPU implementation)."""
    pkg = os.path.join(ROOT, 'dex-ct-sim_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', src, flags=re.M), f
                assert 'liboracle' not in src, f
