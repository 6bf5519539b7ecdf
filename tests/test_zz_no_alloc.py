"""dexct_gn_decompose_static: the decomposition entry point that allocates nothing (its scratch block is a file-scope device
array of the library; include/dexct.h).  Called through the C ABI and checked against the existing entry point,
dexct_gn_decompose, at the single-precision tolerance of tests/test_gpu_gn.py (1e-5 scale-relative), on two calls in a row --
the second with other data and another size, so that anything the first call left in the shared array would show.

Collected last on purpose; passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import scale_rel_err
from dexct import _native as N
from dexct import synthetic as S
from dexct.ops import F32, gn_decompose

pytestmark = pytest.mark.gpu

N_ITERS = 8


def _problem(P, seed, air_fraction):
    g, i0, mus, _ = S.gn_problem(40, 1000)
    r = np.random.default_rng(seed)
    pick = r.permutation(g.shape[1] * g.shape[2])[:P]
    gg = np.ascontiguousarray(g.reshape(2, -1)[:, pick])
    air = r.random(P) < air_fraction
    gg[0, air] = gg[0].max() * 1.1                                        # pixels the air mask zeroes
    return gg, i0, mus


def _static(gd, i0d, musd, mask):
    P = gd.shape[1]
    out = torch.full((P, 2), float('nan'), dtype=torch.float32, device=gd.device)
    rc = N.lib().dexct_gn_decompose_static(N.ptr(gd), N.ptr(i0d), N.ptr(musd), i0d.shape[1], P, N_ITERS, float(mask), N.ptr(out),
                                           N.stream_ptr())
    N.check(rc)
    return out


def test_static_scratch_entry_point_matches_dexct_gn_decompose_twice_in_a_row():
    N.require_device()
    dev = torch.device('cuda', torch.cuda.current_device())
    assert N.lib().dexct_gn_static_max_pixels() == 1048576
    calls = [(30011, 1, 0.3, 0.9), (4097, 2, 0.0, 0.0)]                   # (pixels, seed, air fraction, mask threshold)
    free_before = None
    for k, (P, seed, air, mask) in enumerate(calls):
        g, i0, mus = _problem(P, seed, air)
        gd = torch.as_tensor(g, dtype=torch.float32, device=dev)
        i0d = torch.as_tensor(i0, dtype=torch.float32, device=dev)
        musd = torch.as_tensor(mus, dtype=torch.float32, device=dev)
        want = gn_decompose(gd, i0d, musd, N_ITERS, mask_thresh=mask, dtype=F32).cpu().numpy()      # dexct_gn_decompose
        torch.cuda.synchronize()
        if k == 1:                                                        # (the first call may load the kernels' code)
            free_before = torch.cuda.mem_get_info()[0]
        got_t = _static(gd, i0d, musd, mask)
        torch.cuda.synchronize()
        if k == 1:
            assert torch.cuda.mem_get_info()[0] == free_before, 'the call changed the amount of free device memory'
        got = got_t.cpu().numpy()
        assert np.isfinite(got).all()
        err = scale_rel_err(got, want)
        assert err < 1e-5, f'call {k}: static-scratch entry point vs dexct_gn_decompose {err:.3e}'
        if mask > 0:
            masked = g[0].astype(np.float32) >= np.float32(mask) * g[0].astype(np.float32).max()
            assert masked.any() and (got[masked] == 0).all() and (got[~masked] != 0).all()
        else:
            assert (got != 0).all()


def test_static_scratch_entry_point_refuses_what_does_not_fit():
    N.require_device()
    dev = torch.device('cuda', torch.cuda.current_device())
    tiny = torch.ones((2, 8), dtype=torch.float32, device=dev)
    out = torch.empty((8, 2), dtype=torch.float32, device=dev)
    too_many = N.lib().dexct_gn_static_max_pixels() + 1
    rc = N.lib().dexct_gn_decompose_static(N.ptr(tiny), N.ptr(tiny), N.ptr(tiny), 8, too_many, 1, 0.0, N.ptr(out), N.stream_ptr())
    assert rc == -4                                                       # DEXCT_ECAPACITY
    msg = N.lib().dexct_last_error().decode()
    assert 'DEXCT_GN_STATIC_MAX_PIXELS' in msg and str(too_many) in msg
    with pytest.raises(N.DexctError):
        N.check(N.lib().dexct_gn_decompose_static(None, N.ptr(tiny), N.ptr(tiny), 8, 8, 1, 0.0, N.ptr(out), N.stream_ptr()))
    assert isinstance(C.c_int64(too_many).value, int)
