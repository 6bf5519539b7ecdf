"""ctypes binding of ``libdexct.so`` (C ABI: ``include/dexct.h``).

This is the binding a maintainer of the reference would add next to ``matdecomp.py``.  PyTorch
is used for exactly two things: device buffers (``torch.Tensor.data_ptr()``) and the current CUDA stream.  There is
NO CPU fallback: if the library cannot be loaded, or no B200 is visible, every call raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), 'csrc', 'libdexct.so')

F32, F64 = 0, 1


class DexctError(RuntimeError):
    pass


class DexctScan(C.Structure):
    _fields_ = [('n_views', C.c_int32), ('n_rows', C.c_int32), ('n_ch', C.c_int32), ('reserved', C.c_int32),
                ('sid', C.c_double), ('sdd', C.c_double), ('beta0', C.c_double), ('dbeta', C.c_double),
                ('gamma0', C.c_double), ('dgamma', C.c_double), ('z0', C.c_double), ('dzdet', C.c_double)]


class DexctVolume(C.Structure):
    _fields_ = [('nx', C.c_int32), ('ny', C.c_int32), ('nz', C.c_int32), ('reserved', C.c_int32),
                ('dx', C.c_double), ('dy', C.c_double), ('dz', C.c_double)]


# name -> (restype, argtypes); every symbol declared in include/dexct.h
_P = C.c_void_p
SIGNATURES = {
    'dexct_version': (C.c_int, []),
    'dexct_last_error': (C.c_char_p, []),
    'dexct_device_count': (C.c_int, []),
    'dexct_siddon_workspace_bytes': (C.c_size_t, [C.POINTER(DexctScan), C.POINTER(DexctVolume)]),
    'dexct_siddon_prepare': (C.c_int, [C.POINTER(DexctScan), C.POINTER(DexctVolume), _P, _P, _P]),
    'dexct_siddon_tables_host': (C.c_int, [C.POINTER(DexctScan), _P, _P, _P, _P, _P, _P, _P]),
    'dexct_siddon_poly': (C.c_int, [C.POINTER(DexctScan), C.POINTER(DexctVolume), _P, C.c_int32, C.c_int32,
                                    _P, C.c_int32, C.c_int32, _P, C.c_int32, _P, _P, _P, C.c_int32, _P]),
    'dexct_poly_detect': (C.c_int, [_P, C.c_int64, _P, C.c_int32, C.c_int32, _P, C.c_int32, _P, _P, C.c_int32, _P]),
    'dexct_siddon_trace': (C.c_int, [C.POINTER(DexctScan), C.POINTER(DexctVolume), _P, _P, C.c_int32, C.c_int32,
                                     _P, _P, _P, _P, C.c_int32, _P]),
    'dexct_siddon_walk_trace': (C.c_int, [C.POINTER(DexctScan), C.POINTER(DexctVolume), _P, _P, C.c_int32, C.c_int32, C.c_int32,
                                          _P, _P, _P, _P, _P, _P, _P, _P, C.c_int32, _P]),
    'dexct_gn_decompose': (C.c_int, [_P, _P, _P, C.c_int32, C.c_int64, C.c_int32, C.c_double, _P, _P, C.c_int32,
                                     C.c_int32, _P]),
    'dexct_gn_decompose_general': (C.c_int, [_P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.c_int64,
                                             C.c_int32, _P, C.c_int32, _P]),
    'dexct_gn_scratch_bytes': (C.c_size_t, [C.c_int64, C.c_int32]),
    'dexct_gn_static_max_pixels': (C.c_int64, []),
    'dexct_gn_decompose_static': (C.c_int, [_P, _P, _P, C.c_int32, C.c_int64, C.c_int32, C.c_double, _P, _P]),
    'dexct_gn_air_max': (C.c_int, [_P, C.c_int64, _P, C.c_int32, _P]),
    'dexct_fbp_workspace_bytes': (C.c_size_t, [C.POINTER(DexctScan), C.c_int32, C.c_int32]),
    'dexct_fbp_fan': (C.c_int, [C.POINTER(DexctScan), _P, C.c_int32, C.c_int32, C.c_int32, C.c_double, C.c_double,
                                _P, C.c_int32, _P, C.c_int32, _P]),
    'dexct_fbp_hu': (C.c_int, [_P, C.c_int64, C.c_double, _P, C.c_int32, _P]),
    'dexct_poisson_noise': (C.c_int, [_P, _P, C.c_int64, C.c_uint64, C.c_uint32, C.c_uint64, C.c_double, C.c_double,
                                      C.c_int32, _P]),
    'dexct_peer_ctrl_bytes': (C.c_size_t, []),
    'dexct_peer_alloc': (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p)]),
    'dexct_peer_free': (C.c_int, [_P]),
    'dexct_peer_export': (C.c_int, [_P, _P]),
    'dexct_peer_open': (C.c_int, [_P, C.POINTER(C.c_void_p)]),
    'dexct_peer_close': (C.c_int, [_P]),
    'dexct_peer_barrier': (C.c_int, [_P, C.c_int32, C.c_int32, C.c_uint64, C.c_double, _P]),
    'dexct_peer_max_key': (C.c_int, [_P, C.c_int32, C.c_int32, C.c_uint64, _P, C.c_int32, C.c_double, _P]),
    'dexct_peer_status': (C.c_int, [_P, C.POINTER(C.c_int32), _P]),
    'dexct_peer_push': (C.c_int, [_P, C.c_size_t, _P, C.c_int32, C.c_int32, C.c_int32, _P]),
    'dexct_peer_push_dma': (C.c_int, [_P, C.c_size_t, _P, C.c_int32, C.c_int32, _P]),
    'dexct_gn_decompose_mirrored': (C.c_int, [_P, _P, _P, C.c_int32, C.c_int64, C.c_int32, C.c_double, _P, _P, C.c_int32,
                                              _P, C.c_int32, C.c_int32, _P]),
    'dexct_gn_decompose_fused_push': (C.c_int, [_P, _P, _P, C.c_int32, C.c_int64, C.c_int32, C.c_double, _P, _P, C.c_int32,
                                                _P, C.c_size_t, _P, C.c_int32, _P, C.c_int32, C.c_int32, _P]),
    'dexct_vmi': (C.c_int, [_P, _P, C.c_int64, C.c_double, C.c_double, C.c_double, C.c_int32, _P, C.c_int32, _P]),
}

_lib = None


def lib():
    """Load libdexct.so (once).  Raises DexctError when it has not been built -- never falls back to CPU code."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise DexctError(f'{LIB_PATH} is missing: build it with `python -c "import __graft_entry__ as g; g.build()"` '
                             f'or `make -C {os.path.dirname(LIB_PATH)}`.  There is no CPU fallback.')
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)          # AttributeError here means the library is stale
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise DexctError(f'libdexct error {rc}: {lib().dexct_last_error().decode()}')


def require_device():
    import torch
    if not torch.cuda.is_available() or lib().dexct_device_count() < 1:
        raise DexctError('no B200 (compute capability 10.x) visible: dexct has no CPU fallback')


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return C.c_void_p(0 if t is None else t.data_ptr())
