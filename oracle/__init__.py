"""ORACLE package -- test infrastructure only.

CPU restatements the CUDA path is checked against.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import this package; the product
(``dex-ct-sim_b200/``) never does and fails loudly without its CUDA library instead of falling back to this.

* :mod:`oracle.gn_oracle`  -- numpy float64 restatement of ``/root/reference/matdecomp.py`` (parity PINNED against
  the reference function itself and the golden vectors it produced).
* ``siddon_oracle.c``      -- float64 C restatement of Siddon 1985 + polyenergetic Beer-Lambert detection
  (parity UNPINNED: the reference's tracer source is an absent submodule, see that file's header), loaded here
  through ctypes.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from . import gn_oracle  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, '_build', 'liboracle.so')
_lib = None


def build(force=False):
    """Compile siddon_oracle.c with the recipe in oracle/Makefile (building the checker is not using it)."""
    src = os.path.join(_HERE, 'siddon_oracle.c')
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(['make', '-C', _HERE, '-s'] + (['-B'] if force else []))
    return _LIB_PATH


class _Setup(C.Structure):
    _fields_ = [('n_views', C.c_int32), ('n_rows', C.c_int32), ('n_ch', C.c_int32),
                ('sid', C.c_double), ('sdd', C.c_double),
                ('cos_beta', C.c_void_p), ('sin_beta', C.c_void_p),
                ('cos_gamma', C.c_void_p), ('sin_gamma', C.c_void_p), ('zdet', C.c_void_p),
                ('nx', C.c_int32), ('ny', C.c_int32), ('nz', C.c_int32),
                ('dx', C.c_double), ('dy', C.c_double), ('dz', C.c_double)]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        P = C.c_void_p
        L.dexo_ray_endpoints.argtypes = [C.POINTER(_Setup), C.c_int, C.c_int, C.c_int, P, P]
        L.dexo_classic_lengths_ray.argtypes = [P, P, P, C.c_int, C.c_int, C.c_int,
                                               C.c_double, C.c_double, C.c_double, C.c_int, P]
        L.dexo_classic_lengths_ray.restype = C.c_int
        L.dexo_classic_trace.argtypes = [C.POINTER(_Setup), P, C.c_int, C.c_int, C.c_int, P, P, P, C.c_int]
        L.dexo_classic_trace.restype = C.c_int
        L.dexo_classic_lengths.argtypes = [C.POINTER(_Setup), P, C.c_int, C.c_int, C.c_int, P, P]
        L.dexo_poly_detect.argtypes = [P, C.c_int64, C.c_int, P, C.c_int, P, C.c_int, P, P]
        L.dexo_project_poly.argtypes = [C.POINTER(_Setup), P, C.c_int, P, C.c_int, P, C.c_int,
                                        C.c_int, C.c_int, P, P, P]
        L.dexo_slab_trace.argtypes = [C.POINTER(_Setup), P, C.c_int, C.c_int, C.c_int, C.c_int, P, P, P, C.c_int]
        L.dexo_slab_trace.restype = C.c_int
        L.dexo_slab_lengths.argtypes = [C.POINTER(_Setup), P, C.c_int, C.c_int, C.c_int, C.c_int, P]
        L.dexo_num_threads.restype = C.c_int
        L.dexo_set_num_threads.argtypes = [C.c_int]
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def fan_tables(N_channels, N_proj, fan_angle_total, rotation_angle_total=2.0 * np.pi, beta0=0.0,
               N_rows=1, h_iso=1.0, SID=60.0, SDD=100.0):
    """Project-defined scan conventions, restated independently of the product:
    beta_v = beta0 + v * rotation/N_proj;  gamma_c = (c - (N_ch-1)/2) * fan/N_ch;
    zdet_r = (r - (N_rows-1)/2) * h_iso * SDD/SID."""
    beta = beta0 + np.arange(N_proj, dtype=np.float64) * (rotation_angle_total / N_proj)
    gamma = (np.arange(N_channels, dtype=np.float64) - 0.5 * (N_channels - 1)) * (fan_angle_total / N_channels)
    zdet = (np.arange(N_rows, dtype=np.float64) - 0.5 * (N_rows - 1)) * (h_iso * SDD / SID)
    return dict(cos_beta=np.cos(beta), sin_beta=np.sin(beta), cos_gamma=np.cos(gamma), sin_gamma=np.sin(gamma),
                zdet=zdet)


class Scan:
    """Scanner tables + labelled volume, bound to the C oracle."""

    def __init__(self, labels, voxel, SID, SDD, tables):
        labels = np.ascontiguousarray(labels, dtype=np.uint8)
        if labels.ndim == 2:
            labels = labels[None]
        self.labels = labels                      # [nz, ny, nx]
        self.voxel = tuple(float(x) for x in voxel)     # (dx, dy, dz)
        self.t = {k: np.ascontiguousarray(v, dtype=np.float64) for k, v in tables.items()}
        self.n_views = len(self.t['cos_beta'])
        self.n_ch = len(self.t['cos_gamma'])
        self.n_rows = len(self.t['zdet'])
        nz, ny, nx = labels.shape
        self.s = _Setup(self.n_views, self.n_rows, self.n_ch, float(SID), float(SDD),
                        _ptr(self.t['cos_beta']).value, _ptr(self.t['sin_beta']).value,
                        _ptr(self.t['cos_gamma']).value, _ptr(self.t['sin_gamma']).value,
                        _ptr(self.t['zdet']).value, nx, ny, nz, *self.voxel)

    # --- (A) Siddon 1985 -------------------------------------------------------------------------------
    def endpoints(self, v, r, c):
        S = np.zeros(3)
        T = np.zeros(3)
        lib().dexo_ray_endpoints(C.byref(self.s), v, r, c, _ptr(S), _ptr(T))
        return S, T

    def classic_trace(self, v, r, c, cap=8192):
        vox = np.zeros(cap, np.int64)
        lab = np.zeros(cap, np.uint8)
        ln = np.zeros(cap, np.float64)
        n = lib().dexo_classic_trace(C.byref(self.s), _ptr(self.labels), v, r, c, _ptr(vox), _ptr(lab), _ptr(ln), cap)
        assert n <= cap
        return vox[:n], lab[:n], ln[:n]

    def classic_lengths(self, K, v0=0, nv=None):
        """L [nv, n_rows, n_ch, K] float64 and the total voxel-segment count."""
        nv = self.n_views - v0 if nv is None else nv
        L = np.zeros((nv, self.n_rows, self.n_ch, K), np.float64)
        seg = C.c_int64(0)
        lib().dexo_classic_lengths(C.byref(self.s), _ptr(self.labels), K, v0, nv, _ptr(L), C.byref(seg))
        return L, seg.value

    def project_poly(self, mu, W, v0=0, nv=None, want_log=True):
        """counts, logs [nSpec, nv, n_rows, n_ch] float64 and the segment total (whole forward projection)."""
        nv = self.n_views - v0 if nv is None else nv
        mu = np.ascontiguousarray(mu, np.float64)
        W = np.ascontiguousarray(W, np.float64)
        K, nE = mu.shape
        nS = W.shape[0]
        counts = np.zeros((nS, nv, self.n_rows, self.n_ch), np.float64)
        logs = np.zeros_like(counts) if want_log else None
        seg = C.c_int64(0)
        lib().dexo_project_poly(C.byref(self.s), _ptr(self.labels), K, _ptr(mu), nE, _ptr(W), nS, v0, nv,
                                _ptr(counts), _ptr(logs) if want_log else None, C.byref(seg))
        return counts, logs, seg.value

    # --- (B) twin of the device traversal ---------------------------------------------------------------
    def slab_trace(self, v, r, c, f64, cap=8192):
        vox = np.zeros(cap, np.int64)
        lab = np.zeros(cap, np.uint8)
        ln = np.zeros(cap, np.float64)
        n = lib().dexo_slab_trace(C.byref(self.s), _ptr(self.labels), v, r, c, int(bool(f64)),
                                  _ptr(vox), _ptr(lab), _ptr(ln), cap)
        assert n <= cap
        return vox[:n], lab[:n], ln[:n]

    def slab_lengths(self, K, f64=True, v0=0, nv=None):
        nv = self.n_views - v0 if nv is None else nv
        L = np.zeros((nv, self.n_rows, self.n_ch, K), np.float64)
        lib().dexo_slab_lengths(C.byref(self.s), _ptr(self.labels), K, v0, nv, int(bool(f64)), _ptr(L))
        return L


def classic_lengths_ray(S, T, labels, voxel, K):
    """Per-material lengths of an explicit ray S->T (analytic known-answer tests)."""
    labels = np.ascontiguousarray(labels, dtype=np.uint8)
    if labels.ndim == 2:
        labels = labels[None]
    nz, ny, nx = labels.shape
    S = np.ascontiguousarray(S, np.float64)
    T = np.ascontiguousarray(T, np.float64)
    L = np.zeros(K, np.float64)
    n = lib().dexo_classic_lengths_ray(_ptr(S), _ptr(T), _ptr(labels), nx, ny, nz, *[float(x) for x in voxel], K, _ptr(L))
    return L, n


def poly_detect(L, mu, W, want_log=True):
    """counts, logs [nSpec, n_rays] float64 from L [n_rays, K]."""
    L = np.ascontiguousarray(L, np.float64)
    mu = np.ascontiguousarray(mu, np.float64)
    W = np.ascontiguousarray(W, np.float64)
    K, nE = mu.shape
    L2 = L.reshape(-1, K)
    nS = W.shape[0]
    counts = np.zeros((nS, L2.shape[0]), np.float64)
    logs = np.zeros_like(counts) if want_log else None
    lib().dexo_poly_detect(_ptr(L2), L2.shape[0], K, _ptr(mu), nE, _ptr(W), nS, _ptr(counts),
                           _ptr(logs) if want_log else None)
    return counts, logs


def num_threads():
    return int(lib().dexo_num_threads())


def set_num_threads(n):
    """OpenMP threads of the C oracle from here on (a launcher's OMP_NUM_THREADS=1 would otherwise throttle the timed CPU
    baseline of bench.py)."""
    lib().dexo_set_num_threads(int(n))
