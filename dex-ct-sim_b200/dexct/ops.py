"""Device-level operators over libdexct: torch tensors in, torch tensors out, nothing computed in Python.

``Projector``   -- prepared workspace + Siddon/polyenergetic projection (replaces get_sino, /root/reference/main.py:120)
``poly_detect`` -- detection from stored path lengths (second spectrum of a scan, main.py:109-120)
``gn_decompose``-- Gauss-Newton decomposition (replaces optimize_sino_cpu, /root/reference/matdecomp.py:87-127)
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _native as N
from ._native import DexctScan, DexctVolume, F32, F64

_TORCH_DTYPE = {F32: torch.float32, F64: torch.float64}


def dtype_code(dtype):
    if dtype in (F32, 'f32', 'float32', torch.float32, np.float32):
        return F32
    if dtype in (F64, 'f64', 'float64', torch.float64, np.float64):
        return F64
    raise ValueError(f'dtype must be f32 or f64, got {dtype!r}')


def make_scan(n_views, n_ch, fan_angle_total, SID, SDD, n_rows=1, h_iso=1.0, rotation=2.0 * np.pi, beta0=0.0):
    """Project-defined scan conventions: view angles start at beta0 and advance by rotation/n_views;
    channel centres are symmetric about the central ray; row centres symmetric about z = 0, pitch h_iso at the isocentre."""
    dgamma = float(fan_angle_total) / n_ch
    dzdet = float(h_iso) * SDD / SID
    return DexctScan(n_views=n_views, n_rows=n_rows, n_ch=n_ch, reserved=0, sid=SID, sdd=SDD,
                     beta0=beta0, dbeta=float(rotation) / n_views,
                     gamma0=-0.5 * (n_ch - 1) * dgamma, dgamma=dgamma,
                     z0=-0.5 * (n_rows - 1) * dzdet, dzdet=dzdet)


_small_tables = {}     # (content, shape, dtype, device) -> device tensor: attenuation / spectrum tables of a few KB


def _as_device(a, dtype, device):
    """Device tensor of ``dtype`` for a tensor or host array.  Small host tables (<= 64 KiB: mu, W, i0 -- read-only inputs
    of the kernels) are uploaded once per content: a pageable upload is a synchronous copy, and an end-to-end step makes
    five of them."""
    if isinstance(a, torch.Tensor):
        return a.to(device=device, dtype=dtype).contiguous()
    a = np.ascontiguousarray(a)
    if a.nbytes > 65536:
        return torch.as_tensor(a, device=device).to(dtype).contiguous()
    import hashlib
    key = (hashlib.blake2b(a.tobytes(), digest_size=16).digest(), a.shape, a.dtype.str, dtype, str(device))
    t = _small_tables.get(key)
    if t is None:
        if len(_small_tables) >= 64:
            _small_tables.clear()
        t = torch.as_tensor(np.array(a), device=device).to(dtype).contiguous()
        _small_tables[key] = t
    return t


class Projector:
    """Owns the device copy of the label volume and the prepared workspace for one (scan, volume) pair."""

    def __init__(self, scan: DexctScan, labels, voxel, device=None, max_label=None):
        """``max_label``: the largest label value, if the caller already knows it (saves one reduction + synchronisation at
        the first projection, which checks it against the number of rows of the attenuation table)."""
        N.require_device()
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        lab = labels if isinstance(labels, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(labels, dtype=np.uint8))
        if lab.dtype != torch.uint8:
            raise TypeError('labels must be uint8')
        if lab.dim() == 2:
            lab = lab[None]
        if lab.dim() != 3:
            raise ValueError('labels must be [nz, ny, nx] or [ny, nx]')
        self.labels = lab.to(self.device).contiguous()
        nz, ny, nx = self.labels.shape
        self.scan = scan
        self.vol = DexctVolume(nx=nx, ny=ny, nz=nz, reserved=0, dx=float(voxel[0]), dy=float(voxel[1]), dz=float(voxel[2]))
        nbytes = N.lib().dexct_siddon_workspace_bytes(C.byref(self.scan), C.byref(self.vol))
        if nbytes == 0:
            raise N.DexctError('invalid scan/volume sizes')
        self.workspace = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        self.prepare()
        self._max_label = None if max_label is None else int(max_label)

    def prepare(self):
        """Re-run after the label tensor was modified in place."""
        self._max_label = None
        with torch.cuda.device(self.device):
            N.check(N.lib().dexct_siddon_prepare(C.byref(self.scan), C.byref(self.vol), N.ptr(self.labels),
                                                 N.ptr(self.workspace), N.stream_ptr()))

    @property
    def rays_per_view(self):
        return self.scan.n_rows * self.scan.n_ch

    def tables_host(self):
        """The sin/cos/z tables exactly as the device uses them (for checkers)."""
        s = self.scan
        out = dict(cos_beta=np.zeros(s.n_views), sin_beta=np.zeros(s.n_views), cos_gamma=np.zeros(s.n_ch),
                   sin_gamma=np.zeros(s.n_ch), zdet=np.zeros(s.n_rows))
        with torch.cuda.device(self.device):
            N.check(N.lib().dexct_siddon_tables_host(
                C.byref(s), N.ptr(self.workspace), *[out[k].ctypes.data_as(C.c_void_p) for k in
                                                     ('cos_beta', 'sin_beta', 'cos_gamma', 'sin_gamma', 'zdet')],
                N.stream_ptr()))
        return out

    def project(self, mu, W, view_begin=0, view_count=None, dtype=F32, want_log=True, want_L=False,
                counts=None, logs=None, L=None):
        """counts/logs [nSpec, view_count, n_rows, n_ch], L [view_count, n_rows, n_ch, K] (or None)."""
        code = dtype_code(dtype)
        td = _TORCH_DTYPE[code]
        s = self.scan
        view_count = s.n_views - view_begin if view_count is None else view_count
        mu = _as_device(mu, td, self.device)
        W = _as_device(W, td, self.device)
        if mu.dim() != 2 or W.dim() != 2 or mu.shape[1] != W.shape[1]:
            raise ValueError('mu must be [K, nE] and W [nSpec, nE] on the same energy grid')
        K, nE = mu.shape
        nS = W.shape[0]
        # the kernels index their per-material accumulators with the label itself (include/dexct.h: labels >= K are a
        # caller error), so a phantom that holds a label without a row in ``mu`` is refused here
        if self._max_label is None:
            self._max_label = int(self.labels.max())
        if self._max_label >= K:
            raise ValueError(f'the phantom holds label {self._max_label} but the attenuation table has only {K} rows')
        shape = (nS, view_count, s.n_rows, s.n_ch)
        if counts is None:
            counts = torch.empty(shape, dtype=td, device=self.device)
        if want_log and logs is None:
            logs = torch.empty(shape, dtype=td, device=self.device)
        if want_L and L is None:
            L = torch.empty((view_count, s.n_rows, s.n_ch, K), dtype=td, device=self.device)
        with torch.cuda.device(self.device):
            N.check(N.lib().dexct_siddon_poly(C.byref(s), C.byref(self.vol), N.ptr(self.workspace), view_begin, view_count,
                                              N.ptr(mu), K, nE, N.ptr(W), nS, N.ptr(counts),
                                              N.ptr(logs if want_log else None), N.ptr(L if want_L else None),
                                              code, N.stream_ptr()))
        return counts, (logs if want_log else None), (L if want_L else None)

    def trace(self, ray_ids, max_seg=4096, dtype=F32):
        """Ordered voxel sequence of the given rays: list of (vox int64[], lab uint8[], len float[]) numpy triples."""
        code = dtype_code(dtype)
        td = _TORCH_DTYPE[code]
        ids = torch.as_tensor(np.asarray(ray_ids, dtype=np.int64), device=self.device)
        n = ids.numel()
        vox = torch.empty((n, max_seg), dtype=torch.int64, device=self.device)
        lab = torch.empty((n, max_seg), dtype=torch.uint8, device=self.device)
        ln = torch.empty((n, max_seg), dtype=td, device=self.device)
        ns = torch.empty(n, dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            N.check(N.lib().dexct_siddon_trace(C.byref(self.scan), C.byref(self.vol), N.ptr(self.workspace), N.ptr(ids), n,
                                               max_seg, N.ptr(vox), N.ptr(lab), N.ptr(ln), N.ptr(ns), code, N.stream_ptr()))
        ns_h = ns.cpu().numpy()
        if (ns_h > max_seg).any():
            raise N.DexctError(f'trace overflow: a ray has {ns_h.max()} segments, max_seg={max_seg}')
        vox_h, lab_h, ln_h = vox.cpu().numpy(), lab.cpu().numpy(), ln.cpu().numpy()
        return [(vox_h[t, :ns_h[t]], lab_h[t, :ns_h[t]], ln_h[t, :ns_h[t]]) for t in range(n)]


    def walk_trace(self, ray_ids, K, max_steps=2048, dtype=F32, steps=True):
        """Steps of the PRODUCTION walk over the skip layouts for the given rays (dexct_siddon_walk_trace).

        Returns ``(rays, L)``: ``L`` [n, K] per-material lengths summed by the walk, and per ray a dict with ``meta``
        (xmajor, fwd, mpos, hit) and -- with ``steps`` -- ``col`` int32[s], ``run`` int32[s], ``vox`` int64[s, 3],
        ``lab`` uint8[s, 3], ``len`` float[s, 3]; without ``steps`` only ``n_steps``."""
        code = dtype_code(dtype)
        td = _TORCH_DTYPE[code]
        ids = torch.as_tensor(np.asarray(ray_ids, dtype=np.int64), device=self.device)
        n = ids.numel()
        ns = torch.empty(n, dtype=torch.int32, device=self.device)
        meta = torch.empty(n, dtype=torch.int32, device=self.device)
        L = torch.empty((n, K), dtype=td, device=self.device)
        col = run = vox = lab = ln = None
        if steps:
            col = torch.empty((n, max_steps), dtype=torch.int32, device=self.device)
            run = torch.empty((n, max_steps), dtype=torch.int32, device=self.device)
            vox = torch.empty((n, max_steps, 3), dtype=torch.int64, device=self.device)
            lab = torch.empty((n, max_steps, 3), dtype=torch.uint8, device=self.device)
            ln = torch.empty((n, max_steps, 3), dtype=td, device=self.device)
        with torch.cuda.device(self.device):
            N.check(N.lib().dexct_siddon_walk_trace(C.byref(self.scan), C.byref(self.vol), N.ptr(self.workspace), N.ptr(ids), n,
                                                    max_steps if steps else 0, int(K), N.ptr(col), N.ptr(run), N.ptr(vox), N.ptr(lab),
                                                    N.ptr(ln), N.ptr(ns), N.ptr(meta), N.ptr(L), code, N.stream_ptr()))
        ns_h, meta_h = ns.cpu().numpy(), meta.cpu().numpy()
        if steps and (ns_h > max_steps).any():
            raise N.DexctError(f'walk trace overflow: a ray takes {ns_h.max()} steps, max_steps={max_steps}')
        out = []
        if steps:
            col_h, run_h, vox_h, lab_h, ln_h = (x.cpu().numpy() for x in (col, run, vox, lab, ln))
        for t in range(n):
            m = int(meta_h[t])
            rec = dict(xmajor=bool(m & 1), fwd=bool(m & 2), mpos=bool(m & 4), hit=bool(m & 8), n_steps=int(ns_h[t]))
            if steps:
                k = ns_h[t]
                rec.update(col=col_h[t, :k], run=run_h[t, :k], vox=vox_h[t, :k], lab=lab_h[t, :k], len=ln_h[t, :k])
            out.append(rec)
        return out, L


def count_segments(pj, view_begin=0, view_count=None, chunk=1 << 22):
    """Total number of voxel segments (pieces of positive length) of the rays of a view range -- the unit of the
    bytes-per-ray roofline."""
    s = pj.scan
    view_count = s.n_views - view_begin if view_count is None else view_count
    lo = view_begin * pj.rays_per_view
    hi = lo + view_count * pj.rays_per_view
    total = 0
    with torch.cuda.device(pj.device):
        for a in range(lo, hi, chunk):
            ids = torch.arange(a, min(a + chunk, hi), dtype=torch.int64, device=pj.device)
            ns = torch.empty(ids.numel(), dtype=torch.int32, device=pj.device)
            N.check(N.lib().dexct_siddon_trace(C.byref(s), C.byref(pj.vol), N.ptr(pj.workspace), N.ptr(ids), ids.numel(), 0,
                                               None, None, None, N.ptr(ns), F32, N.stream_ptr()))
            total += int(ns.sum(dtype=torch.int64))
    return total


def poly_detect(L, mu, W, dtype=F32, want_log=True):
    """counts, logs [nSpec, *L.shape[:-1]] from stored lengths L [..., K] (device tensor)."""
    N.require_device()
    code = dtype_code(dtype)
    td = _TORCH_DTYPE[code]
    dev = L.device
    L2 = L.to(td).contiguous()
    mu = _as_device(mu, td, dev)
    W = _as_device(W, td, dev)
    K, nE = mu.shape
    nS = W.shape[0]
    n_rays = L2.numel() // K
    counts = torch.empty((nS,) + tuple(L2.shape[:-1]), dtype=td, device=dev)
    logs = torch.empty_like(counts) if want_log else None
    with torch.cuda.device(dev):
        N.check(N.lib().dexct_poly_detect(N.ptr(L2), n_rays, N.ptr(mu), K, nE, N.ptr(W), nS, N.ptr(counts), N.ptr(logs),
                                          code, N.stream_ptr()))
    return counts, logs


def gn_air_key(g1, dtype=F32, reduce_max_over_ranks=False, peers=None):
    """Monotone integer key of ``max(g1)`` -- the maximum of the first measurement, which sets the air threshold of
    get_basismat_sinos (/root/reference/matdecomp.py:196) -- as an 8-byte device tensor for ``gn_decompose(air_key=...)``.
    Lets a caller decompose a sinogram in several pieces (e.g. to start sending finished pieces to other ranks) against
    the threshold of the WHOLE sinogram.  ``reduce_max_over_ranks``: MAX over all ranks (``peers``: through flags in peer
    memory, else one NCCL all-reduce)."""
    N.require_device()
    code = dtype_code(dtype)
    if not (isinstance(g1, torch.Tensor) and g1.is_cuda and g1.dtype == _TORCH_DTYPE[code] and g1.is_contiguous()):
        raise ValueError('g1 must be a contiguous CUDA tensor of the compute dtype')
    scratch = torch.empty(16, dtype=torch.uint8, device=g1.device)
    with torch.cuda.device(g1.device):
        N.check(N.lib().dexct_gn_air_max(N.ptr(g1), g1.numel(), N.ptr(scratch), code, N.stream_ptr()))
        if reduce_max_over_ranks:
            if peers is not None:
                peers.max_key(scratch, code)
            else:
                import torch.distributed as dist
                key = scratch[:8].view(torch.int64) if code == F64 else scratch[:4].view(torch.int32)
                dist.all_reduce(key, op=dist.ReduceOp.MAX)
    return scratch[:8]


def gn_decompose(g, i0, mus, n_iters, mask_thresh=0.0, dtype=F32, out=None, reduce_max_over_ranks=False, peers=None,
                 mirrors=None, push=None, air_key=None):
    """g [2, ...] counts (device tensor or array), i0 [2, nE], mus [2, nE] -> a [..., 2] device tensor.

    ``reduce_max_over_ranks``: the pixels are one shard of a larger sinogram; the air threshold then uses the maximum of the
    first measurement over ALL ranks (one 8-byte integer MAX all-reduce), as the unsharded reference does
    (matdecomp.py:196) -- through ``peers`` (a dexct.peer.PeerGroup: flags in peer memory) when given, else NCCL.
    ``mirrors``: ctypes table of further device addresses (peer copies of the basis sinogram, SymBuffer.table) that the
    kernel's epilogue stores every pixel to as well.  ``push``: ``(src_tensor, table)`` -- a finished contiguous buffer
    (the count/log shard) that the same kernel copies to the addresses in ``table`` while it computes.
    ``air_key``: result of :func:`gn_air_key` -- the threshold then comes from that maximum instead of ``max(g[0])``."""
    N.require_device()
    code = dtype_code(dtype)
    td = _TORCH_DTYPE[code]
    dev = g.device if isinstance(g, torch.Tensor) and g.is_cuda else torch.device('cuda', torch.cuda.current_device())
    g = _as_device(g, td, dev)
    i0 = _as_device(i0, td, dev)
    mus = _as_device(mus, td, dev)
    if g.shape[0] != 2 or i0.shape[0] != 2 or mus.shape[0] != 2 or i0.shape != mus.shape:
        raise ValueError('g must be [2, ...], i0 and mus [2, nE]')
    P = g[0].numel()
    nE = i0.shape[1]
    if out is None:
        out = torch.empty(tuple(g.shape[1:]) + (2,), dtype=td, device=dev)
    scratch = torch.empty(N.lib().dexct_gn_scratch_bytes(P, code), dtype=torch.uint8, device=dev)
    have_max = 0
    with torch.cuda.device(dev):
        if air_key is not None and mask_thresh > 0:
            scratch[:8].copy_(air_key)
            have_max = 1
        elif reduce_max_over_ranks and mask_thresh > 0:
            import torch.distributed as dist
            N.check(N.lib().dexct_gn_air_max(N.ptr(g), P, N.ptr(scratch), code, N.stream_ptr()))
            if peers is not None:
                peers.max_key(scratch, code)
            else:
                key = scratch[:8].view(torch.int64) if code == F64 else scratch[:4].view(torch.int32)
                dist.all_reduce(key, op=dist.ReduceOp.MAX)
            have_max = 1
        if push is not None and len(push[1]):
            src, tab = push
            if not src.is_contiguous():
                raise ValueError('push source must be contiguous')
            nm = 0 if mirrors is None else len(mirrors)
            N.check(N.lib().dexct_gn_decompose_fused_push(N.ptr(g), N.ptr(i0), N.ptr(mus), nE, P, int(n_iters), float(mask_thresh),
                                                          N.ptr(out), mirrors if nm else None, nm, N.ptr(src),
                                                          src.numel() * src.element_size(), tab, len(tab), N.ptr(scratch),
                                                          have_max, code, N.stream_ptr()))
        elif mirrors is not None and len(mirrors):
            N.check(N.lib().dexct_gn_decompose_mirrored(N.ptr(g), N.ptr(i0), N.ptr(mus), nE, P, int(n_iters), float(mask_thresh),
                                                        N.ptr(out), mirrors, len(mirrors), N.ptr(scratch), have_max, code,
                                                        N.stream_ptr()))
        else:
            N.check(N.lib().dexct_gn_decompose(N.ptr(g), N.ptr(i0), N.ptr(mus), nE, P, int(n_iters), float(mask_thresh),
                                               N.ptr(out), N.ptr(scratch), have_max, code, N.stream_ptr()))
    gn_decompose.last_scratch = scratch          # diagnostics: bytes 8..12 = pixels redone in float64 (single precision)
    return out


def gn_decompose_general(g, i0, mus, n_iters, dtype=F32):
    """optimize_sino_cpu in its general form (/root/reference/matdecomp.py:87-127): g [nMeas, nViews, nBins] counts,
    i0 [nMeas, nBins, nE] (a spectrum per detector bin) or [nMeas, nE], mus [nMats, nE], 1..4 measurements and materials
    -> a [nViews, nBins, nMats] device tensor.  float64 arithmetic inside; no air mask."""
    N.require_device()
    code = dtype_code(dtype)
    td = _TORCH_DTYPE[code]
    dev = g.device if isinstance(g, torch.Tensor) and g.is_cuda else torch.device('cuda', torch.cuda.current_device())
    g = _as_device(g, td, dev)
    i0 = _as_device(i0, td, dev)
    mus = _as_device(mus, td, dev)
    if g.dim() != 3:
        raise ValueError('g must be [nMeas, nViews, nBins]')
    n_meas, n_views, n_bins = g.shape
    n_mats, nE = mus.shape
    per_bin = i0.dim() == 3
    if i0.shape[0] != n_meas or i0.shape[-1] != nE or (per_bin and i0.shape[1] != n_bins):
        raise ValueError(f'i0 must be [{n_meas}, {n_bins}, {nE}] or [{n_meas}, {nE}], got {tuple(i0.shape)}')
    out = torch.empty((n_views, n_bins, n_mats), dtype=td, device=dev)
    with torch.cuda.device(dev):
        N.check(N.lib().dexct_gn_decompose_general(N.ptr(g), N.ptr(i0), N.ptr(mus), n_meas, n_mats, nE, n_bins, int(per_bin),
                                                   n_views * n_bins, int(n_iters), N.ptr(out), code, N.stream_ptr()))
    return out


def gn_fixup_count():
    """Pixels the most recent single-precision ``gn_decompose`` call handed to its float64 pass (synchronises)."""
    scratch = getattr(gn_decompose, 'last_scratch', None)
    if scratch is None or scratch.numel() <= 16:
        return 0
    return int(scratch[8:12].view(torch.int32).item())


def fbp_fan(scan, sino, n_matrix, fov, ramp, view_begin=0, dtype=F32, image=None, accumulate=False):
    """Fan-beam FBP of the views ``[view_begin, view_begin + sino.shape[0])`` of ``scan`` (replaces the numerics of
    get_recon, /root/reference/main.py:134,168).

    sino: device tensor ``[views, n_ch]`` or ``[views, n_rows, n_ch]`` holding exactly those views.
    Returns the image ``[n_rows, n_matrix, n_matrix]`` on the device (added to ``image`` when ``accumulate``).
    """
    N.require_device()
    code = dtype_code(dtype)
    td = _TORCH_DTYPE[code]
    if not (isinstance(sino, torch.Tensor) and sino.is_cuda):
        raise TypeError('sino must be a CUDA tensor')
    dev = sino.device
    if sino.dim() == 2:
        sino = sino[:, None, :]
    if sino.dim() != 3 or sino.shape[1] != scan.n_rows or sino.shape[2] != scan.n_ch:
        raise ValueError(f'sino must be [views, {scan.n_rows}, {scan.n_ch}], got {tuple(sino.shape)}')
    sino = sino.to(td).contiguous()
    nv = sino.shape[0]
    shape = (scan.n_rows, int(n_matrix), int(n_matrix))
    if image is None:
        if accumulate:
            raise ValueError('accumulate needs an image to add to')
        image = torch.empty(shape, dtype=td, device=dev)
    elif tuple(image.shape) != shape or image.dtype != td or not image.is_contiguous():
        raise ValueError(f'image must be a contiguous {td} tensor of shape {shape}')
    with torch.cuda.device(dev):
        nbytes = N.lib().dexct_fbp_workspace_bytes(C.byref(scan), max(nv, 1), code)
        scratch = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev)
        N.check(N.lib().dexct_fbp_fan(C.byref(scan), N.ptr(sino), int(view_begin), nv, int(n_matrix), float(fov),
                                      float(ramp), N.ptr(image), int(bool(accumulate)), N.ptr(scratch), code,
                                      N.stream_ptr()))
    return image


def hounsfield(image, mu_water, dtype=F32, out=None):
    """1000 (image - mu_water) / mu_water on the device."""
    N.require_device()
    code = dtype_code(dtype)
    if image.dtype != _TORCH_DTYPE[code] or not image.is_contiguous():
        raise ValueError('image must be a contiguous tensor of the compute dtype')
    out = torch.empty_like(image) if out is None else out
    with torch.cuda.device(image.device):
        N.check(N.lib().dexct_fbp_hu(N.ptr(image), image.numel(), float(mu_water), N.ptr(out), code, N.stream_ptr()))
    return out


def poisson_noise(counts, seed, stream_id=0, offset=0, logs=None, air_counts=1.0, floor_counts=0.5, dtype=F32):
    """In place: every element of ``counts`` becomes a Poisson variate of that mean (the
    spectra are scaled to photons per pixel per view at /root/reference/main.py:68).  ``offset`` is the index of
    ``counts.view(-1)[0]`` in the whole sinogram, so view shards draw what a single GPU would.  ``logs`` (optional, same
    shape) receives ``-ln(max(counts, floor_counts) / air_counts)``."""
    N.require_device()
    code = dtype_code(dtype)
    if not (counts.is_cuda and counts.is_contiguous() and counts.dtype == _TORCH_DTYPE[code]):
        raise ValueError('counts must be a contiguous CUDA tensor of the compute dtype')
    if logs is not None and (logs.shape != counts.shape or logs.dtype != counts.dtype or not logs.is_contiguous()
                             or logs.device != counts.device):
        raise ValueError('logs must match counts in shape, dtype and device')
    with torch.cuda.device(counts.device):
        N.check(N.lib().dexct_poisson_noise(N.ptr(counts), N.ptr(logs) if logs is not None else None, counts.numel(),
                                            int(seed) & 0xFFFFFFFFFFFFFFFF, int(stream_id) & 0xFFFFFFFF, int(offset),
                                            float(air_counts), float(floor_counts), code, N.stream_ptr()))
    return counts


def vmi(m1, m2, u1, u2, u_water, hu=True, dtype=F32):
    """Virtual monoenergetic image ``u1 m1 + u2 m2`` (in HU when ``hu``) from two basis-material images on the device;
    float32 result like the reference's make_vmi (/root/reference/plots.py:136-144)."""
    N.require_device()
    code = dtype_code(dtype)
    td = _TORCH_DTYPE[code]
    if not (m1.is_cuda and m1.shape == m2.shape and m1.dtype == td and m2.dtype == td and m1.device == m2.device):
        raise ValueError('m1 and m2 must be CUDA tensors of the compute dtype and equal shape')
    m1, m2 = m1.contiguous(), m2.contiguous()
    out = torch.empty(m1.shape, dtype=torch.float32, device=m1.device)
    with torch.cuda.device(m1.device):
        N.check(N.lib().dexct_vmi(N.ptr(m1), N.ptr(m2), m1.numel(), float(u1), float(u2), float(u_water), int(bool(hu)),
                                  N.ptr(out), code, N.stream_ptr()))
    return out
