"""Drop-in ``matdecomp``: dual-energy basis-material decomposition on the B200.

Same public names as /root/reference/matdecomp.py (``get_basismat_sinos``, ``do_matdecomp_gn``, ``optimize_sino``,
``optimize_sino_cpu``, the basis-material constants) so that /root/reference/main.py:22,153 and plots.py:18 work
unchanged.  The per-pixel Newton iteration of matdecomp.py:87-127 runs in csrc/gn.cu; this module only assembles the
small spectral tables on the host (matdecomp.py:138-160) and moves arrays across the boundary.
"""
from __future__ import annotations

import numpy as np
import torch

import xcompy as xc
import xtomosim.xcompy as _xc_impl
from dexct import config, dist
from dexct.ops import dtype_code, gn_decompose

# Basis materials -- the same two ICRU compositions the reference hard-codes (matdecomp.py:11-17).
mat1 = 'ICRU tissue'
matcomp1 = 'H(10.2)C(14.3)N(3.4)O(70.8)Na(0.2)P(0.3)S(0.3)Cl(0.2)K(0.3)'
density1 = 1.06  # [g/cm3]
mat2 = 'ICRU bone'
matcomp2 = 'H(3.4)C(15.5)N(4.2)O(43.5)Na(0.1)Mg(0.2)P(10.3)S(0.3)Ca(22.5)'
density2 = 1.92  # [g/cm3]


_tables_memo = {}      # content key -> (ee, i0, mus); a handful of entries (one per spectrum pair of a run)


def _content_key(*arrays):
    import zlib
    return tuple((a.shape, zlib.crc32(np.ascontiguousarray(a, dtype=np.float64).tobytes())) for a in map(np.asarray, arrays))


def spectral_tables(ct, spec1, spec2):
    """Host-side model tables of the decomposition (matdecomp.py:138-160).

    Union energy grid of the two spectra; first bin width runs from 0; detector efficiency interpolated onto the grid
    (times E for an energy-integrating detector); spectra interpolated with numpy's end-value clamping; basis rows are
    MASS attenuation (the reference leaves the density multiply commented out, so the outputs are g/cm^2).
    Returns ee [nE], i0 [2, nE], mus [2, nE] (memoised on the CONTENT of the inputs -- a few hundred numbers -- so that
    repeated decompositions of one acquisition do not rebuild them; the caller gets its own copies).
    """
    key = (_content_key(spec1.E, spec1.I0, spec2.E, spec2.I0, ct.det_E, ct.det_eta_E), bool(ct.eid), matcomp1, matcomp2,
           id(xc.mixatten), _xc_impl.generation)
    hit = _tables_memo.get(key)
    if hit is not None:
        return tuple(a.copy() for a in hit)
    ee = np.unique(np.concatenate([np.asarray(spec1.E), np.asarray(spec2.E)]))
    dE = np.diff(ee, prepend=0.0)
    resp = np.interp(ee, ct.det_E, ct.det_eta_E)
    if ct.eid:
        resp = resp * ee
    i0 = np.stack([np.interp(ee, s.E, s.I0) * resp * dE for s in (spec1, spec2)])
    mus = np.stack([xc.mixatten(mc, ee) for mc in (matcomp1, matcomp2)])
    if len(_tables_memo) >= 16:
        _tables_memo.clear()
    _tables_memo[key] = (ee.copy(), i0.copy(), mus.copy())
    return ee, i0, mus


def _counts_to_device(sinos, td):
    """Two count sinograms (device tensors or host arrays of any float type) -> one device tensor [2, ...] of dtype
    ``td``.  Host arrays are uploaded as they are (one DMA each when they sit in page-locked memory, as the arrays
    returned by get_sino do) and converted on the device; nothing is stacked or converted on the host."""
    dev = torch.device('cuda', torch.cuda.current_device())
    first = sinos[0]
    g = torch.empty((2,) + tuple(first.shape), dtype=td, device=dev)
    for m, s in enumerate(sinos):
        if not isinstance(s, torch.Tensor):
            s = np.asarray(s)
            if s.dtype not in (np.float32, np.float64) or not s.flags.c_contiguous:
                s = np.ascontiguousarray(s, dtype=np.float64)
            s = torch.from_numpy(s)
        if tuple(s.shape) != tuple(first.shape):
            raise ValueError('the two sinograms must have the same shape')
        g[m].copy_(s, non_blocking=True)
    return g


def _host_f64(t):
    """Device tensor -> float64 numpy array in page-locked memory (conversion on the device, one DMA)."""
    h = torch.empty(t.shape, dtype=torch.float64, pin_memory=True)
    h.copy_(t, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return h.numpy()


PIPELINE_MIN_BYTES = 512 << 20       # host sinograms at least this large are decomposed in chunks (see below)


def _decompose_host_pipelined(sinos, i0, mus, n_iters, mask_thresh, code, reduce_max_over_ranks=False):
    """Two LARGE host count sinograms -> float64 host result, in chunks of the leading axis: the upload of the next chunks
    (stream `up`), the Newton iterations of the current one and the float64 download of the finished ones (stream `down`)
    run side by side -- the reference-shaped API moves 8 bytes per pixel and sinogram in and out (17.7 GB for configuration
    C5), which costs more time on the PCIe link than the kernel takes.  The air threshold needs the maximum of the WHOLE
    first measurement (matdecomp.py:196), so the first measurement goes up first and its chunk maxima are combined
    (gn_air_key) before the first chunk is decomposed.  Pixels are independent: same results as the one-shot path."""
    from dexct.ops import gn_air_key
    td = torch.float32 if code == 0 else torch.float64
    dev = torch.device('cuda', torch.cuda.current_device())
    host = []
    for sg in sinos:
        a = np.asarray(sg)
        if a.dtype not in (np.float32, np.float64) or not a.flags.c_contiguous:
            a = np.ascontiguousarray(a, dtype=np.float64)
        host.append(torch.from_numpy(a))
    if tuple(host[0].shape) != tuple(host[1].shape):
        raise ValueError('the two sinograms must have the same shape')
    shape = tuple(host[0].shape)
    n0 = shape[0]
    n_chunks = int(min(16, max(2, -(-host[0].numel() * 8 // (1 << 30)))))
    step = -(-n0 // n_chunks)
    spans = [(a, min(step, n0 - a)) for a in range(0, n0, step)]
    out_h = torch.empty(shape + (2,), dtype=torch.float64, pin_memory=True)
    up, down = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    main = torch.cuda.current_stream(dev)
    bufs = [torch.empty((2, n) + shape[1:], dtype=td, device=dev) for _, n in spans]
    ev = [[torch.cuda.Event(), torch.cuda.Event()] for _ in spans]
    with torch.cuda.stream(up):
        up.wait_stream(main)
        for m in range(2):                                # all of measurement 1 first: its maximum sets the air threshold
            for k, (a, n) in enumerate(spans):
                bufs[k][m].copy_(host[m][a:a + n], non_blocking=True)
                ev[k][m].record()
    for b in bufs:
        b.record_stream(up)
    key = None
    if mask_thresh > 0:
        for k in range(len(spans)):
            main.wait_event(ev[k][0])
            kk = gn_air_key(bufs[k][0], dtype=code)
            kv = kk.view(torch.int64) if code == 1 else kk[:4].view(torch.int32)
            if key is None:
                key, key_v = kk, kv
            else:
                torch.maximum(key_v, kv, out=key_v)
        if reduce_max_over_ranks:                         # view shards: the threshold of the unsharded sinogram
            import torch.distributed as tdist
            tdist.all_reduce(key_v, op=tdist.ReduceOp.MAX)
    for k, (a, n) in enumerate(spans):
        main.wait_event(ev[k][0])
        main.wait_event(ev[k][1])
        o = gn_decompose(bufs[k], i0, mus, n_iters, mask_thresh=mask_thresh, dtype=code, air_key=key)
        done = torch.cuda.Event()
        done.record()
        with torch.cuda.stream(down):
            down.wait_event(done)
            out_h[a:a + n].copy_(o, non_blocking=True)
        o.record_stream(down)
    down.synchronize()
    return out_h.numpy()


def optimize_sino(Sino_gg, ee, i0, mus, n_iters, verbose=True, dtype=None, mask_thresh=0.0, to_host=True,
                  reduce_max_over_ranks=False):
    """GPU Newton iterations; signature of matdecomp.py:20 (``ee`` is unused there too).

    Sino_gg [nMeas, nViews, nBins] counts (array, device tensor, or a sequence of them); i0 [nMeas, nBins, nE] or
    [nMeas, nE]; mus [nMats, nE].  Returns [nViews, nBins, nMats] -- float64 numpy by default, the device tensor with
    ``to_host=False``.  Two measurements, two basis materials and one spectrum for all bins (what do_matdecomp_gn
    builds, matdecomp.py:151) take the fast kernels; a spectrum per detector bin or other counts of measurements /
    materials (1..4; matdecomp.py:95,101-105,118,125 allow them) take the general float64 kernel.
    """
    del ee, verbose
    from dexct import _native
    _native.require_device()
    code = dtype_code(config.dtype if dtype is None else dtype)
    td = torch.float32 if code == 0 else torch.float64
    i0 = np.asarray(i0, dtype=np.float64)
    mus = np.asarray(mus, dtype=np.float64)
    per_bin = i0.ndim == 3 and i0.shape[1] > 1 and not np.array_equal(i0[:, :1, :].repeat(i0.shape[1], axis=1), i0)
    n_meas = Sino_gg.shape[0] if hasattr(Sino_gg, 'shape') else len(Sino_gg)
    if per_bin or mus.shape[0] != 2 or n_meas != 2:
        # the reference's general form (matdecomp.py:95,101-105,118,125): a spectrum per detector bin, any number of
        # measurements / basis materials (1..4) -- the general kernel; get_basismat_sinos never comes here
        from dexct.ops import gn_decompose_general
        if mask_thresh > 0 or reduce_max_over_ranks:
            raise ValueError('the air mask belongs to get_basismat_sinos (two materials, one spectrum for all bins)')
        g = Sino_gg if isinstance(Sino_gg, torch.Tensor) else np.ascontiguousarray(np.stack([np.asarray(s) for s in Sino_gg]))
        if g.ndim != 3:
            raise ValueError('Sino_gg must be [nMeas, nViews, nBins]')
        if not per_bin and i0.ndim == 3:
            i0 = i0[:, 0, :]
        out = gn_decompose_general(g, i0, mus, n_iters, dtype=code)
        return _host_f64(out) if to_host else out
    if i0.ndim == 3:
        i0 = i0[:, 0, :]
    if isinstance(Sino_gg, torch.Tensor) and Sino_gg.is_cuda:
        g = Sino_gg
    else:
        if len(Sino_gg) != 2:
            raise ValueError('Sino_gg must hold two measurements')
        first = Sino_gg[0]
        if (to_host and not isinstance(first, torch.Tensor) and not isinstance(Sino_gg[1], torch.Tensor)
                and np.ndim(first) >= 2 and np.shape(first)[0] >= 2 and np.size(first) * 8 >= PIPELINE_MIN_BYTES):
            return _decompose_host_pipelined([Sino_gg[0], Sino_gg[1]], i0, mus, n_iters, mask_thresh, code, reduce_max_over_ranks)
        g = _counts_to_device([Sino_gg[0], Sino_gg[1]], td)
    out = gn_decompose(g, i0, mus, n_iters, mask_thresh=mask_thresh, dtype=code,
                       reduce_max_over_ranks=reduce_max_over_ranks)
    return _host_f64(out) if to_host else out


# The reference keeps a CPU twin because its cupy version misbehaved (matdecomp.py:89-91); here both names run the
# same kernel -- there is no CPU implementation in this package.
optimize_sino_cpu = optimize_sino


def do_matdecomp_gn(ct, sino1, sino2, spec1, spec2, n_iters, mask_thresh=0.0, to_host=True):
    """matdecomp.py:130-164.  With a process group and ``config.shard_views`` the pixels are split by view range,
    decomposed without communication and the result all-gathered."""
    ee, i0, mus = spectral_tables(ct, spec1, spec2)
    rank, ws = dist.world() if config.shard_views else (0, 1)
    if ws > 1:
        n_views = ct.N_proj
        v0, nv = dist.shard_range(n_views, rank, ws)
        if sino1.shape[0] == n_views:
            part = (sino1[v0:v0 + nv], sino2[v0:v0 + nv])
        elif sino1.shape[0] == nv and not config.gather_views:
            part = (sino1, sino2)                        # the caller holds only its own view range (config.gather_views off)
        else:
            raise ValueError(f'sinograms have {sino1.shape[0]} views: expected the whole scan ({n_views})'
                             + ('' if config.gather_views else f' or this rank\'s range ({nv})'))
        # the air threshold needs the maximum over the whole sinogram: one integer MAX all-reduce between the shards
        if to_host and not config.gather_views:          # the rank's own range, host to host (chunked when large)
            return optimize_sino(part, ee, i0, mus, n_iters, mask_thresh=mask_thresh, to_host=True, reduce_max_over_ranks=True)
        local = optimize_sino(part, ee, i0, mus, n_iters, mask_thresh=mask_thresh, to_host=False, reduce_max_over_ranks=True)
        full = dist.all_gather_views(local, n_views, dim=0) if config.gather_views else local
        return _host_f64(full) if to_host else full
    return optimize_sino((sino1, sino2), ee, i0, mus, n_iters, mask_thresh=mask_thresh, to_host=to_host)


def get_basismat_sinos(ct, sino_raw_1, sino_raw_2, spec1, spec2, n_iters=30, mask_thresh=0.95):
    """Basis-material sinograms of a dual-energy acquisition (matdecomp.py:167-207).

    ``sino_raw_1/2``: photon counts ``[N_proj, N_channels]``; ``spec1/2``: the spectra they were acquired with.  Pixels
    whose first measurement reaches ``mask_thresh`` of its maximum (air) are zeroed, as in the reference
    (matdecomp.py:195-196,204-205) -- inside the kernel.  Returns two float64 arrays ``[N_proj, N_channels]``.
    """
    a = do_matdecomp_gn(ct, sino_raw_1, sino_raw_2, spec1, spec2, n_iters, mask_thresh=mask_thresh)
    return a[..., 0], a[..., 1]


def make_vmi(E0, M1, M2, HU=True, matcomp1=matcomp1, matcomp2=matcomp2):
    """Virtual monoenergetic image at ``E0`` keV from the two basis-material images (the reference's helper of the same
    name and signature, /root/reference/plots.py:136-144): ``u1 M1 + u2 M2``, in Hounsfield units when ``HU``; float32.
    The three attenuation lookups stay on the host (three numbers); the image arithmetic is one kernel."""
    from dexct.ops import vmi
    E = np.array([E0]).astype(np.float64)
    u1, u2 = float(xc.mixatten(matcomp1, E)[0]), float(xc.mixatten(matcomp2, E)[0])
    uw = float(1.0 * xc.mixatten('H(11.2)O(88.8)', E)[0])
    td = torch.float32 if config.dtype == 'f32' else torch.float64
    dev = torch.device('cuda', torch.cuda.current_device())
    m1 = torch.as_tensor(np.ascontiguousarray(M1), dtype=td, device=dev)
    m2 = torch.as_tensor(np.ascontiguousarray(M2), dtype=td, device=dev)
    return vmi(m1, m2, u1, u2, uw, hu=HU, dtype=dtype_code(config.dtype)).cpu().numpy()
