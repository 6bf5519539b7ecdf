"""Drop-in ``xtomosim.forward_project``: ``get_sino(ct, phantom, spec) -> (sino_raw, sino_log)``.

Call site in the reference: /root/reference/main.py:120; output contract (photon counts, ``[N_proj, N_channels]``)
from matdecomp.py:132-136,178-179.  The reference's implementation sits in an absent submodule; the algorithm is
Siddon 1985 (README.md:28,41) followed by the polyenergetic forward model the decomposition inverts
(matdecomp.py:116-118,146-150).  Here both run in one CUDA kernel (csrc/siddon.cu); numpy only crosses the boundary.
"""
from __future__ import annotations

import os
import weakref
import zlib

import numpy as np
import torch

from dexct import config, dist
from dexct.ops import Projector, make_scan, poisson_noise, poly_detect, dtype_code

_state = {}      # (id(ct), id(phantom)) -> _ScanState


class _ScanState:
    """Prepared device workspace of one (ct, phantom) pair and, optionally, its traced path lengths."""

    def __init__(self, ct, phantom, signature, staged=None):
        self.signature = signature
        self.ct_ref = weakref.ref(ct)
        self.ph_ref = weakref.ref(phantom)
        labels, voxel = _labels_for(ct, phantom, staged)
        scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, n_rows=ct.N_rows, h_iso=ct.h_iso,
                         rotation=ct.rotation, beta0=ct.beta0)
        # host label arrays of moderate size: their maximum costs microseconds here and saves the projector a device
        # reduction + synchronisation at its first projection
        hint = int(labels.max()) if isinstance(labels, np.ndarray) and 0 < labels.nbytes <= (64 << 20) else None
        self.projector = Projector(scan, labels, voxel, max_label=hint)
        self.L = None            # [views_local, n_rows, n_ch, K] on the device
        self.L_dtype = None


def _labels_for(ct, phantom, staged=None):
    """Single-row scans image slice ``z_index`` of the phantom (params.txt:12,16: Nz = 1, z_index = 0).  ``staged``: the
    device copy of ``phantom.M`` that _signature already made (large host volumes)."""
    M = phantom.M if staged is None else staged
    if ct.N_rows == 1:
        return M[phantom.z_index:phantom.z_index + 1], (phantom.dx, phantom.dy, phantom.dz)
    return M, (phantom.dx, phantom.dy, phantom.dz)


def _content_digest(lab):
    """Digest of the WHOLE label volume: an in-place edit of any voxel must invalidate the prepared layouts and the cached
    path lengths (a sampled digest once let a changed C4/C5 phantom reuse stale layouts without an error).

    Host arrays: CRC-32 of every byte; volumes above 64 MiB are cut into 32 MiB pieces whose CRCs are computed on a thread
    pool (zlib releases the GIL) and combined in order -- about 0.1 s for the 1 GiB labels of configuration C5.
    Device tensors: a position-weighted wrapping sum and an xor over the 8-byte words, computed on the device."""
    if isinstance(lab, torch.Tensor):
        flat = lab.reshape(-1)
        if not flat.is_contiguous():
            flat = flat.contiguous()
        n8 = flat.numel() // 8
        words = flat[:n8 * 8].view(torch.int64)
        tail = flat[n8 * 8:].to(torch.int64)
        acc = []
        if n8:
            w = torch.arange(1, 2 * n8, 2, dtype=torch.int64, device=flat.device)      # odd weights: position-dependent
            acc.append(int((words * w).sum().item()))
            x = words
            while x.numel() > 1:                                                   # xor fold (no xor reduction in torch)
                h = x.numel() // 2
                x = torch.cat([x[:h] ^ x[h:2 * h], x[2 * h:]])
            acc.append(int(x.item()))
        acc.extend(int(v) for v in tail.cpu())
        return ('device', tuple(acc))
    buf = memoryview(np.ascontiguousarray(lab)).cast('B')
    piece = 32 << 20
    if len(buf) <= 2 * piece:
        return zlib.crc32(buf)
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1)) as pool:
        crcs = list(pool.map(lambda o: zlib.crc32(buf[o:o + piece]), range(0, len(buf), piece)))
    return zlib.crc32(np.asarray(crcs, dtype=np.uint32).tobytes())


STAGE_MIN_BYTES = 64 << 20


def _signature(ct, phantom):
    """(signature, staged): large HOST volumes are digested on the device -- the upload (20 ms for the 1 GiB labels of
    configuration C5, from page-locked memory) is needed anyway whenever the workspace has to be rebuilt, and the device
    digest of the copy costs a millisecond where the threaded CRC of the host array cost 100 ms per get_sino call.  The copy
    is handed on to the projector when the signature turns out to be new, dropped otherwise."""
    lab = phantom.M
    staged = None
    if isinstance(lab, np.ndarray) and lab.nbytes >= STAGE_MIN_BYTES and lab.dtype == np.uint8 and torch.cuda.is_available():
        staged = torch.from_numpy(np.ascontiguousarray(lab)).to(torch.device('cuda', torch.cuda.current_device()), non_blocking=True)
        digest = _content_digest(staged)
    else:
        digest = _content_digest(lab)
    return (ct.N_proj, ct.N_channels, ct.N_rows, ct.gamma_fan, ct.SID, ct.SDD, ct.h_iso, ct.rotation, ct.beta0,
            tuple(lab.shape), phantom.dx, phantom.dy, phantom.dz, phantom.z_index, digest), staged


def _get_state(ct, phantom):
    key = (id(ct), id(phantom))
    sig, staged = _signature(ct, phantom)
    st = _state.get(key)
    if st is None or st.signature != sig or st.ct_ref() is not ct or st.ph_ref() is not phantom:
        for k in [k for k, v in _state.items() if v.ct_ref() is None or v.ph_ref() is None]:
            del _state[k]
        st = _ScanState(ct, phantom, sig, staged)
        _state[key] = st
    return st


def clear_cache():
    _state.clear()


def spectral_weights(ct, spec):
    """Detected weight per energy bin, W(E) = I0 * eta(E) * (E if energy-integrating) * dE -- the same expression the
    decomposition uses for its i0 (matdecomp.py:142,146-150), evaluated on the spectrum's own grid."""
    return spec.I0 * ct.det_response(spec.E) * spec.bin_widths()


def project_device(ct, phantom, specs, dtype=None, want_log=True):
    """Device-resident core of get_sino for one or several spectra.

    Returns (counts, logs) lists of tensors ``[views_local, n_rows, n_ch]`` -- the local view shard when a process group
    is initialised and ``config.shard_views`` is on, the whole scan otherwise -- plus (view_begin, view_count).
    """
    groups, span = _project_groups(ct, phantom, specs, dtype, want_log)
    counts_out, logs_out = [], []
    for c, lg in groups:
        for n in range(c.shape[0]):
            counts_out.append(c[n])
            logs_out.append(lg[n] if want_log else None)
    return counts_out, logs_out, span


def _project_groups(ct, phantom, specs, dtype=None, want_log=True):
    """One fused traversal per group of spectra that share an energy grid: list of (counts, logs) tensors
    ``[n_spectra_in_group, views_local, n_rows, n_ch]`` in the order of ``specs``."""
    code = dtype_code(config.dtype if dtype is None else dtype)
    st = _get_state(ct, phantom)
    rank, ws = dist.world() if config.shard_views else (0, 1)
    v0, nv = dist.shard_range(ct.N_proj, rank, ws)
    out = []
    same_grid = all(len(s.E) == len(specs[0].E) and np.array_equal(s.E, specs[0].E) for s in specs)
    groups = [list(range(len(specs)))] if (same_grid and len(specs) <= 4) else [[i] for i in range(len(specs))]
    for grp in groups:
        E = specs[grp[0]].E
        mu = phantom.mu_table(E)
        W = np.stack([spectral_weights(ct, specs[i]) for i in grp])
        have_L = config.cache_lengths and st.L is not None and st.L_dtype == code and st.L.shape[0] == nv
        if have_L:
            c, lg = poly_detect(st.L, mu, W, dtype=code, want_log=want_log)
        else:
            # a single-shot call (all spectra at once) has no second call to serve from stored lengths
            keep_L = config.cache_lengths and (len(groups) > 1 or len(specs) == 1)
            c, lg, L = st.projector.project(mu, W, view_begin=v0, view_count=nv, dtype=code, want_log=want_log,
                                            want_L=keep_L)
            if keep_L:
                st.L, st.L_dtype = L, code
        if noise_enabled(ct):
            _add_noise(ct, c, lg if want_log else None, W, v0, code)
        out.append((c, lg))
    return out, (v0, nv)


def noise_enabled(ct):
    flag = getattr(ct, 'noise', None)
    return config.noise if flag is None else bool(flag)


def _add_noise(ct, counts, logs, W, v0, code):
    """Poisson statistics on the counts ``[n_spectra, views_local, n_rows, n_ch]`` of one group, in place; the log
    sinograms are recomputed from the noisy counts.  Every acquisition (spectrum) takes the scanner's next random stream;
    elements are addressed by their index in the whole scan, so the result is independent of the view sharding."""
    seed = getattr(ct, 'noise_seed', None)
    seed = config.noise_seed if seed is None else int(seed)
    per_view = ct.N_rows * ct.N_channels
    for n in range(counts.shape[0]):
        stream_id = getattr(ct, 'noise_stream', 0)
        poisson_noise(counts[n], seed, stream_id=stream_id, offset=v0 * per_view, logs=None if logs is None else logs[n],
                      air_counts=float(np.sum(W[n])), floor_counts=config.noise_floor, dtype=code)
        ct.noise_stream = stream_id + 1


def _to_host(tensors, ct):
    """Device tensors ``[..., views_local, n_rows, n_ch]`` -> float64 numpy arrays like the reference's.

    View shards are gathered first (NCCL all-gather).  The conversion to float64 runs on the device and the copy lands
    in page-locked host memory (torch's caching host allocator: the block is recycled once the caller drops the array),
    one DMA per tensor and a single synchronisation for all of them.
    """
    hosts = []
    for t in tensors:
        if config.shard_views and config.gather_views:
            t = dist.all_gather_views(t, ct.N_proj, dim=t.dim() - 3)
        h = torch.empty(t.shape, dtype=torch.float64, pin_memory=True)
        h.copy_(t, non_blocking=True)
        hosts.append(h)
    torch.cuda.current_stream().synchronize()
    out = []
    for h in hosts:
        a = h.numpy()
        out.append(a[..., 0, :] if ct.N_rows == 1 else a)
    return out


PIPELINE_MIN_BYTES = 512 << 20       # host results at least this large are produced in view chunks (see _project_to_host)


def _project_to_host(ct, phantom, specs):
    """get_sino / get_sinos for LARGE scans: the projection runs in chunks of views and the float64 copy of every finished
    chunk travels to page-locked host memory on a side stream while the next chunk is being traced -- the reference-shaped
    API hands back 8 bytes per ray, sinogram and spectrum (17.7 GB for configuration C5), which costs more time on the PCIe
    link than the kernels take.  Same arithmetic, same results as the one-shot path (a view range on its own equals the same
    views inside the whole scan, tests/test_gpu_siddon.py).  Returns None when the call should take the one-shot path:
    small results, view shards that are gathered, counting noise, several energy grids."""
    code = dtype_code(config.dtype)
    rank, ws = dist.world() if config.shard_views else (0, 1)
    if ws > 1 and config.gather_views:
        return None
    if noise_enabled(ct):
        return None
    if not all(len(s.E) == len(specs[0].E) and np.array_equal(s.E, specs[0].E) for s in specs) or len(specs) > 4:
        return None
    v0, nv = dist.shard_range(ct.N_proj, rank, ws)
    per_view = ct.N_rows * ct.N_channels
    total = 2 * len(specs) * nv * per_view * 8
    if total < PIPELINE_MIN_BYTES or nv < 2:
        return None
    st = _get_state(ct, phantom)
    keep_L = config.cache_lengths and len(specs) == 1
    have_L = config.cache_lengths and st.L is not None and st.L_dtype == code and st.L.shape[0] == nv   # a second get_sino call
    E = specs[0].E
    mu = phantom.mu_table(E)
    W = np.stack([spectral_weights(ct, sp) for sp in specs])
    nS = len(specs)
    n_chunks = int(min(16, max(2, -(-total // (2 << 30)))))
    step = -(-nv // n_chunks)
    shape = (nS, nv, ct.N_rows, ct.N_channels)
    h_c = torch.empty(shape, dtype=torch.float64, pin_memory=True)
    h_l = torch.empty(shape, dtype=torch.float64, pin_memory=True)
    dev = st.projector.device
    td = torch.float32 if code == 0 else torch.float64
    L_full = torch.empty((nv, ct.N_rows, ct.N_channels, mu.shape[0]), dtype=td, device=dev) if (keep_L and not have_L) else None
    side = torch.cuda.Stream(device=dev)
    with torch.cuda.device(dev):
        for a in range(0, nv, step):
            n = min(step, nv - a)
            if have_L:                                   # detection alone, from the path lengths the first call stored
                c, lg = poly_detect(st.L[a:a + n], mu, W, dtype=code, want_log=True)
            else:
                c, lg, _ = st.projector.project(mu, W, view_begin=v0 + a, view_count=n, dtype=code, want_log=True,
                                                want_L=keep_L, L=None if L_full is None else L_full[a:a + n])
            done = torch.cuda.Event()
            done.record()
            with torch.cuda.stream(side):
                side.wait_event(done)
                for sp in range(nS):                     # conversion to float64 on the device, then one DMA per block
                    h_c[sp, a:a + n].copy_(c[sp], non_blocking=True)
                    h_l[sp, a:a + n].copy_(lg[sp], non_blocking=True)
            c.record_stream(side)
            lg.record_stream(side)
        side.synchronize()
    if L_full is not None:
        st.L, st.L_dtype = L_full, code
    raw, log = h_c.numpy(), h_l.numpy()
    if ct.N_rows == 1:
        raw, log = raw[..., 0, :], log[..., 0, :]
    return [(raw[sp], log[sp]) for sp in range(nS)]


def get_sino(ct, phantom, spec):
    """Forward-project ``phantom`` with scanner ``ct`` and spectrum ``spec``.

    Returns ``(sino_raw, sino_log)``: detected counts and ``-ln(counts / unattenuated counts)``, each a float64 numpy
    array ``[N_proj, N_channels]`` (``[N_proj, N_rows, N_channels]`` for multi-row scanners).
    """
    piped = _project_to_host(ct, phantom, [spec])
    if piped is not None:
        return piped[0]
    (c, lg), = _project_groups(ct, phantom, [spec])[0]
    raw, log = _to_host([c[0], lg[0]], ct)
    return raw, log


def get_sinos(ct, phantom, specs):
    """Extension: all spectra of a multi-energy acquisition in one traversal (one fused kernel when they share an
    energy grid).  Returns a list of ``(sino_raw, sino_log)``."""
    piped = _project_to_host(ct, phantom, list(specs))
    if piped is not None:
        return piped
    groups = _project_groups(ct, phantom, list(specs))[0]
    flat = _to_host([t for c, lg in groups for t in (c, lg)], ct)      # [n_in_group, views, (rows,) ch] each
    out = []
    for g in range(len(groups)):
        raw, log = flat[2 * g], flat[2 * g + 1]
        out.extend((raw[n], log[n]) for n in range(raw.shape[0]))
    return out
