"""Peer-memory groups for view-sharded scans on one NVLink/NVSwitch node (csrc/peer.cu, include/dexct.h "peer memory").

One process per GPU (torch.distributed).  ``PeerGroup`` allocates *symmetric* device buffers -- the same size on every
rank, each mapped into every other rank's address space through CUDA IPC -- so that kernels can store their results into
all ranks' copies directly: the decomposition kernel's epilogue (``gn_decompose(..., mirrors=...)``) and ``push`` (a plain
store kernel for a finished buffer).  That replaces the all-gathers of SURVEY.md section 8e for steps that are too short
to hide NCCL's fixed per-call latency; ``barrier`` and ``max_key`` are the only synchronisation.  torch.distributed is
used once per buffer, to exchange the 64-byte handles.
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.distributed as dist

from . import _native as N
from . import dist as ddist


class _RawCuda:
    """Exposes a raw device allocation to torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {'shape': (int(nbytes),), 'typestr': '|u1', 'data': (int(ptr), False), 'version': 2}


class SymBuffer:
    """One symmetric allocation: ``local`` is this rank's copy as a uint8 tensor, ``ptrs[r]`` the address of rank r's copy
    in THIS process (``ptrs[rank]`` is the local one)."""

    def __init__(self, group, nbytes):
        """Collective over the group.  A failure on ANY rank -- allocation, export, or mapping a peer's handle -- raises
        DexctError on EVERY rank, after everything that did succeed was undone: the callers fall back together (bench.py:
        NCCL all-gathers) instead of one rank waiting for the others in the next collective."""
        self.group = group
        self.nbytes = int(nbytes)
        self._own = None
        self._opened = []
        self.local = None
        ws = group.world_size
        err, mine = None, None
        try:
            if os.environ.get('DEXCT_PEER_FAIL') == f'alloc:{group.rank}':         # fault injection (tests/peer_worker.py)
                raise N.DexctError('injected allocation failure')
            p = C.c_void_p()
            N.check(N.lib().dexct_peer_alloc(self.nbytes, C.byref(p)))
            self._own = p.value
            handle = (C.c_ubyte * 64)()
            N.check(N.lib().dexct_peer_export(C.c_void_p(self._own), handle))
            mine = bytes(handle)
        except N.DexctError as exc:
            err = exc
        handles = [mine]
        if ws > 1:
            handles = [None] * ws
            dist.all_gather_object(handles, mine, group=group.group)
        failed = [r for r, h in enumerate(handles) if h is None]
        if failed:
            self._undo(collective=False)                   # nobody has mapped anything yet
            raise N.DexctError(f'peer allocation of {self.nbytes} bytes failed on rank(s) {failed}'
                               + (f': {err}' if err is not None else ''))
        self.ptrs = []
        try:
            for r in range(ws):
                if r == group.rank:
                    self.ptrs.append(self._own)
                    continue
                if os.environ.get('DEXCT_PEER_FAIL') == f'open:{group.rank}':
                    raise N.DexctError('injected mapping failure')
                q = C.c_void_p()
                buf = (C.c_ubyte * 64).from_buffer_copy(handles[r])
                N.check(N.lib().dexct_peer_open(buf, C.byref(q)))
                self._opened.append(q.value)
                self.ptrs.append(q.value)
        except N.DexctError as exc:
            err = exc
        if ws > 1 and not ddist.all_agree(err is None, group=group.group):
            self._undo(collective=True)
            raise N.DexctError('peer memory could not be mapped on every rank' + (f': {err}' if err is not None else ''))
        self._raw = _RawCuda(self._own, self.nbytes)
        self.local = torch.as_tensor(self._raw, device=group.device)

    def _undo(self, collective):
        for q in self._opened:
            N.lib().dexct_peer_close(C.c_void_p(q))
        self._opened = []
        if collective:
            dist.barrier(group=self.group.group)           # nobody frees memory a peer still has mapped
        if self._own is not None:
            N.lib().dexct_peer_free(C.c_void_p(self._own))
            self._own = None

    def view(self, dtype, shape):
        """This rank's copy as a tensor of ``dtype`` and ``shape`` (must fill the buffer exactly or less)."""
        n = 1
        for s in shape:
            n *= int(s)
        nb = n * torch.empty((), dtype=dtype).element_size()
        if nb > self.nbytes:
            raise ValueError('view larger than the buffer')
        return self.local[:nb].view(dtype).view(*shape)

    def table(self, offset_bytes=0, exclude_self=False):
        """ctypes array of the copies' addresses (+ offset), in rank order."""
        ps = [p + int(offset_bytes) for r, p in enumerate(self.ptrs) if not (exclude_self and r == self.group.rank)]
        return (C.c_void_p * len(ps))(*ps)

    def close(self):
        if self._own is None:
            return
        torch.cuda.synchronize(self.group.device)
        self.local = None
        for q in self._opened:
            N.lib().dexct_peer_close(C.c_void_p(q))
        self._opened = []
        if self.group.world_size > 1:
            dist.barrier(group=self.group.group)           # nobody frees memory a peer still has mapped
        N.lib().dexct_peer_free(C.c_void_p(self._own))
        self._own = None


class PeerGroup:
    """All ranks of ``group`` (default: the world) must sit on one node with peer access between their GPUs."""

    def __init__(self, group=None, device=None, timeout_s=20.0):
        N.require_device()
        self.group = group
        if dist.is_available() and dist.is_initialized():
            self.rank, self.world_size = dist.get_rank(group), dist.get_world_size(group)
        else:
            self.rank, self.world_size = 0, 1
        if self.world_size > 16:
            raise ValueError('at most DEXCT_MAX_PEERS = 16 ranks')
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        self.timeout_s = float(timeout_s)
        self._bar_epoch = 0
        self._max_epoch = 0
        self.buffers = []
        with torch.cuda.device(self.device):
            self.ctrl = SymBuffer(self, N.lib().dexct_peer_ctrl_bytes())
        self._ctrl_table = self.ctrl.table()

    def alloc(self, nbytes):
        with torch.cuda.device(self.device):
            b = SymBuffer(self, (int(nbytes) + 255) // 256 * 256)
        self.buffers.append(b)
        return b

    def barrier(self):
        """Enqueue on the current stream: returns (on the device) once every rank has reached this point and all stores
        the ranks issued before it -- into anyone's copy -- are visible."""
        self._bar_epoch += 1
        with torch.cuda.device(self.device):
            N.check(N.lib().dexct_peer_barrier(self._ctrl_table, self.rank, self.world_size, self._bar_epoch, self.timeout_s,
                                               N.stream_ptr()))

    def max_key(self, scratch, code):
        """Integer MAX all-reduce of the air-maximum key at the head of a gn_decompose scratch buffer (in place)."""
        self._max_epoch += 1
        with torch.cuda.device(self.device):
            N.check(N.lib().dexct_peer_max_key(self._ctrl_table, self.rank, self.world_size, self._max_epoch, N.ptr(scratch),
                                               int(code), self.timeout_s, N.stream_ptr()))

    def push(self, src, buf, offset_bytes, max_ctas=0, dma=False):
        """Store the contiguous tensor ``src`` at ``offset_bytes`` of every OTHER rank's copy of ``buf`` -- with a store
        kernel on at most ``max_ctas`` CTAs (0: one per SM), or, ``dma=True``, on the copy engines (no SM time: for a copy
        that runs beside a kernel occupying every SM)."""
        if self.world_size == 1:
            return
        if not src.is_contiguous():
            raise ValueError('src must be contiguous')
        nb = src.numel() * src.element_size()
        if offset_bytes % 16 or nb % 16 or offset_bytes + nb > buf.nbytes:
            raise ValueError('push needs 16-byte aligned offset and size inside the buffer')
        tab = buf.table(offset_bytes, exclude_self=True)
        with torch.cuda.device(self.device):
            if dma:
                N.check(N.lib().dexct_peer_push_dma(N.ptr(src), nb, tab, len(tab), -1, N.stream_ptr()))
            else:
                N.check(N.lib().dexct_peer_push(N.ptr(src), nb, tab, len(tab), -1, int(max_ctas), N.stream_ptr()))

    def status(self):
        """0 when every wait so far was answered in time, else 1 + the rank that never arrived (synchronises)."""
        s = C.c_int32(0)
        with torch.cuda.device(self.device):
            N.check(N.lib().dexct_peer_status(C.c_void_p(self.ctrl.ptrs[self.rank]), C.byref(s), N.stream_ptr()))
        return s.value

    def close(self):
        for b in self.buffers:
            b.close()
        self.buffers = []
        self.ctrl.close()
