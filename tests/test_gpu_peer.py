"""Peer-memory path of view-sharded scans (csrc/peer.cu, dexct/peer.py): symmetric buffers mapped through CUDA IPC, store
kernels in place of all-gathers, flag barriers and the integer MAX exchange.  Multi-rank cases run as subprocesses -- on
distinct GPUs when the box has them, otherwise sharing cuda:0."""
import ctypes as C
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch

from dexct import synthetic as S
from dexct.ops import F32, F64, gn_decompose
from dexct.peer import PeerGroup

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_single_rank_group_and_mirrored_decomposition():
    pg = PeerGroup()
    assert (pg.rank, pg.world_size) == (0, 1)
    buf = pg.alloc(1000)
    assert buf.nbytes == 1024 and buf.ptrs == [buf.local.data_ptr()] and int(buf.local.sum()) == 0
    v = buf.view(torch.float32, (4, 8))
    v.fill_(2.0)
    assert float(buf.view(torch.float32, (32,)).sum()) == 64.0
    for _ in range(3):
        pg.barrier()
    g, i0, mus, _ = S.gn_problem(6, 50, 30)
    for code, td in ((F32, torch.float32), (F64, torch.float64)):
        gd = torch.as_tensor(g, dtype=td, device='cuda')
        scratch_key = torch.zeros(16, dtype=torch.uint8, device='cuda')
        pg.max_key(scratch_key, code)                      # a world of one: the key comes back unchanged
        assert int(scratch_key.sum()) == 0
        plain = gn_decompose(gd, i0, mus, 15, mask_thresh=0.95, dtype=code)
        m1, m2 = torch.full_like(plain, -1.0), torch.full_like(plain, -1.0)
        tab = (C.c_void_p * 2)(m1.data_ptr(), m2.data_ptr())
        again = gn_decompose(gd, i0, mus, 15, mask_thresh=0.95, dtype=code, mirrors=tab)
        assert torch.equal(again, plain) and torch.equal(m1, plain) and torch.equal(m2, plain)
        # the same kernel can carry a finished buffer along (here: its own input) to further addresses
        for iters in (15, 1, 0):
            d1, d2, m3 = torch.zeros_like(gd), torch.zeros_like(gd), torch.zeros_like(plain)
            out = gn_decompose(gd, i0, mus, iters, mask_thresh=0.95, dtype=code, mirrors=(C.c_void_p * 1)(m3.data_ptr()),
                               push=(gd, (C.c_void_p * 2)(d1.data_ptr(), d2.data_ptr())))
            assert torch.equal(d1, gd) and torch.equal(d2, gd) and torch.equal(m3, out)
            if iters == 15:
                assert torch.equal(out, plain)
    # more pixels than one resident wave covers (a CTA then owns two groups, some only one): every slice must still go out
    gb, i0b, musb, _ = S.gn_problem(700, 700, 30)
    gd = torch.as_tensor(gb, dtype=torch.float32, device='cuda')
    d1 = torch.zeros_like(gd)
    out = gn_decompose(gd, i0b, musb, 3, dtype=F32, push=(gd, (C.c_void_p * 1)(d1.data_ptr())))
    assert torch.equal(d1, gd) and torch.equal(out, gn_decompose(gd, i0b, musb, 3, dtype=F32))
    assert pg.status() == 0
    pg.close()


def test_starved_pixels_reach_the_mirrors_too():
    # pixels redone by the float64 pass (photon-starved, built as in test_gpu_gn.py) must be rewritten in every copy
    E = np.arange(1.0, 141.0)
    lo, hi = S.gaussian_spectrum('lo', 55.0, 18.0, E, 1e8), S.gaussian_spectrum('hi', 95.0, 25.0, E, 1e8)
    i0 = np.stack([lo.I0, hi.I0])
    mus = np.stack([0.2 * (60.0 / E) ** 2.2 + 0.15, 0.9 * (60.0 / E) ** 2.9 + 0.12])
    r = np.random.default_rng(9)
    a = np.stack([r.uniform(0, 25, 400), r.uniform(0, 4, 400)], axis=-1)
    att = np.exp(-(a[:, 0:1] * mus[0] + a[:, 1:2] * mus[1]))
    g = np.stack([att @ i0[0], att @ i0[1]])
    g[1, :40] *= r.uniform(1e-9, 1e-3, 40)
    gd = torch.as_tensor(g, dtype=torch.float32, device='cuda')
    plain = gn_decompose(gd, i0, mus, 30, dtype=F32)
    redone = int(gn_decompose.last_scratch[8:12].view(torch.int32)[0])
    assert redone > 0
    m = torch.zeros_like(plain)
    gn_decompose(gd, i0, mus, 30, dtype=F32, mirrors=(C.c_void_p * 1)(m.data_ptr()))
    assert torch.equal(torch.nan_to_num(m, nan=-7.0), torch.nan_to_num(plain, nan=-7.0))


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


@pytest.mark.parametrize('ws', [2, 3])
def test_ranks_store_into_each_others_copies(ws):
    port = _free_port()
    procs = []
    for r in range(ws):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(ws), MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, 'peer_worker.py')], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=240)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(out)
    for r, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and f'peer worker {r}/{ws} ok' in out, out[-3000:]
