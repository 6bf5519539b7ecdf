#!/usr/bin/env python3
"""Host side of the end-to-end step under view sharding: what every rank spends in the numpy entry points on its own view
range of C5, and what the box's host memory / PCIe roots give when ALL ranks copy at once (the ceiling of the e2e leg).
usage: python -m torch.distributed.run --nproc-per-node N tools/prof_e2e_ranks.py [C5|C4]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'dex-ct-sim_b200')):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from dexct import config as cfg  # noqa: E402
from dexct import synthetic as S  # noqa: E402
from xtomosim import forward_project as FP  # noqa: E402
import matdecomp  # noqa: E402

rank, ws, lr = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('LOCAL_RANK', 0))
torch.cuda.set_device(lr)
dev = torch.device('cuda', lr)
os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
if ws > 1:
    dist.init_process_group('nccl', device_id=dev)
name = sys.argv[1] if len(sys.argv) > 1 else 'C5'


def sync():
    torch.cuda.synchronize()
    if ws > 1:
        dist.barrier()
    torch.cuda.synchronize()


def gather(x):
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    out = [torch.zeros_like(t) for _ in range(ws)]
    if ws > 1:
        dist.all_gather(out, t)
    else:
        out = [t]
    return [float(o) for o in out]


# ---- bare copies, all ranks at once
n = 1 << 28                                                               # 2 GiB of float64
h = torch.empty(n, dtype=torch.float64, pin_memory=True)
d = torch.empty(n, dtype=torch.float64, device=dev)
res = {}
for label, fn in (('down', lambda: h.copy_(d, non_blocking=True)), ('up', lambda: d.copy_(h, non_blocking=True))):
    fn()
    sync()
    t0 = time.perf_counter()
    fn()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    res[label] = gather(n * 8 / dt / 1e9)
    sync()
if rank == 0:
    for k, v in res.items():
        print(f'all {ws} ranks copying 2 GiB {k} at once: per rank {min(v):.1f} .. {max(v):.1f} GB/s, sum {sum(v):.1f} GB/s', flush=True)
del h, d

# ---- the entry points on this rank's view range
cfg.dtype = 'f32'
cfg.gather_views = ws == 1
ct, ph, lo, hi = S.make_config(name, device=dev)
host = torch.empty(ph.M.shape, dtype=torch.uint8, pin_memory=True)
host.copy_(ph.M)
torch.cuda.synchronize()
ph.M = host.numpy()
for rep in range(3):
    FP.clear_cache()
    sync()
    t0 = time.perf_counter()
    FP._get_state(ct, ph)
    torch.cuda.synchronize()
    t_state = 1e3 * (time.perf_counter() - t0)
    FP.clear_cache()
    sync()
    t0 = time.perf_counter()
    (raw_lo, log_lo), (raw_hi, log_hi) = FP.get_sinos(ct, ph, [lo, hi])
    torch.cuda.synchronize()
    t_sin = 1e3 * (time.perf_counter() - t0)
    sync()
    t0 = time.perf_counter()
    m1, m2 = matdecomp.get_basismat_sinos(ct, raw_lo, raw_hi, lo, hi, n_iters=10, mask_thresh=0.95)
    torch.cuda.synchronize()
    t_gn = 1e3 * (time.perf_counter() - t0)
    a, b, c = gather(t_state), gather(t_sin), gather(t_gn)
    if rank == 0 and rep > 0:
        print(f'rep {rep}: workspace (upload + digest + layouts of the whole phantom, every rank) {min(a):.0f} .. {max(a):.0f} ms; '
              f'get_sinos on {raw_lo.shape[0]} views {min(b):.0f} .. {max(b):.0f} ms; get_basismat_sinos {min(c):.0f} .. {max(c):.0f} ms', flush=True)
    del raw_lo, raw_hi, log_lo, log_hi, m1, m2
if ws > 1:
    dist.destroy_process_group()
