#!/usr/bin/env python3
"""Wall time of the two numpy entry points on a full-size configuration (C5 by default), call by call, and the bare PCIe
transfers they contain.  usage: prof_e2e_c5.py [C5|C4]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'dex-ct-sim_b200')):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from dexct import synthetic as S  # noqa: E402
from xtomosim import forward_project as FP  # noqa: E402
import matdecomp  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'C5'
dev = torch.device('cuda', 0)
ct, ph, lo, hi = S.make_config(name, device=dev)
host = torch.empty(ph.M.shape, dtype=torch.uint8, pin_memory=True)
host.copy_(ph.M)
torch.cuda.synchronize()
ph.M = host.numpy()


def t(f):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = f()
    torch.cuda.synchronize()
    return r, 1e3 * (time.perf_counter() - t0)


for rep in range(3):
    FP.clear_cache()
    _, t_sig = t(lambda: FP._signature(ct, ph))
    st, t_state = t(lambda: FP._get_state(ct, ph))
    FP.clear_cache()
    sin, t_sinos = t(lambda: FP.get_sinos(ct, ph, [lo, hi]))
    (raw_lo, log_lo), (raw_hi, log_hi) = sin
    bas, t_gn = t(lambda: matdecomp.get_basismat_sinos(ct, raw_lo, raw_hi, lo, hi, n_iters=10, mask_thresh=0.95))
    print(f'rep {rep}: signature {t_sig:.1f} ms, state (signature + upload + layouts) {t_state:.1f} ms, get_sinos {t_sinos:.1f} ms, '
          f'get_basismat_sinos {t_gn:.1f} ms', flush=True)
    del sin, bas, raw_lo, raw_hi, log_lo, log_hi
# bare transfers
n = 1 << 30
h = torch.empty(n, dtype=torch.float64, pin_memory=True)
d = torch.empty(n, dtype=torch.float64, device=dev)
for _ in range(2):
    _, t_up = t(lambda: d.copy_(h, non_blocking=True))
    _, t_dn = t(lambda: h.copy_(d, non_blocking=True))
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
d2 = torch.empty_like(d)
h2 = torch.empty(n, dtype=torch.float64, pin_memory=True)


def both():
    with torch.cuda.stream(s1):
        d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)


_, t_both = t(both)
print(f'8 GiB pinned: up {8.59 / t_up * 1e3:.1f} GB/s, down {8.59 / t_dn * 1e3:.1f} GB/s, both at once {2 * 8.59 / t_both * 1e3:.1f} GB/s in total')
_, t_alloc = t(lambda: torch.empty(n, dtype=torch.float64, pin_memory=True))
print(f'page-locked allocation of 8 GiB (caching host allocator, after the first): {t_alloc:.1f} ms')
