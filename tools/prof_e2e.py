#!/usr/bin/env python3
"""Host-side breakdown of the end-to-end step bench.py times (`e2e`): cProfile of get_sinos + get_basismat_sinos on C2
with host buffers in and out.  Prints a profile, not a bench value."""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'dex-ct-sim_b200')):
    sys.path.insert(0, p)
import torch  # noqa: E402

from dexct import synthetic as S  # noqa: E402
from xtomosim import forward_project as FP  # noqa: E402
import matdecomp  # noqa: E402


def main():
    ct, ph, lo, hi = S.make_config('C2')

    def step():
        FP.clear_cache()
        (raw_lo, log_lo), (raw_hi, log_hi) = FP.get_sinos(ct, ph, [lo, hi])
        m1, m2 = matdecomp.get_basismat_sinos(ct, raw_lo, raw_hi, lo, hi, n_iters=10, mask_thresh=0.95)
        return m1

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        step()
    torch.cuda.synchronize()
    print('ms per e2e step: %.3f' % ((time.perf_counter() - t0) * 100))
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(10):
        step()
    torch.cuda.synchronize()
    pr.disable()
    pstats.Stats(pr).sort_stats('cumulative').print_stats(45)


if __name__ == '__main__':
    main()
