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

    # the two entry points on their own, and the bare transfers they contain (page-locked float64 arrays, as returned)
    def wall(f, n=20):
        f()
        torch.cuda.synchronize()
        t = time.perf_counter()
        for _ in range(n):
            f()
        torch.cuda.synchronize()
        return (time.perf_counter() - t) / n * 1e3

    (raw_lo, log_lo), (raw_hi, log_hi) = FP.get_sinos(ct, ph, [lo, hi])
    print('get_sinos (phantom upload + prepare + kernel + 4 arrays down)  %.3f ms' % wall(
        lambda: (FP.clear_cache(), FP.get_sinos(ct, ph, [lo, hi]))))
    print('get_sinos, phantom layouts kept                                 %.3f ms' % wall(lambda: FP.get_sinos(ct, ph, [lo, hi])))
    print('get_basismat_sinos (2 arrays up, kernel, 1 array down)          %.3f ms' % wall(
        lambda: matdecomp.get_basismat_sinos(ct, raw_lo, raw_hi, lo, hi, n_iters=10, mask_thresh=0.95)))
    d32 = torch.rand(2, 1000, 1, 900, device='cuda')
    d64 = d32.double()
    h64 = torch.empty(d32.shape, dtype=torch.float64, pin_memory=True)

    def down(src):
        h64.copy_(src, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def down_alloc(src):
        h = torch.empty(src.shape, dtype=torch.float64, pin_memory=True)
        h.copy_(src, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def up():
        d32.copy_(h64, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    mb = h64.numel() * 8 / 1e6
    for name, f in [('down f64 -> pinned f64 (pure DMA)', lambda: down(d64)), ('down f32 -> pinned f64 (convert + DMA)', lambda: down(d32)),
                    ('same + pinned allocation per call', lambda: down_alloc(d32)), ('up pinned f64 -> f32 (DMA + convert)', up)]:
        t = wall(f)
        print('%-45s %.3f ms  (%.1f MB, %.1f GB/s)' % (name, t, mb, mb / t))
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(10):
        step()
    torch.cuda.synchronize()
    pr.disable()
    pstats.Stats(pr).sort_stats('cumulative').print_stats(45)


if __name__ == '__main__':
    main()
