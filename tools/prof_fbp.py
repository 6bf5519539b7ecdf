#!/usr/bin/env python3
"""Per-kernel device times of one multi-row FBP (torch profiler, CUDA activities): filter versus back-projection.
usage: prof_fbp.py [n_rows] [n_ch] [n_views] [n_matrix]    (defaults: the C5 detector at a quarter of its views)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'dex-ct-sim_b200')):
    sys.path.insert(0, p)
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

from dexct.ops import fbp_fan, make_scan  # noqa: E402


def main():
    a = [int(x) for x in sys.argv[1:]]
    n_rows, n_ch, n_views, n = (a + [128, 1800, 600, 1024][len(a):])[:4]
    scan = make_scan(n_views, n_ch, 0.8230337, 60.0, 100.0, n_rows=n_rows, h_iso=0.05)
    sino = torch.rand((n_views, n_rows, n_ch), device='cuda')
    img = torch.empty((n_rows, n, n), device='cuda')
    fbp_fan(scan, sino, n, 51.2, 0.8, image=img)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(2):
            fbp_fan(scan, sino, n, 51.2, 0.8, image=img)
        torch.cuda.synchronize()
    print(f'rows {n_rows} ch {n_ch} views {n_views} -> {n}^2, DEXCT_FBP_RB={os.environ.get("DEXCT_FBP_RB", "auto")}')
    for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:6]:
        print(f'  {e.device_time_total / e.count / 1e3:10.3f} ms x {e.count}  {e.key[:110]}')


if __name__ == '__main__':
    main()
