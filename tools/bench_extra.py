#!/usr/bin/env python3
"""Timings of the pieces bench.py's headline does not cover: workspace preparation, the 3-D (multi-row) projector on a
C4 view range, and fan-beam FBP at C2 size.  Tuning aid; prints plain lines."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'dex-ct-sim_b200')):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from dexct import synthetic as S  # noqa: E402
from dexct.ops import Projector, count_segments, fbp_fan, make_scan  # noqa: E402
from xtomosim import forward_project as FP  # noqa: E402


def timeit(fn, reps=10, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    which = sys.argv[1:] or ['prep', 'fbp', 'c4']
    dev = torch.device('cuda', 0)
    if 'prep' in which or 'fbp' in which:
        ct, ph, lo, hi = S.make_config('C2')
        scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, rotation=ct.rotation)
        pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz), device=dev)
        if 'prep' in which:
            print('C2 prepare (tables + pair layouts + skip layouts): %.4f ms' % timeit(pj.prepare))
        if 'fbp' in which:
            mu = torch.as_tensor(ph.mu_table(lo.E), dtype=torch.float32, device=dev)
            W = torch.as_tensor(np.stack([FP.spectral_weights(ct, lo)]), dtype=torch.float32, device=dev)
            _, logs, _ = pj.project(mu, W)
            sino = logs[0, :, 0, :].contiguous()
            for dt in ('f32', 'f64'):
                t = timeit(lambda: fbp_fan(scan, sino, 512, 51.2, 0.8, dtype=dt), reps=5)
                print('C2 FBP 1000 views x 900 ch -> 512^2 (%s): %.4f ms' % (dt, t))
    if 'c4' in which:
        n_views = int(os.environ.get('C4_VIEWS', '48'))
        t0 = time.time()
        ct, ph, lo, hi = S.make_config('C4', device=dev)
        print('C4 phantom built in %.1f s' % (time.time() - t0))
        scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, n_rows=ct.N_rows, h_iso=ct.h_iso, rotation=ct.rotation)
        pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz), device=dev)
        print('C4 prepare: %.3f ms' % timeit(pj.prepare, reps=3, warm=1))
        mu = torch.as_tensor(ph.mu_table(lo.E), dtype=torch.float32, device=dev)
        W = torch.as_tensor(np.stack([FP.spectral_weights(ct, lo), FP.spectral_weights(ct, hi)]), dtype=torch.float32, device=dev)
        counts = torch.empty((2, n_views, ct.N_rows, ct.N_channels), dtype=torch.float32, device=dev)
        logs = torch.empty_like(counts)
        for v0 in (0, 150):       # 0: axis-aligned views, 150: 45 degrees
            t = timeit(lambda: pj.project(mu, W, view_begin=v0, view_count=n_views, counts=counts, logs=logs), reps=3, warm=1)
            rays = n_views * ct.N_rows * ct.N_channels
            seg = count_segments(pj, v0, 4) / (4 * ct.N_rows * ct.N_channels)
            bpr = seg + 8
            print('C4 views [%d, %d): %.3f ms for %d rays -> %.3e rays/s, %.1f seg/ray, %.1f GB/s algorithmic (frac %.3f of 6452.8)'
                  % (v0, v0 + n_views, t, rays, rays / t * 1e3, seg, rays * bpr / t * 1e3 / 1e9, rays * bpr / t * 1e3 / 1e9 / 6452.8))


if __name__ == '__main__':
    main()
