#!/usr/bin/env python3
"""Profiling harness for the 3-D projector: a few launches of `siddon_skip3_kernel` on one view range of C4 / C5 (for `ncu`;
the time it prints is a plain CUDA-event time of the same launches, a tuning aid and not a bench value).
usage: prof_sid3.py [C4|C5] [first view] [views] [reps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'dex-ct-sim_b200')):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from dexct import _native  # noqa: E402
if os.environ.get('DEXCT_LIB_VARIANT'):          # a tuning build (make -C dex-ct-sim_b200/csrc variant_sid VARFLAGS=...)
    _native.LIB_PATH = os.path.join(ROOT, 'dex-ct-sim_b200', 'csrc', '_build', os.environ['DEXCT_LIB_VARIANT'])
from dexct import synthetic as S  # noqa: E402
from dexct.ops import Projector, count_segments, make_scan  # noqa: E402
from xtomosim import forward_project as FP  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else 'C4'
    v0 = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    nv = int(sys.argv[3]) if len(sys.argv) > 3 else 48
    reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
    dev = torch.device('cuda', 0)
    ct, ph, lo, hi = S.make_config(name, device=dev)
    scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, n_rows=ct.N_rows, h_iso=ct.h_iso, rotation=ct.rotation)
    pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz), device=dev)
    mu = torch.as_tensor(ph.mu_table(lo.E), dtype=torch.float32, device=dev)
    W = torch.as_tensor(np.stack([FP.spectral_weights(ct, lo), FP.spectral_weights(ct, hi)]), dtype=torch.float32, device=dev)
    counts = torch.empty((2, nv, ct.N_rows, ct.N_channels), dtype=torch.float32, device=dev)
    logs = torch.empty_like(counts)
    pj.project(mu, W, view_begin=v0, view_count=nv, counts=counts, logs=logs)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        pj.project(mu, W, view_begin=v0, view_count=nv, counts=counts, logs=logs)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    rays = nv * ct.N_rows * ct.N_channels
    seg = count_segments(pj, v0, min(nv, 4)) / (min(nv, 4) * ct.N_rows * ct.N_channels)
    # steps of the production walk (homogeneous stretches and ordinary columns) on one view of the range
    ids = np.arange(v0 * pj.rays_per_view, (v0 + 1) * pj.rays_per_view, 7, dtype=np.int64)
    recs, _ = pj.walk_trace(ids, 8, dtype='f32', steps=False)
    steps = float(np.mean([r['n_steps'] for r in recs]))
    print(f'{name} view {v0}: {steps:.1f} walk steps per ray (sample of {len(ids)} rays)')
    print(f'{name} views [{v0}, {v0 + nv}): {ms:.4f} ms per launch, {rays / ms / 1e6:.3f} Grays/s, {seg:.1f} segments per ray, '
          f'{rays * (seg + 8) / ms / 1e6:.1f} GB/s algorithmic, checksum {float(counts.sum()):.6e}')


if __name__ == '__main__':
    main()
