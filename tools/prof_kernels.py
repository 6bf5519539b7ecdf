#!/usr/bin/env python3
"""Profiling harness: a few launches of the hot kernels on config C2 (for `ncu`; prints nothing that is a bench value)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'dex-ct-sim_b200')):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from dexct import synthetic as S  # noqa: E402
from dexct.ops import Projector, gn_decompose, make_scan  # noqa: E402
from xtomosim import forward_project as FP  # noqa: E402
import matdecomp  # noqa: E402


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else 'all'
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    dev = torch.device('cuda', 0)
    ct, ph, lo, hi = S.make_config('C2')
    scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, rotation=ct.rotation)
    pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz), device=dev)
    mu = torch.as_tensor(ph.mu_table(lo.E), dtype=torch.float32, device=dev)
    W = torch.as_tensor(np.stack([FP.spectral_weights(ct, lo), FP.spectral_weights(ct, hi)]), dtype=torch.float32, device=dev)
    _, i0, mus = matdecomp.spectral_tables(ct, lo, hi)
    counts = torch.empty((2, ct.N_proj, 1, ct.N_channels), dtype=torch.float32, device=dev)
    logs = torch.empty_like(counts)
    basis = torch.empty((ct.N_proj * ct.N_channels, 2), dtype=torch.float32, device=dev)
    for _ in range(reps):
        if which in ('all', 'siddon'):
            pj.project(mu, W, counts=counts, logs=logs)
        else:
            pj.project(mu, W, counts=counts, logs=logs) if _ == 0 else None
        if which in ('all', 'gn'):
            gn_decompose(counts.view(2, -1), i0, mus, 10, mask_thresh=0.95, out=basis)
    torch.cuda.synchronize()
    print('ok', float(counts.sum()), float(basis.abs().sum()))


if __name__ == '__main__':
    main()
