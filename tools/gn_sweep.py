#!/usr/bin/env python3
"""Time the GN kernel on the C2 pixel set for the tuning knobs read from the environment (DEXCT_GN_PX, ...).
Prints timings for tuning, not bench values."""
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
    dev = torch.device('cuda', 0)
    ct, ph, lo, hi = S.make_config('C2')
    scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, rotation=ct.rotation)
    pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz), device=dev)
    mu = torch.as_tensor(ph.mu_table(lo.E), dtype=torch.float32, device=dev)
    W = torch.as_tensor(np.stack([FP.spectral_weights(ct, lo), FP.spectral_weights(ct, hi)]), dtype=torch.float32, device=dev)
    _, i0, mus = matdecomp.spectral_tables(ct, lo, hi)
    i0 = torch.as_tensor(i0, dtype=torch.float32, device=dev)
    mus = torch.as_tensor(mus, dtype=torch.float32, device=dev)
    counts, _, _ = pj.project(mu, W)
    g = counts.view(2, -1)
    basis = torch.empty((g.shape[1], 2), dtype=torch.float32, device=dev)
    ref = None
    knobs = [dict(DEXCT_GN_PX=str(px)) for px in (4,)] + [dict(a.split('=') for a in k.split(',')) for k in sys.argv[1:]]
    for kn in knobs:
        for k in list(os.environ):
            if k.startswith('DEXCT_GN_'):
                del os.environ[k]
        os.environ.update(kn)
        for _ in range(3):
            gn_decompose(g, i0, mus, 10, mask_thresh=0.95, out=basis)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            gn_decompose(g, i0, mus, 10, mask_thresh=0.95, out=basis)
        e1.record()
        torch.cuda.synchronize()
        if ref is None:
            ref = basis.clone()
        err = float((basis - ref).abs().max() / ref.abs().max())
        nfl = int(gn_decompose.last_scratch[8:12].view(torch.int32)[0])
        print('flagged for the float64 pass:', nfl, 'of', g.shape[1])
        print(kn, 'ms per call (incl. max + table kernels): %.4f' % (e0.elapsed_time(e1) / 20), 'max diff vs first: %.2e' % err)


if __name__ == '__main__':
    main()
