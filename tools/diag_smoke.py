#!/usr/bin/env python3
"""Diagnostic: where does the single-precision decomposition of smoke()'s scan differ most from the float64 oracle?"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'dex-ct-sim_b200')):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import matdecomp  # noqa: E402
from dexct import synthetic as S  # noqa: E402
from dexct.ops import F32, F64, gn_decompose  # noqa: E402
from oracle import gn_oracle  # noqa: E402
from xtomosim.forward_project import get_sinos  # noqa: E402

ct, ph, lo, hi = S.make_config('C2', n_views=24, n=128)
(raw_lo, _), (raw_hi, _) = get_sinos(ct, ph, [lo, hi])
ee, i0, mus = gn_oracle.gn_model_tables(ct, lo, hi, matdecomp.xc.mixatten)
r32 = lambda a: np.asarray(a).astype(np.float32).astype(np.float64)      # noqa: E731
g, i0, mus = r32(np.stack([raw_lo, raw_hi])), r32(i0), r32(mus)
nv = 4
i0_t = np.tile(i0, (1, ct.N_channels)).reshape(2, ct.N_channels, len(ee))
for iters in (10, 25, 40):
    ref = gn_oracle.optimize_sino_cpu(g[:, :nv], ee, i0_t, mus, iters)
    a32 = gn_decompose(g[:, :nv], i0, mus, iters, dtype=F32)
    sc = gn_decompose.last_scratch
    n_flag = int(sc[8:12].view(torch.int32)[0])
    flagged = sc[16:16 + 4 * n_flag].view(torch.int32).cpu().numpy()
    a32 = a32.cpu().numpy()
    a64 = gn_decompose(g[:, :nv], i0, mus, iters, dtype=F64).cpu().numpy()
    scale = np.max(np.abs(ref))
    e32 = np.abs(a32 - ref).max(-1).reshape(-1) / scale
    e64 = np.abs(a64 - ref).max(-1).reshape(-1) / scale
    order = np.argsort(-e32)[:6]
    print(f'iters {iters}: scale {scale:.3f}  max e32 {e32.max():.3e}  max e64 {e64.max():.3e}  flagged {n_flag} {sorted(flagged.tolist())[:8]}')
    for p in order:
        v, c = divmod(int(p), ct.N_channels)
        print(f'   px {p} (view {v}, ch {c}): e32 {e32[p]:.2e} e64 {e64[p]:.2e} counts {g[0, v, c]:.4g} {g[1, v, c]:.4g} ref {ref[v, c]} a32 {a32[v, c]}'
              f' flagged {p in set(flagged.tolist())}')
