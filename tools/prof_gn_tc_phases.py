#!/usr/bin/env python3
"""Phase clocks of gn_tc_kernel (CTA 0, compute warp 0 and the MMA warp), from the -DGT_PROF build of the library
(`make -C dex-ct-sim_b200/csrc prof`).  usage: prof_gn_tc_phases.py [ctas_per_sm]"""
import ctypes as C
import os
import sys

os.environ['DEXCT_GN_TC'] = '1'
if len(sys.argv) > 1:
    os.environ['DEXCT_GN_TC_CTAS'] = sys.argv[1]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'dex-ct-sim_b200')):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from dexct import _native  # noqa: E402
_native.LIB_PATH = os.path.join(ROOT, 'dex-ct-sim_b200', 'csrc', '_build', 'libdexct_prof.so')
from dexct import synthetic as S  # noqa: E402
from dexct.ops import F32, gn_decompose  # noqa: E402

g, i0, mus, _ = S.gn_problem(300, 1000)
gd = torch.as_tensor(g.reshape(2, -1), dtype=torch.float32).cuda().repeat(1, 3).contiguous()
i0d, musd = torch.as_tensor(i0, dtype=torch.float32).cuda(), torch.as_tensor(mus, dtype=torch.float32).cuda()
lib = _native.lib()
out = (C.c_ulonglong * 16)()
gn_decompose(gd, i0d, musd, 10, dtype=F32)
torch.cuda.synchronize()
lib.dexct_debug_gn_tc_prof(out, 1)
gn_decompose(gd, i0d, musd, 10, dtype=F32)
torch.cuda.synchronize()
lib.dexct_debug_gn_tc_prof(out, 1)
v = np.array(list(out), dtype=np.float64)
chunks, iters, mchunks = v[5], v[6], v[11]
print(f'CTA 0: {int(iters)} tile-iterations, {int(chunks)} chunks (compute warp 0), {int(mchunks)} chunks (MMA warp)')
print(f'compute warp 0, cycles per chunk: free-wait {v[0] / chunks:.0f}, k-steps {v[1] / chunks:.0f}, st-wait + arrive {v[2] / chunks:.0f}')
print(f'compute warp 0, cycles per iteration: chunks {(v[0] + v[1] + v[2]) / iters:.0f}, done-wait {v[3] / iters:.0f}, '
      f'tcgen05.ld + Newton update {v[4] / iters:.0f}; done seen {v[7] / iters:.0f} cycles after the MMA warp committed it')
print(f'MMA warp, cycles per chunk: full-wait {v[8] / mchunks:.0f}, issue {v[9] / mchunks:.0f}, commit {v[10] / mchunks:.0f}')
print(f'tcgen05.alloc, all CTAs: mean wait {v[13] / 592:.0f} cycles (if 592 CTAs), max {v[14]:.0f}, CTAs that waited > 20000 cycles: {int(v[15])}')
print(f'CTA 0 lived {v[12]:.0f} cycles; sum of the MMA warp phases {v[8] + v[9] + v[10]:.0f}, of the compute warp phases {v[0] + v[1] + v[2] + v[3] + v[4]:.0f}')
n = 592
buf = (C.c_ulonglong * (3 * n))()
lib.dexct_debug_gn_tc_ctas(buf, n)
t = np.array(list(buf), dtype=np.float64).reshape(n, 3)
t0 = t[:, 0].min()
birth, death, sm = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3, t[:, 2].astype(int)
print(f'CTA births [us]: min {birth.min():.1f} median {np.median(birth):.1f} max {birth.max():.1f}; deaths: min {death.min():.1f} median {np.median(death):.1f} max {death.max():.1f}')
print('CTAs born later than 20 us after the first:', int((birth > 20).sum()), '; CTAs per SM: min', np.bincount(sm).min(), 'max', np.bincount(sm).max())
late = birth > 20
if late.any():
    print(f'late CTAs: born {birth[late].min():.1f}..{birth[late].max():.1f} us, live {np.median(death[late] - birth[late]):.1f} us (median); early CTAs live {np.median(death[~late] - birth[~late]):.1f} us')
print(f'CTA 0: born {birth[0]:.1f} us, died {death[0]:.1f} us -> SM clock {v[12] / (death[0] - birth[0]) / 1e3:.3f} GHz during the kernel')
q = np.percentile(death, [0, 10, 25, 50, 75, 90, 100])
print('CTA deaths [us], percentiles 0/10/25/50/75/90/100:', ' '.join(f'{x:.0f}' for x in q))
per_sm = np.array([death[sm == k].max() for k in np.unique(sm)])
print(f'last death per SM [us]: min {per_sm.min():.0f} median {np.median(per_sm):.0f} max {per_sm.max():.0f}')
