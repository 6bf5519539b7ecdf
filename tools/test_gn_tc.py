#!/usr/bin/env python3
"""Bring-up check of the tensor-core decomposition kernel (DEXCT_GN_TC=1): golden problems against the float64 kernel,
ragged sizes, and a timing at C2 size.  usage: test_gn_tc.py [swap]"""
import os
import sys

os.environ['DEXCT_GN_TC'] = '1'
if len(sys.argv) > 1:
    os.environ['DEXCT_GN_TC_SWAP'] = sys.argv[1]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'dex-ct-sim_b200')):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from dexct import _native  # noqa: E402
if os.environ.get('DEXCT_LIB_VARIANT'):          # a tuning build (make -C dex-ct-sim_b200/csrc variant VARFLAGS=...)
    _native.LIB_PATH = os.path.join(ROOT, 'dex-ct-sim_b200', 'csrc', '_build', os.environ['DEXCT_LIB_VARIANT'])
from dexct import synthetic as S  # noqa: E402
from dexct.ops import F32, F64, gn_decompose  # noqa: E402


def sre(a, r):
    return float(np.max(np.abs(a - r)) / np.max(np.abs(r)))


def main():
    for name in ('gn_synth_25it', 'gn_synth_10it', 'gn_c1_tables'):
        d = np.load(os.path.join(ROOT, 'tests', 'golden', name + '.npz'))
        g32 = d['g'].astype(np.float32).astype(np.float64)
        i032 = d['i0'].astype(np.float32).astype(np.float64)
        mus32 = d['mus'].astype(np.float32).astype(np.float64)
        n = int(d['n_iters'])
        ref = gn_decompose(g32, i032, mus32, n, dtype=F64).cpu().numpy()
        out = gn_decompose(d['g'], d['i0'], d['mus'], n, dtype=F32).cpu().numpy()
        os.environ['DEXCT_GN_TC'] = '0'
        old = gn_decompose(d['g'], d['i0'], d['mus'], n, dtype=F32).cpu().numpy()
        os.environ['DEXCT_GN_TC'] = '1'
        one = gn_decompose(d['g'], d['i0'], d['mus'], 1, dtype=F32).cpu().numpy()
        one64 = gn_decompose(g32, i032, mus32, 1, dtype=F64).cpu().numpy()
        print(f'{name}: nE={d["i0"].shape[1]} pixels={out.size // 2} tc vs f64 {sre(out, ref):.3e}  fp32-pipe vs f64 {sre(old, ref):.3e}  '
              f'one iteration {sre(one, one64):.3e}  finite {np.isfinite(out).all()}', flush=True)
    g, i0, mus, a_true = S.gn_problem(300, 1000)
    for P in (1, 37, 128, 129, 1000, 300000):
        gg = np.ascontiguousarray(g.reshape(2, -1)[:, :P])
        ref = gn_decompose(gg, i0, mus, 10, dtype=F64).cpu().numpy()
        out = gn_decompose(gg, i0, mus, 10, dtype=F32).cpu().numpy()
        print(f'P={P}: tc vs f64 {sre(out, ref):.3e}', flush=True)
    gd = torch.as_tensor(g.reshape(2, -1), dtype=torch.float32).cuda()
    gd = gd.repeat(1, 3).contiguous()
    i0d, musd = torch.as_tensor(i0, dtype=torch.float32).cuda(), torch.as_tensor(mus, dtype=torch.float32).cuda()
    for tc in ('1', '0'):
        os.environ['DEXCT_GN_TC'] = tc
        for _ in range(3):
            gn_decompose(gd, i0d, musd, 10, dtype=F32)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            gn_decompose(gd, i0d, musd, 10, dtype=F32)
        e1.record()
        torch.cuda.synchronize()
        print(f'TC={tc}: {e0.elapsed_time(e1) / 20:.4f} ms per call, {gd.shape[1]} pixels x {i0.shape[1]} bins x 10 iterations', flush=True)


if __name__ == '__main__':
    main()
