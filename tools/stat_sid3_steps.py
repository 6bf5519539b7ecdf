#!/usr/bin/env python3
"""Step statistics of the production 3-D walk (walk_skip3 through dexct_siddon_walk_trace) on one view of C4 / C5:
steps per ray, share of homogeneous stretches, run-length histogram, and what a warp pays (the maximum over its 32 rays).
usage: stat_sid3_steps.py [C4|C5] [view ...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'dex-ct-sim_b200')):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from dexct import synthetic as S  # noqa: E402
from dexct.ops import Projector, make_scan  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else 'C4'
    views = [int(v) for v in sys.argv[2:]] or [0, 75, 150]
    dev = torch.device('cuda', 0)
    ct, ph, lo, hi = S.make_config(name, device=dev)
    scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, n_rows=ct.N_rows, h_iso=ct.h_iso, rotation=ct.rotation)
    pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz), device=dev)
    for v in views:
        rows = sorted(set([0, ct.N_rows // 4, ct.N_rows // 2, ct.N_rows - 1]))
        n_steps, n_hom, runs, warp_max, warp_mean = [], [], [], [], []
        for r in rows:
            ids = (v * ct.N_rows + r) * ct.N_channels + np.arange(ct.N_channels, dtype=np.int64)
            recs, _ = pj.walk_trace(ids, 8, max_steps=2048, dtype='f32')
            ns = np.array([rec['n_steps'] for rec in recs])
            n_steps.append(ns)
            for rec in recs:
                if rec['n_steps']:
                    run = rec['run']
                    n_hom.append((run > 0).sum())
                    runs.append(run[run > 0])
            for w in range(0, len(ns) - 31, 32):
                warp_max.append(ns[w:w + 32].max())
                warp_mean.append(ns[w:w + 32].mean())
        # cost models of a warp (instruction counts measured on the SASS of walk_skip3: 97 per branch-free step; a stretch-only
        # step would take ~30, an ordinary column ~95): (A) today, every lane pays the general step; (B) two phases per round --
        # every lane first runs through its stretches up to the next ordinary column, then all lanes do that column together
        costA = costB = 0.0
        for r in rows:
            ids = (v * ct.N_rows + r) * ct.N_channels + np.arange(ct.N_channels, dtype=np.int64)
            recs, _ = pj.walk_trace(ids, 8, max_steps=2048, dtype='f32')
            for w in range(0, len(recs) - 31, 32):
                lanes = [rec['run'] > 0 for rec in recs[w:w + 32]]          # per lane: is step t a stretch?
                costA += 97.0 * max(len(x) for x in lanes)
                pos = [0] * 32
                while any(pos[q] < len(lanes[q]) for q in range(32)):
                    longest = 0
                    for q in range(32):
                        n = 0
                        while pos[q] < len(lanes[q]) and lanes[q][pos[q]]:
                            pos[q] += 1
                            n += 1
                        longest = max(longest, n)
                    costB += 30.0 * longest + 8.0
                    if any(pos[q] < len(lanes[q]) for q in range(32)):
                        costB += 95.0
                        for q in range(32):
                            if pos[q] < len(lanes[q]):
                                pos[q] += 1
        print(f'    warp cost model: today {costA:.3e} instructions, two-phase rounds {costB:.3e} ({costA / costB:.2f}x)')
        # (C) lockstep: all lanes of a warp always stand at the same column and advance by the SHORTEST run any of them has
        # there (a stretch may be cut short; an ordinary column anywhere makes it a one-column step) -- loads coalesce, no
        # lane idles, but a warp takes more steps
        stepsA = stepsC = 0
        for r in rows:
            ids = (v * ct.N_rows + r) * ct.N_channels + np.arange(ct.N_channels, dtype=np.int64)
            recs, _ = pj.walk_trace(ids, 8, max_steps=2048, dtype='f32')
            for w in range(0, len(recs) - 31, 32):
                group = [rec for rec in recs[w:w + 32] if rec['n_steps'] and rec['xmajor'] == recs[w]['xmajor']]
                if not group:
                    continue
                stepsA += max(rec['n_steps'] for rec in group)
                lo_c = min(int(rec['col'][0]) for rec in group)
                hi_c = max(int(rec['col'][-1]) + max(int(rec['run'][-1]), 1) for rec in group)
                # remaining run of every lane at every column (0 outside its range, 1 on ordinary columns)
                rem = np.zeros((len(group), hi_c - lo_c), dtype=np.int32)
                for q, rec in enumerate(group):
                    for c0, rn in zip(rec['col'], rec['run']):
                        n = max(int(rn), 1)
                        rem[q, c0 - lo_c:c0 - lo_c + n] = np.arange(n, 0, -1) if rn > 0 else 1
                c = 0
                while c < hi_c - lo_c:
                    live = rem[:, c][rem[:, c] > 0]
                    c += int(live.min()) if live.size else 1
                    stepsC += 1
        print(f'    lockstep model: {stepsC} warp steps against {stepsA} today ({stepsC / max(stepsA, 1):.2f}x)')
        ns = np.concatenate(n_steps)
        runs = np.concatenate(runs) if runs else np.zeros(0)
        hit = ns > 0
        print(f'{name} view {v}, rows {rows}: {hit.mean() * 100:.0f} % of the rays hit; steps per hitting ray mean {ns[hit].mean():.1f} '
              f'median {np.median(ns[hit]):.0f} max {ns.max()}; homogeneous stretches {np.sum(n_hom) / ns.sum() * 100:.0f} % of the steps, '
              f'mean run {runs.mean():.1f} columns (median {np.median(runs):.0f}, 90 % {np.percentile(runs, 90):.0f}, max {runs.max():.0f}); '
              f'columns covered by stretches {runs.sum() / max(1, runs.sum() + (ns.sum() - np.sum(n_hom))) * 100:.0f} %')
        print(f'    per warp of 32 adjacent channels: sum of maxima / sum of means = {np.sum(warp_max) / np.sum(warp_mean):.2f} '
              f'(steps a warp issues / steps its rays need)')
        h = np.bincount(np.minimum(runs.astype(int), 127), minlength=128)
        edges = [1, 2, 3, 5, 9, 14, 27, 41, 54, 128]
        print('    run-length histogram (share of stretches): ' + ', '.join(
            f'{a}-{b - 1}: {h[a:b].sum() / max(1, h.sum()) * 100:.0f}%' for a, b in zip(edges[:-1], edges[1:])))


if __name__ == '__main__':
    main()
