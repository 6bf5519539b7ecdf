#!/usr/bin/env python3
"""Device-side timings of BASELINE.json's other configurations at FULL size (bench.py's headline is C2):

  C3  Gauss-Newton sweep: 2048 x 2048 dual-energy pixel pairs, 120 bins, 10 iterations
  C4  512^3 phantom, 64 rows x 900 ch x 1200 views, 120 bins: projection (both spectra fused) + decomposition
  C5  1024^3 phantom, 128 rows x 1800 ch x 2400 views, 150 bins: the whole pipeline of /root/reference/main.py:100-178 --
      projection, FBP of both log sinograms, decomposition, FBP of both basis sinograms

    python tools/bench_configs.py C3 C4 C5                                  # one GPU
    python -m torch.distributed.run --nproc-per-node N ... tools/bench_configs.py C4 C5    # views sharded over N GPUs

Under torchrun the total work is fixed (strong scaling): rank r traces, decomposes and back-projects its own contiguous view
range; sinograms are assembled through peer memory (the GN kernel carries the count/log shard and mirrors its results, as in
bench.py; DEXCT_COLLECTIVE=nccl: all-gathers) and the partial images summed with one all-reduce per volume (SURVEY.md
section 8e).  Times are CUDA-event times on the launching stream, max over ranks; inputs are resident before the timed
region; every working set is far larger than the 126 MB L2 (C3: 268 MB of counts in, C4/C5: GBs).  One JSON line per
configuration, printed by rank 0.
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'dex-ct-sim_b200')):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import matdecomp  # noqa: E402
from dexct import dist as ddist  # noqa: E402
from dexct import synthetic as S  # noqa: E402
from dexct.ops import Projector, count_segments, fbp_fan, gn_decompose, make_scan  # noqa: E402
from xtomosim import forward_project as FP  # noqa: E402

GN_ITERS = 10
MASK_THRESH = 0.95


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:                                                     # noqa: BLE001
        return {}


class Stages:
    """CUDA-event stopwatch over named stages of a step; keeps the per-stage mean over the timed repetitions."""

    def __init__(self):
        self.marks = []
        self.cur = None

    def begin(self):
        self.cur = [('start', self._ev())]

    def mark(self, name):
        self.cur.append((name, self._ev()))

    def end(self):
        self.marks.append(self.cur)

    @staticmethod
    def _ev():
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def means(self):
        out = {}
        for run in self.marks:
            for (_, a), (name, b) in zip(run[:-1], run[1:]):
                out[name] = out.get(name, 0.0) + a.elapsed_time(b) / len(self.marks)
        out['total'] = sum(run[0][1].elapsed_time(run[-1][1]) for run in self.marks) / len(self.marks)
        return out


def max_over_ranks(d, dev, ws):
    keys = sorted(d)
    t = torch.tensor([d[k] for k in keys], dtype=torch.float64, device=dev)
    if ws > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return dict(zip(keys, [float(x) for x in t.cpu()]))


def run_c3(dev, rank, ws, reps):
    g, i0, mus, a_true = S.gn_problem(2048, 2048, 120)
    P = 2048 * 2048
    p0, pn = ddist.shard_range(2048, rank, ws)                # rows of the pixel grid
    gd = torch.as_tensor(g[:, p0:p0 + pn].reshape(2, -1), dtype=torch.float32, device=dev).contiguous()
    i0d = torch.as_tensor(i0, dtype=torch.float32, device=dev)
    musd = torch.as_tensor(mus, dtype=torch.float32, device=dev)
    out = torch.empty((gd.shape[1], 2), dtype=torch.float32, device=dev)
    st = Stages()
    for it in range(reps + 2):
        if it >= 2:
            st.begin()
        gn_decompose(gd, i0d, musd, GN_ITERS, mask_thresh=0.0, out=out)
        if it >= 2:
            st.mark('gn')
            st.end()
    torch.cuda.synchronize()
    t = max_over_ranks(st.means(), dev, ws)
    err = float((out.double().cpu() - torch.as_tensor(a_true[p0:p0 + pn].reshape(-1, 2))).abs().max())
    flop = 28.0 * 120 * GN_ITERS
    sm_mhz = peaks().get('sm_max_mhz', 1965.0)
    fma_peak = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12 * ws
    ach = flop * P / (t['gn'] * 1e-3) / 1e12
    return {'workload': 'C3: GN sweep, 2048 x 2048 pixel pairs, 120 bins, 10 iterations, water/bone-like bases', 'n_gpus': ws,
            'scaling': 'strong', 'pixels': P, 'ms': t['gn'], 'gn_pixels_per_s': P / (t['gn'] * 1e-3),
            'roofline_gn': {'bound': 'fp32_fma', 'achieved': ach, 'peak': fma_peak, 'unit': 'TFLOP/s', 'frac': ach / fma_peak,
                            'flop_per_pixel': flop, 'clock_mhz_assumed': sm_mhz},
            'hbm_bytes_per_pixel': 16, 'max_abs_err_vs_a_true_after_10_iters': err}


def run_scan(name, dev, rank, ws, reps, with_fbp):
    t0 = time.time()
    ct, ph, lo, hi = S.make_config(name, device=dev)
    t_phantom = time.time() - t0
    n_views = ct.N_proj
    v0, nv = ddist.shard_range(n_views, rank, ws)
    scan = make_scan(n_views, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, n_rows=ct.N_rows, h_iso=ct.h_iso, rotation=ct.rotation)
    pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz), device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    pj.prepare()
    e1.record()
    torch.cuda.synchronize()
    t_prepare = e0.elapsed_time(e1)
    td = torch.float32
    mu = torch.as_tensor(ph.mu_table(lo.E), dtype=td, device=dev)
    W = torch.as_tensor(np.stack([FP.spectral_weights(ct, lo), FP.spectral_weights(ct, hi)]), dtype=td, device=dev)
    _, i0_h, mus_h = matdecomp.spectral_tables(ct, lo, hi)
    i0 = torch.as_tensor(i0_h, dtype=td, device=dev)
    mus = torch.as_tensor(mus_h, dtype=td, device=dev)
    R, Cn = ct.N_rows, ct.N_channels
    even = n_views % ws == 0
    pg = None
    if ws > 1 and even and os.environ.get('DEXCT_COLLECTIVE', 'peer') == 'peer':
        # sinograms assembled through peer memory: the GN kernel carries the count/log shard to the other ranks' copies
        # and stores its own results into all of them (dexct/peer.py), as bench.py does
        from dexct.peer import PeerGroup
        pg = PeerGroup()
        slot_s, slot_b = 4 * nv * R * Cn * 4, nv * R * Cn * 2 * 4
        sbuf, bbuf = pg.alloc(ws * slot_s), pg.alloc(ws * slot_b)
        g_sino = sbuf.view(td, (ws, 4, nv, R, Cn))
        g_basis = bbuf.view(td, (ws, nv, R, Cn, 2))
        counts, logs, basis = g_sino[rank, 0:2], g_sino[rank, 2:4], g_basis[rank]
        mirrors = bbuf.table(rank * slot_b, exclude_self=True)
        push_table = sbuf.table(rank * slot_s, exclude_self=True)
    else:
        counts = torch.empty((2, nv, R, Cn), dtype=td, device=dev)
        logs = torch.empty_like(counts)
        basis = torch.empty((nv, R, Cn, 2), dtype=td, device=dev)
        if ws > 1 and even:
            g_counts = torch.empty((ws, 2, nv, R, Cn), dtype=td, device=dev)
            g_logs = torch.empty_like(g_counts)
            g_basis = torch.empty((ws, nv, R, Cn, 2), dtype=td, device=dev)
    n_matrix, fov, ramp = ph.Nx, ph.Nx * ph.dx, 0.8
    images = [torch.empty((R, n_matrix, n_matrix), dtype=td, device=dev) for _ in range(4)] if with_fbp else []

    st = Stages()

    def step():
        st.begin()
        pj.project(mu, W, view_begin=v0, view_count=nv, counts=counts, logs=logs)
        st.mark('siddon_poly')
        pending = []
        if pg is None and ws > 1 and even:
            pending.append(dist.all_gather_into_tensor(g_counts, counts, async_op=True))
            pending.append(dist.all_gather_into_tensor(g_logs, logs, async_op=True))
        if with_fbp:
            for s in range(2):
                fbp_fan(scan, logs[s], n_matrix, fov, ramp, view_begin=v0, image=images[s])
            st.mark('fbp_2_log_sinograms')
        if pg is not None:
            gn_decompose(counts.reshape(2, -1), i0, mus, GN_ITERS, mask_thresh=MASK_THRESH, out=basis.view(-1, 2),
                         reduce_max_over_ranks=True, peers=pg, mirrors=mirrors, push=(g_sino[rank], push_table))
        else:
            gn_decompose(counts.view(2, -1), i0, mus, GN_ITERS, mask_thresh=MASK_THRESH, out=basis.view(-1, 2),
                         reduce_max_over_ranks=ws > 1)
        st.mark('gn')
        if pg is None and ws > 1 and even:
            pending.append(dist.all_gather_into_tensor(g_basis, basis, async_op=True))
        if with_fbp:
            for m in range(2):
                fbp_fan(scan, basis[..., m].contiguous(), n_matrix, fov, ramp, view_begin=v0, image=images[2 + m])
            st.mark('fbp_2_basis_sinograms')
            if ws > 1:
                for im in images:
                    dist.all_reduce(im)
        for w in pending:
            w.wait()
        if pg is not None:
            pg.barrier()
        st.mark('collectives_tail')
        st.end()

    step()                                                   # warm-up (allocator, NCCL channels)
    st.marks.clear()
    if ws > 1:
        dist.barrier()
    torch.cuda.synchronize()
    for _ in range(reps):
        step()
    if ws > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t = max_over_ranks(st.means(), dev, ws)

    # segments per ray on a sample of this rank's views (the unit of the bytes-per-ray roofline)
    sample = list(range(v0, v0 + nv, max(1, nv // 8)))[:8]
    seg = sum(count_segments(pj, v, 1) for v in sample) / (len(sample) * R * Cn)
    segt = torch.tensor([seg], dtype=torch.float64, device=dev)
    if ws > 1:
        dist.all_reduce(segt)
        segt /= ws
    seg = float(segt[0])
    rays = n_views * R * Cn
    hbm = float(peaks().get('hbm_gbs', 6650.0))
    bpr = seg + 2 * 4
    ach = rays * bpr / (t['siddon_poly'] * 1e-3) / 1e9
    sm_mhz = peaks().get('sm_max_mhz', 1965.0)
    fma_peak = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12 * ws
    gn_ach = 28.0 * len(lo.E) * GN_ITERS * rays / (t['gn'] * 1e-3) / 1e12
    finite = bool(torch.isfinite(basis).all()) and bool(torch.isfinite(counts).all())
    c = S.CONFIGS[name]
    line = {'workload': f"{name}: {c['n']}x{c['n']}x{c['nz']} phantom (K=8), {R} rows x {Cn} ch x {n_views} views, {c['n_bins']} bins, "
                        f"2 spectra fused, GN {GN_ITERS} iters" + (f', 4 x FBP -> {R} x {n_matrix}^2 (ramp {ramp})' if with_fbp else ''),
            'n_gpus': ws, 'scaling': 'strong', 'reps': reps, 'rays': rays, 'ms': t, 'rays_per_s_pipeline': rays / (t['total'] * 1e-3),
            'siddon_rays_per_s': rays / (t['siddon_poly'] * 1e-3), 'gn_pixels_per_s': rays / (t['gn'] * 1e-3),
            'roofline': {'bound': 'hbm', 'achieved': ach, 'peak': hbm * ws, 'unit': 'GB/s', 'frac': ach / (hbm * ws),
                         'bytes_per_ray': bpr, 'segments_per_ray': seg, 'kernel': 'siddon_skip3d_kernel'},
            'roofline_gn': {'bound': 'fp32_fma', 'achieved': gn_ach, 'peak': fma_peak, 'unit': 'TFLOP/s', 'frac': gn_ach / fma_peak},
            'prepare_ms': t_prepare, 'phantom_rasterise_s': t_phantom, 'all_finite': finite,
            'device_mem_gb': torch.cuda.max_memory_allocated(dev) / 1e9,
            'collective': (('peer memory (GN kernel with fused push and mirrored epilogue)' if pg is not None else 'all-gather counts/logs/basis')
                           + (' + all-reduce of 4 partial volumes' if with_fbp else '')) if ws > 1 and even
                          else ('all-reduce of 4 partial volumes' if ws > 1 and with_fbp else 'none')}
    if pg is not None:
        assert pg.status() == 0
        del counts, logs, basis, g_sino, g_basis
        pg.close()
    if with_fbp:
        pv = 2.0 * R * n_matrix * n_matrix * n_views
        line['fbp_pixel_views_per_s'] = pv / (t['fbp_2_log_sinograms'] * 1e-3)
        line['recon_centre_mu_low'] = float(images[0][R // 2, n_matrix // 2, n_matrix // 2])
    return line


def main():
    which = [a for a in sys.argv[1:] if not a.startswith('-')] or ['C3', 'C4']
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    ws = int(os.environ.get('WORLD_SIZE', '1'))
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if ws > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)
    for name in which:
        torch.cuda.reset_peak_memory_stats(dev)
        if name == 'C3':
            line = run_c3(dev, rank, ws, reps=5)
        else:
            line = run_scan(name, dev, rank, ws, reps=3 if name == 'C4' else 2, with_fbp=(name == 'C5'))
        if rank == 0:
            print(json.dumps(line), flush=True)
        torch.cuda.empty_cache()
    if ws > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
