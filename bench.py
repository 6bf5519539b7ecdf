#!/usr/bin/env python3
"""Benchmark of the hot path: Siddon polyenergetic forward projection + Gauss-Newton two-basis decomposition.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1], "C2"): synthetic 512x512 single-slice phantom, 900 channels x 1000 views, 120 energy
bins, dual-spectrum sinograms (both spectra in ONE traversal) followed by the GN decomposition of every sinogram pixel
(10 iterations, the fixed count configs[2] names).  One step = one whole scan: 900 000 rays -> 2 count sinograms ->
900 000 decomposed pixel pairs.  Unit = rays (= pixels) per second through the whole path.

N > 1 (weak scaling): the scan has N x 1000 views; rank r owns the contiguous view range [1000 r, 1000 (r+1)), traces and
decomposes it without communication, and the view-major sinograms are assembled with one NCCL all-gather per array.

`value`   : device throughput, inputs resident in HBM, CUDA events around every step, max over ranks.
`e2e`     : the same work through the reference-facing Python entry points (get_sino x2 / get_basismat_sinos) with host
            numpy arrays in and out; every step uploads the phantom again and reads the results back.
`roofline`: the Siddon kernel against the bytes-per-ray HBM roofline of SURVEY.md section 8(d); `roofline_gn`: the GN kernel
            against the FP32 FMA pipe.
`cpu_baseline` / `--impl reference`: the CPU oracle (oracle/: C restatement of Siddon 1985 with OpenMP + the numpy
            restatement of the reference's optimize_sino_cpu, one process per core) on a bounded sample of the same scan.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, 'dex-ct-sim_b200')
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

VIEWS_PER_GPU = 1000
N_CH = 900
GN_ITERS = 10
MASK_THRESH = 0.95
METRIC = 'siddon_poly_gn_rays_per_s'
UNIT = 'rays/s'
WORKLOAD = ('C2: 512x512x1 phantom (K=8 materials), 900 ch x 1000 views per GPU, 120 energy bins, 2 spectra fused, '
            'GN 10 iters water/bone, air mask 0.95')


# N > 1: how the finished count/log shard reaches the peers' copies.  'fused' (default): the decomposition kernel's own
# threads copy it, a slice per Newton iteration (dexct_gn_decompose_fused_push) -- spread over all SMs and over the kernel's
# run time.  'sm': a separate store kernel on DEXCT_PUSH_CTAS CTAs on a side stream; 'dma': seven copy-engine transfers.
# Measured at N = 8 (profiles/r01h_scaling.txt): sm/16 CTAs 0.855 ms per step, 48 CTAs 0.876, 148 CTAs 0.920, dma 0.905,
# NCCL all-gathers 1.195.
PUSH_MODE = os.environ.get('DEXCT_PUSH', 'fused')
PUSH_CTAS = int(os.environ.get('DEXCT_PUSH_CTAS', '16'))


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--dtype', default='f32', choices=['f32', 'f64'])
    ap.add_argument('--cpu-views', type=int, default=0, help='views in the CPU sample (0 = auto)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-extra', action='store_true', help='skip the secondary measurements (3-D projector, FBP)')
    ap.add_argument('--collective', default='peer', choices=['peer', 'nccl'],
                    help='N > 1: assemble the sinograms by stores into peer memory (default) or by NCCL all-gathers')
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------------------------
# CPU arm (oracle).  Used for `cpu_baseline` and `--impl reference`.

def _gn_view_job(args):
    from oracle import gn_oracle
    g, i0, mus, n_iters = args
    nB = g.shape[2]
    i0_t = np.tile(i0, (1, nB)).reshape(2, nB, i0.shape[1])
    return gn_oracle.optimize_sino_cpu(g, None, i0_t, mus, n_iters)


class CpuArm:
    """The oracle's CPU implementation of one bounded sample of the workload: `n_views` views spread evenly over the
    rotation; Siddon + detection in C with OpenMP over all host threads, GN in numpy with one worker process per core."""

    def __init__(self, n_views):
        import multiprocessing as mp
        import oracle
        from oracle import gn_oracle
        from dexct import synthetic as S
        from xtomosim.forward_project import spectral_weights
        oracle.build()
        self.oracle = oracle
        ct, ph, lo, hi = S.make_config('C2', n_views=VIEWS_PER_GPU)
        tabs = oracle.fan_tables(ct.N_channels, ct.N_proj, ct.gamma_fan, ct.rotation, ct.beta0, ct.N_rows, ct.h_iso, ct.SID, ct.SDD)
        pick = np.linspace(0, ct.N_proj, n_views, endpoint=False).astype(int)
        tabs = {k: (v[pick] if k.endswith('beta') else v) for k, v in tabs.items()}
        self.scan = oracle.Scan(ph.M, (ph.dx, ph.dy, ph.dz), ct.SID, ct.SDD, tabs)
        self.mu = ph.mu_table(lo.E)
        self.W = np.stack([spectral_weights(ct, lo), spectral_weights(ct, hi)])
        import matdecomp
        _, self.i0, self.mus = gn_oracle.gn_model_tables(ct, lo, hi, matdecomp.xc.mixatten)
        self.n_views = n_views
        self.cores = os.cpu_count() or 1
        self.threads = oracle.num_threads()
        self.pool = mp.get_context('spawn').Pool(self.cores, initializer=_worker_init) if self.cores > 1 else None
        self.seg_per_ray = None

    def step(self):
        counts, _, seg = self.scan.project_poly(self.mu, self.W, want_log=True)
        self.seg_per_ray = seg / (self.n_views * N_CH)
        g = counts[:, :, 0, :]                                            # [2, views, ch]
        mask = g[0] >= MASK_THRESH * g[0].max()
        jobs = [(g[:, v:v + 1], self.i0, self.mus, GN_ITERS) for v in range(self.n_views)]
        outs = self.pool.map(_gn_view_job, jobs, chunksize=1) if self.pool else [_gn_view_job(j) for j in jobs]
        a = np.concatenate(outs, axis=0)
        a[mask] = 0
        return a

    def close(self):
        if self.pool:
            self.pool.close()
            self.pool.join()

    def describe(self, value):
        return {'value': value, 'unit': UNIT, 'cores': self.cores, 'kind': 'port',
                'sample': f'{self.n_views} of {VIEWS_PER_GPU} views of the C2 scan, evenly spaced ({self.n_views * N_CH} rays): '
                          f'oracle/siddon_oracle.c float64 Siddon 1985 + Beer-Lambert (OpenMP, {self.threads} threads) then '
                          f'oracle/gn_oracle.py optimize_sino_cpu float64 numpy ({GN_ITERS} iters, {self.cores} worker processes)'}


def auto_cpu_views():
    # ~80 ms per view and core for the numpy Newton loop at 900 x 120 x 10: 24 views per core
    # make a step of about 2-3 s
    return int(min(VIEWS_PER_GPU, 24 * (os.cpu_count() or 1)))


def _worker_init():
    os.environ['OMP_NUM_THREADS'] = '1'


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    arm = CpuArm(args.cpu_views or auto_cpu_views())
    for _ in range(min(args.warmup, 1)):
        arm.step()
    t = []
    for _ in range(max(1, min(args.steps, 5))):
        t0 = time.perf_counter()
        arm.step()
        t.append(time.perf_counter() - t0)
    arm.close()
    steps = len(t)
    ms = 1e3 * sum(t) / steps
    value = arm.n_views * N_CH / (ms * 1e-3)
    line = {'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': steps,
            'warmup': min(args.warmup, 1), 'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': {'workload': WORKLOAD, 'step': 'bounded sample: see cpu_baseline.sample'},
            'cpu_baseline': arm.describe(value),
            'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------------------
# clocks

class ClockSampler(threading.Thread):
    def __init__(self, index, period=0.02):
        super().__init__(daemon=True)
        self.period = period
        self.samples = []
        self.reasons = set()
        self.stop_flag = threading.Event()
        self.ok = False
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            self.nv = pynvml
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(('GPU-' + str(torch.cuda.get_device_properties(index).uuid)).encode())
            except Exception:                                             # noqa: BLE001
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:                                            # noqa: BLE001
            self.err = repr(e)

    NAMES = {0x1: 'gpu_idle', 0x2: 'applications_clocks_setting', 0x4: 'sw_power_cap', 0x8: 'hw_slowdown', 0x10: 'sync_boost',
             0x20: 'sw_thermal_slowdown', 0x40: 'hw_thermal_slowdown', 0x80: 'hw_power_brake_slowdown', 0x100: 'display_clock_setting'}

    def sample(self):
        nv = self.nv
        mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
        try:
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:                                                 # noqa: BLE001
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        util = nv.nvmlDeviceGetUtilizationRates(self.h).gpu
        self.samples.append((mhz, util))
        for bit, name in self.NAMES.items():
            if r & bit and name != 'gpu_idle':
                self.reasons.add(name)

    def run(self):
        while self.ok and not self.stop_flag.is_set():
            try:
                self.sample()
            except Exception:                                             # noqa: BLE001
                pass
            time.sleep(self.period)

    def finish(self):
        self.stop_flag.set()
        if self.ok:
            self.join(timeout=2)
        if not self.ok or not self.samples:
            return {'sm_mhz': None, 'sm_max_mhz': getattr(self, 'max_mhz', None), 'reasons': [], 'note': 'no samples'}
        mhz = sorted(m for m, _ in self.samples)
        return {'sm_mhz': mhz[len(mhz) // 2], 'sm_max_mhz': self.max_mhz, 'reasons': sorted(self.reasons),
                'samples': len(mhz)}


# ----------------------------------------------------------------------------------------------------------------
# B200 arm

def run_b200(args):
    import torch
    import torch.distributed as dist

    from dexct import _native as N
    from dexct import synthetic as S
    from dexct import dist as ddist
    from dexct.ops import Projector, count_segments, dtype_code, gn_decompose, make_scan
    from xtomosim import forward_project as FP
    import matdecomp

    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    ws = int(os.environ.get('WORLD_SIZE', '1'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device -- this framework has no CPU fallback (use --impl reference for the CPU arm)')
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if ws > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)
    N.require_device()
    code = dtype_code(args.dtype)
    td = torch.float32 if args.dtype == 'f32' else torch.float64

    # ---- the scan: ws x 1000 views, this rank owns [v0, v0 + nv)
    n_views = VIEWS_PER_GPU * ws
    ct, ph, lo, hi = S.make_config('C2', n_views=n_views)
    v0, nv = ddist.shard_range(n_views, rank, ws)
    scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, rotation=ct.rotation)
    pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz), device=dev)
    mu = torch.as_tensor(ph.mu_table(lo.E), dtype=td, device=dev)
    W = torch.as_tensor(np.stack([FP.spectral_weights(ct, lo), FP.spectral_weights(ct, hi)]), dtype=td, device=dev)
    ee, i0_h, mus_h = matdecomp.spectral_tables(ct, lo, hi)
    i0 = torch.as_tensor(i0_h, dtype=td, device=dev)
    mus = torch.as_tensor(mus_h, dtype=td, device=dev)
    rays_local = nv * N_CH
    # N > 1: every rank ends the step holding the whole view-major sinograms.  'peer': the buffers are symmetric
    # allocations mapped into every rank (dexct/peer.py); the projection writes its shard in place, the decomposition
    # kernel carries that shard into the other ranks' copies while it iterates (DEXCT_PUSH=sm / dma: a store kernel or the
    # copy engines on a side stream instead), its epilogue stores each basis pixel into all copies, and the air maximum is
    # exchanged through flags in peer memory -- no NCCL call in the step.  'nccl': the same with three all-gathers and
    # one all-reduce.
    pg = None
    collective = 'none'
    esz = 4 if args.dtype == 'f32' else 8
    slot_s, slot_b = 4 * nv * N_CH * esz, nv * N_CH * 2 * esz
    if ws > 1 and args.collective == 'peer':
        from dexct.dist import all_agree
        from dexct.peer import PeerGroup
        try:                                                              # (a failure on one rank raises on all: peer.py)
            pg = PeerGroup()
            sbuf, bbuf = pg.alloc(ws * slot_s), pg.alloc(ws * slot_b)
        except Exception as exc:                                          # noqa: BLE001
            print(f'bench.py: rank {rank}: peer memory unavailable ({exc})', file=sys.stderr)
            sbuf = None
        if not all_agree(pg is not None and sbuf is not None):            # belt and braces: all ranks take the same path
            pg = None             # (not closed: closing is collective, and the ranks may not agree on what exists)
            if rank == 0:
                print('bench.py: using NCCL all-gathers on every rank', file=sys.stderr)
        else:
            collective = 'peer'
    if ws > 1 and pg is None:
        collective = 'nccl'
    if pg is not None:
        g_sino = sbuf.view(td, (ws, 4, nv, 1, N_CH))                      # [rank][counts lo, hi, logs lo, hi]
        g_basis = bbuf.view(td, (ws, nv, 1, N_CH, 2))
        counts, logs, basis = g_sino[rank, 0:2], g_sino[rank, 2:4], g_basis[rank]
        mirrors = bbuf.table(rank * slot_b, exclude_self=True)
        push_table = sbuf.table(rank * slot_s, exclude_self=True)
        side = torch.cuda.Stream(device=dev)
    else:
        counts = torch.empty((2, nv, 1, N_CH), dtype=td, device=dev)
        logs = torch.empty_like(counts)
        basis = torch.empty((nv, 1, N_CH, 2), dtype=td, device=dev)
        if ws > 1:
            g_counts = torch.empty((ws, 2, nv, 1, N_CH), dtype=td, device=dev)
            g_logs = torch.empty_like(g_counts)
            g_basis = torch.empty((ws, nv, 1, N_CH, 2), dtype=td, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2

    ev = lambda: torch.cuda.Event(enable_timing=True)                   # noqa: E731

    def step(timers=None):
        if timers is not None:
            timers[0].record()
        pj.project(mu, W, view_begin=v0, view_count=nv, dtype=code, want_log=True, counts=counts, logs=logs)
        if timers is not None:
            timers[1].record()
        if pg is not None:
            if PUSH_MODE == 'fused':
                # one kernel: Newton iterations + the copy of the count/log shard into the peers, a slice per iteration
                gn_decompose(counts.reshape(2, -1), i0, mus, GN_ITERS, mask_thresh=MASK_THRESH, dtype=code, out=basis.view(-1, 2),
                             reduce_max_over_ranks=True, peers=pg, mirrors=mirrors, push=(g_sino[rank], push_table))
                if timers is not None:
                    timers[2].record()
            else:
                main = torch.cuda.current_stream()
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    pg.push(g_sino[rank], sbuf, rank * slot_s, dma=PUSH_MODE == 'dma', max_ctas=PUSH_CTAS)
                gn_decompose(counts.reshape(2, -1), i0, mus, GN_ITERS, mask_thresh=MASK_THRESH, dtype=code, out=basis.view(-1, 2),
                             reduce_max_over_ranks=True, peers=pg, mirrors=mirrors)
                if timers is not None:
                    timers[2].record()
                main.wait_stream(side)
            pg.barrier()                                    # every rank's stores have landed in every copy
            return
        pending = []
        if ws > 1:
            # the sinogram gathers ride NCCL's stream behind the projection and overlap the decomposition, which only
            # needs the local view shard
            pending.append(dist.all_gather_into_tensor(g_counts, counts, async_op=True))
            pending.append(dist.all_gather_into_tensor(g_logs, logs, async_op=True))
        gn_decompose(counts.view(2, -1), i0, mus, GN_ITERS, mask_thresh=MASK_THRESH, dtype=code, out=basis.view(-1, 2),
                     reduce_max_over_ranks=ws > 1)
        if timers is not None:
            timers[2].record()
        if ws > 1:
            pending.append(dist.all_gather_into_tensor(g_basis, basis, async_op=True))
            for w in pending:
                w.wait()

    def barrier():
        if ws > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        step()
    barrier()

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    marks = []
    barrier()
    for _ in range(args.steps):
        flush.zero_()
        t = [ev() for _ in range(4)]
        step(t)
        t[3].record()
        marks.append(t)
    barrier()
    clocks = sampler.finish() if sampler else None
    peer_ok = None
    if pg is not None:
        # the peer-assembled buffers against NCCL's all-gather of the same shards (outside the timed region)
        chk_s, chk_b = torch.empty_like(g_sino), torch.empty_like(g_basis)
        dist.all_gather_into_tensor(chk_s, g_sino[rank].contiguous())
        dist.all_gather_into_tensor(chk_b, g_basis[rank].contiguous())
        ok = torch.tensor([int(torch.equal(chk_s, g_sino) and torch.equal(chk_b, g_basis))], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        peer_ok = bool(int(ok[0]))
        del chk_s, chk_b
    tot = sum(t[0].elapsed_time(t[3]) for t in marks)
    t_sid = sum(t[0].elapsed_time(t[1]) for t in marks) / args.steps
    t_gn = sum(t[1].elapsed_time(t[2]) for t in marks) / args.steps
    tt = torch.tensor([tot, t_sid, t_gn], dtype=torch.float64, device=dev)
    if ws > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    tot, t_sid, t_gn = [float(x) for x in tt.cpu()]
    ms = tot / args.steps
    value = rays_local * ws / (ms * 1e-3)

    # ---- end to end through the reference-facing entry points, host buffers in and out
    e2e = None
    if not args.no_e2e:
        from dexct import config as cfg
        cfg.dtype = args.dtype
        k_e2e = max(3, min(args.steps, 10))

        # N > 1: every rank keeps the host arrays of ITS OWN view range (config.gather_views off) -- the data-parallel form
        # in which each rank writes its part of the output; with the default (every rank returns the whole arrays, like
        # the single-process reference) each rank would move the whole scan over its PCIe link and the host side could
        # not scale.  The air-mask maximum is still reduced over all ranks.
        cfg.gather_views = ws == 1

        def e2e_step():
            FP.clear_cache()                                            # the phantom is uploaded and re-laid-out every step
            (raw_lo, log_lo), (raw_hi, log_hi) = FP.get_sinos(ct, ph, [lo, hi])
            m1, m2 = matdecomp.get_basismat_sinos(ct, raw_lo, raw_hi, lo, hi, n_iters=GN_ITERS, mask_thresh=MASK_THRESH)
            return raw_lo, m1

        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(k_e2e):
            e2e_step()
        barrier()
        dt = torch.tensor([(time.perf_counter() - t0) / k_e2e], dtype=torch.float64, device=dev)
        if ws > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        esz = 4 if args.dtype == 'f32' else 8
        # up: labels + tables + the two float64 count sinograms handed to get_basismat_sinos (they sit in page-locked
        # memory: get_sinos returned them); down: 2 x (raw, log) + 2 basis sinograms, float64 like the reference's arrays
        # summed over ranks; the KB-sized tables are uploaded once per content
        h2d = ws * ph.M.nbytes + 2 * n_views * N_CH * 8
        d2h = 6 * n_views * N_CH * 8
        e2e = {'value': n_views * N_CH / float(dt[0]), 'unit': UNIT, 'h2d_bytes_per_step': int(h2d), 'd2h_bytes_per_step': int(d2h),
               'ms_per_step': 1e3 * float(dt[0]), 'steps': k_e2e,
               'api': 'xtomosim.forward_project.get_sinos + matdecomp.get_basismat_sinos (float64 numpy in/out through page-locked '
                      'buffers, phantom re-uploaded and re-laid-out every step'
                      + (')' if ws == 1 else '; every rank keeps the host arrays of its own view range, bytes summed over ranks)')}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        except Exception:                                                 # noqa: BLE001
            pass
        hbm = float(peaks.get('hbm_gbs', 6650.0))
        seg = count_segments(pj, v0, min(nv, 100)) / (min(nv, 100) * N_CH)
        traffic, traffic_src = None, None
        try:                                                              # measured under ncu, committed with its source
            tr = json.load(open(os.path.join(ROOT, 'profiles', 'ncu_traffic.json')))['siddon_skip_kernel']
            traffic, traffic_src = tr['dram_bytes_read'] + tr['dram_bytes_write'], tr['source']
        except Exception:                                                 # noqa: BLE001
            pass
        bytes_per_ray = seg * 1 + 2 * (4 if args.dtype == 'f32' else 8)
        ach = rays_local * bytes_per_ray / (t_sid * 1e-3) / 1e9
        sm_mhz = (clocks or {}).get('sm_mhz') or peaks.get('sm_max_mhz', 1965.0)
        fma_peak = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12                    # TFLOP/s fp32 at the clock seen during the run
        nE = int(i0.shape[1])
        gn_flops = 28.0 * nE * GN_ITERS * rays_local
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': ws, 'steps': args.steps, 'warmup': max(args.warmup, 3),
            'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': args.dtype,
            'data': 'synthetic',
            'config': {'workload': WORKLOAD, 'views_total': n_views, 'rays_per_step': rays_local * ws, 'gn_iters': GN_ITERS,
                       'l2': 'flushed between timed steps (256 MiB write)', 'parallelism': f'view-sharded x{ws}',
                       'collective': {'none': 'none',
                                      'peer': 'none from a library: counts/logs pushed into the peers\' copies by a store kernel beside the '
                                              'decomposition, basis sinograms stored to all copies by the GN epilogue, air maximum and end-of-step '
                                              'barrier through flags in peer memory (CUDA IPC over NVLink)',
                                      'nccl': 'all-gather of counts/logs (overlapping the decomposition) and basis sinograms + MAX all-reduce '
                                              'inside the step'}[collective]},
            'kernel_ms': {'siddon_poly': t_sid, 'gn': t_gn},
            **({'peer_buffers_equal_nccl_all_gather': peer_ok} if peer_ok is not None else {}),
            'siddon_rays_per_s': rays_local * ws / (t_sid * 1e-3), 'gn_pixels_per_s': rays_local * ws / (t_gn * 1e-3),
            'roofline': {'bound': 'hbm', 'achieved': ach, 'peak': hbm, 'unit': 'GB/s', 'frac': ach / hbm, 'traffic': traffic,
                         'traffic_source': traffic_src, 'algorithmic_bytes_per_launch': rays_local * bytes_per_ray,
                         'kernel': 'siddon_skip_kernel', 'bytes_per_ray': bytes_per_ray, 'segments_per_ray': seg,
                         'peak_source': 'MEASURED_PEAKS.json hbm_gbs' if peaks else 'fallback'},
            'roofline_gn': {'bound': 'fp32_fma', 'achieved': gn_flops / (t_gn * 1e-3) / 1e12, 'peak': fma_peak, 'unit': 'TFLOP/s',
                            'frac': gn_flops / (t_gn * 1e-3) / 1e12 / fma_peak, 'kernel': 'gn_kernel',
                            'flop_per_pixel': 28.0 * nE * GN_ITERS},
            # per step: siddon_skip, gn_max, gn_table, gn_kernel, gn_fixup (+ peer_push, peer_max, peer_barrier at N > 1)
            'gpu_launches': ((7 if PUSH_MODE == 'fused' else 8) if pg is not None else 5) * args.steps, 'clocks': clocks,
        }
        if e2e:
            line['e2e'] = e2e
        if ws == 1 and not args.no_cpu_baseline:
            arm = CpuArm(args.cpu_views or auto_cpu_views())
            arm.step()
            t0 = time.perf_counter()
            arm.step()
            dtc = time.perf_counter() - t0
            arm.close()
            line['cpu_baseline'] = arm.describe(arm.n_views * N_CH / dtc)
        if ws == 1 and not args.no_extra:
            line['extra'] = extra_measurements(dev, args.dtype)
        print(json.dumps(line), flush=True)
    if pg is not None:
        status = pg.status()
        pg.close()
        if status:
            raise SystemExit(f'bench.py: rank {rank} waited in vain for rank {status - 1} (peer flags)')
    if ws > 1:
        dist.barrier()
        dist.destroy_process_group()


def extra_measurements(dev, dtype):
    """Secondary device timings, reported beside the headline (never part of `value`): the 3-D (cone-beam) projector on
    two 48-view ranges of config C4 against the same bytes-per-ray roofline, and fan-beam FBP of one C2 sinogram."""
    import torch
    from dexct import synthetic as S
    from dexct.ops import Projector, count_segments, fbp_fan, make_scan
    from xtomosim import forward_project as FP
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:                                                     # noqa: BLE001
        pass
    hbm = float(peaks.get('hbm_gbs', 6650.0))
    td = torch.float32 if dtype == 'f32' else torch.float64
    esz = 4 if dtype == 'f32' else 8

    def timed(fn, reps):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    out = {}
    ct, ph, lo, hi = S.make_config('C2')
    scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, rotation=ct.rotation)
    pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz), device=dev)
    mu = torch.as_tensor(ph.mu_table(lo.E), dtype=td, device=dev)
    W = torch.as_tensor(np.stack([FP.spectral_weights(ct, lo)]), dtype=td, device=dev)
    _, logs, _ = pj.project(mu, W, dtype=dtype)
    sino = logs[0, :, 0, :].contiguous()
    t = timed(lambda: fbp_fan(scan, sino, 512, 51.2, 0.8, dtype=dtype), 10)
    out['fbp_c2'] = {'ms': t, 'workload': '1000 views x 900 ch -> 512 x 512, ramp 0.8', 'pixel_views_per_s': 512 * 512 * 1000 / (t * 1e-3)}
    out['prepare_c2_ms'] = timed(pj.prepare, 10)
    del pj
    n_views = 48
    ct, ph, lo, hi = S.make_config('C4', device=dev)            # labels rasterised on the device (12 s in numpy)
    scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, n_rows=ct.N_rows, h_iso=ct.h_iso, rotation=ct.rotation)
    pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz), device=dev)
    out['prepare_c4_ms'] = timed(pj.prepare, 3)
    mu = torch.as_tensor(ph.mu_table(lo.E), dtype=td, device=dev)
    W = torch.as_tensor(np.stack([FP.spectral_weights(ct, lo), FP.spectral_weights(ct, hi)]), dtype=td, device=dev)
    counts = torch.empty((2, n_views, ct.N_rows, ct.N_channels), dtype=td, device=dev)
    logs = torch.empty_like(counts)
    rays = n_views * ct.N_rows * ct.N_channels
    out['siddon_3d_c4'] = []
    for v0 in (0, 150):
        t = timed(lambda: pj.project(mu, W, view_begin=v0, view_count=n_views, dtype=dtype, counts=counts, logs=logs), 5)
        seg = count_segments(pj, v0, 4) / (4 * ct.N_rows * ct.N_channels)
        ach = rays * (seg + 2 * esz) / (t * 1e-3) / 1e9
        out['siddon_3d_c4'].append({'views': [v0, v0 + n_views], 'ms': t, 'rays_per_s': rays / (t * 1e-3), 'segments_per_ray': seg,
                                    'roofline': {'bound': 'hbm', 'achieved': ach, 'peak': hbm, 'unit': 'GB/s', 'frac': ach / hbm}})
    out['workload_c4'] = '512^3 phantom, 64 rows x 900 ch, 48 of 1200 views per range, 120 bins, 2 spectra fused'
    return out


def main():
    args = parse()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_b200(args)


if __name__ == '__main__':
    main()
