#!/usr/bin/env python3
"""Benchmark of the hot path: Siddon polyenergetic forward projection + Gauss-Newton two-basis decomposition.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config C5|C4|C2]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Headline workload (BASELINE.json configs[4], "C5" -- the largest configuration that fits one GPU): synthetic 1024^3
phantom, 128 detector rows x 1800 channels x 2400 views, 150 energy bins, both spectra in ONE traversal, followed by the GN
decomposition of every sinogram pixel (10 iterations, the fixed count configs[2] names).  One step = one whole scan:
552 960 000 rays -> 2 count + 2 log sinograms -> 552 960 000 decomposed pixel pairs.  Unit = rays (= pixels) per second
through projection + decomposition.  The FBP stage of configs[4] ("simulate -> decompose -> FBP") is timed beside it under
`extra.c5_pipeline`, the full 1200-view C4 scan under `extra.c4`, and round 1's headline (C2, weak scaling) under `extra.c2`.

N > 1 (STRONG scaling, as configs[4] says "at 1/2/4/8 GPUs"): the 2400 views are split over the ranks; rank r traces and
decomposes its own contiguous view range without communication, and every rank ends the step holding the whole view-major
sinograms (north_star: "assembled with all-gather"): through peer memory (CUDA IPC over NVLink) -- small shards ride inside
the decomposition kernel (`fused`), large ones go out on the copy engines beside it (`dma`) -- or NCCL (`--collective nccl`).

`value`   : device throughput, inputs resident in HBM, CUDA events around every step, max over ranks.
`e2e`     : the same work through the reference-facing Python entry points (get_sinos / get_basismat_sinos) with host
            numpy arrays in and out; every step uploads the phantom again and reads the results back.
`roofline`: the 3-D Siddon kernel against the bytes-per-ray HBM roofline of SURVEY.md section 8(d); `roofline_gn`: the GN
            kernel against the MUFU unit (tensor-core kernel: one ex2 per pixel, bin and iteration is what stays on the SM)
            or, with DEXCT_GN_TC=0, against the FP32 FMA pipe.
`cpu_baseline` / `--impl reference`: the CPU oracle (oracle/: C restatement of Siddon 1985 with OpenMP + the numpy
            restatement of the reference's optimize_sino_cpu, one process per core) on a bounded view sample of the SAME scan.
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

GN_ITERS = 10
MASK_THRESH = 0.95
METRIC = 'siddon_poly_gn_rays_per_s'
UNIT = 'rays/s'
HEADLINE = 'C5'

WORKLOADS = {
    'C5': 'C5: 1024x1024x1024 phantom (K=8 materials), 128 rows x 1800 ch x 2400 views, 150 energy bins, 2 spectra fused, '
          'GN 10 iters tissue/bone, air mask 0.95',
    'C4': 'C4: 512x512x512 phantom (K=8 materials), 64 rows x 900 ch x 1200 views, 120 energy bins, 2 spectra fused, '
          'GN 10 iters tissue/bone, air mask 0.95',
    'C2': 'C2: 512x512x1 phantom (K=8 materials), 900 ch x 1000 views per GPU, 120 energy bins, 2 spectra fused, '
          'GN 10 iters tissue/bone, air mask 0.95',
}
SCALING = {'C5': 'strong', 'C4': 'strong', 'C2': 'weak'}

# N > 1: how a finished shard reaches the peers' copies.  'fused': the decomposition kernel's own threads copy the count/log
# shard, a slice per Newton iteration, and its epilogue stores every basis pixel into all copies -- no extra launch, the
# choice for steps of ~1 ms (C2: profiles/r01h_scaling.txt).  'dma': the copy engines carry the count/log shard beside the
# decomposition, which runs in view chunks whose basis pixels follow on the copy engines as they finish -- no SM time, the
# choice for GB-sized shards (C5 at N = 8: the rider stretched the GN kernel from 52 to 72 ms, profiles/r01k_configs_n8.jsonl).
# 'auto' picks by shard size.  'sm': a separate store kernel on a side stream -- the choice for small shards since the
# tensor-core decomposition kernel, whose compute warps have no slack for the rider (C2 at N = 8: 0.588 against 0.639 ms).
# kernels the tensor-core decomposition launches beyond table / kernel / fix-up: gn_first_sums_kernel, and gn_compact_kernel when
# the air mask is on and the list of non-air pixels is not switched off
GN_EXTRA = (0 if os.environ.get('DEXCT_GN_TC', '1') == '0' else
            1 + (1 if os.environ.get('DEXCT_GN_COMPACT', '1') != '0' else 0))
PUSH_MODE = os.environ.get('DEXCT_PUSH', 'auto')
PUSH_CTAS = int(os.environ.get('DEXCT_PUSH_CTAS', '16'))
DMA_THRESHOLD = int(os.environ.get('DEXCT_DMA_THRESHOLD', str(64 << 20)))     # count/log shard bytes
GN_CHUNKS = int(os.environ.get('DEXCT_GN_CHUNKS', '4'))


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--config', default=HEADLINE, choices=sorted(WORKLOADS))
    ap.add_argument('--dtype', default='f32', choices=['f32', 'f64'])
    ap.add_argument('--cpu-views', type=int, default=0, help='views in the CPU sample (0 = auto)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-extra', action='store_true', help='skip the secondary measurements (C5 pipeline with FBP, C4, C2)')
    ap.add_argument('--collective', default='peer', choices=['peer', 'nccl'],
                    help='N > 1: assemble the sinograms by stores into peer memory (default) or by NCCL all-gathers')
    return ap.parse_args()


def config_dict(name, ws):
    """`config` of the JSON line -- the same for both arms, so that the driver compares like with like."""
    from dexct import synthetic as S
    c = S.CONFIGS[name]
    views = c['n_views'] * (ws if SCALING[name] == 'weak' else 1)
    return {'workload': WORKLOADS[name], 'views_total': views, 'rays_per_step': views * c['n_rows'] * c['n_ch'],
            'gn_iters': GN_ITERS, 'parallelism': f'view-sharded x{ws}',
            'l2': 'flushed between timed steps (256 MiB write)' if name == 'C2'
                  else 'inputs and outputs of a step (GBs) are far larger than the 126 MB L2'}


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:                                                     # noqa: BLE001
        return {}


# ----------------------------------------------------------------------------------------------------------------
# CPU arm (oracle).  Used for `cpu_baseline` and `--impl reference`.

def _gn_job(args):
    from oracle import gn_oracle
    g, i0, mus, n_iters = args                                            # g [2, 1, n_ch]: one detector line
    nB = g.shape[2]
    i0_t = np.tile(i0, (1, nB)).reshape(2, nB, i0.shape[1])
    return gn_oracle.optimize_sino_cpu(g, None, i0_t, mus, n_iters)


def _worker_init():
    os.environ['OMP_NUM_THREADS'] = '1'
    try:                                                                   # numpy's BLAS/OpenMP pools, if already loaded
        import threadpoolctl
        threadpoolctl.threadpool_limits(1)
    except Exception:                                                     # noqa: BLE001
        pass


class CpuArm:
    """The oracle's CPU implementation of one bounded sample of the workload: `n_views` views spread evenly over the
    rotation, all detector rows and channels; Siddon + detection in C with OpenMP over all host threads, GN in numpy with
    one worker process per core (one detector line per job)."""

    def __init__(self, name, n_views):
        import multiprocessing as mp
        self.cores = os.cpu_count() or 1
        # a launcher may have exported OMP_NUM_THREADS=1 (torchrun does): the C oracle would then trace on ONE thread and
        # every ratio against it would be inflated (round 1, SCALE N >= 2).  Ask for all cores, and check below.
        os.environ['OMP_NUM_THREADS'] = str(self.cores)
        import oracle
        from oracle import gn_oracle
        from dexct import synthetic as S
        from xtomosim.forward_project import spectral_weights
        oracle.build()
        oracle.set_num_threads(self.cores)
        self.threads = oracle.num_threads()
        if self.threads < self.cores:
            print(f'bench.py: WARNING: the C oracle runs on {self.threads} of {self.cores} cores', file=sys.stderr)
        self.name = name
        c = S.CONFIGS[name]
        self.full_views = c['n_views']
        dev = None
        try:                                     # input generation only: the 1024^3 labels take minutes in numpy
            import torch
            dev = 'cuda' if torch.cuda.is_available() else ('cpu' if c['nz'] > 1 else None)
        except Exception:                                                 # noqa: BLE001
            pass
        ct, ph, lo, hi = S.make_config(name, device=dev)
        labels = ph.M.cpu().numpy() if dev is not None else ph.M
        tabs = oracle.fan_tables(ct.N_channels, ct.N_proj, ct.gamma_fan, ct.rotation, ct.beta0, ct.N_rows, ct.h_iso, ct.SID, ct.SDD)
        pick = np.linspace(0, ct.N_proj, n_views, endpoint=False).astype(int)
        tabs = {k: (v[pick] if k.endswith('beta') else v) for k, v in tabs.items()}
        self.scan = oracle.Scan(labels, (ph.dx, ph.dy, ph.dz), ct.SID, ct.SDD, tabs)
        self.mu = ph.mu_table(lo.E)
        self.W = np.stack([spectral_weights(ct, lo), spectral_weights(ct, hi)])
        import matdecomp
        _, self.i0, self.mus = gn_oracle.gn_model_tables(ct, lo, hi, matdecomp.xc.mixatten)
        self.n_views = n_views
        self.n_rows, self.n_ch = ct.N_rows, ct.N_channels
        self.rays = n_views * ct.N_rows * ct.N_channels
        self.pool = mp.get_context('spawn').Pool(self.cores, initializer=_worker_init) if self.cores > 1 else None
        if self.pool:                                                      # workers imported and warm before the clock starts
            self.pool.map(_gn_job, [(np.ones((2, 1, 4)), self.i0, self.mus, 1)] * self.cores, chunksize=1)
        self.seg_per_ray = None

    def step(self):
        counts, _, seg = self.scan.project_poly(self.mu, self.W, want_log=True)     # [2, views, rows, ch]
        self.seg_per_ray = seg / self.rays
        g = counts.reshape(2, -1, self.n_ch)                                          # [2, views * rows, ch]
        mask = g[0] >= MASK_THRESH * g[0].max()
        jobs = [(g[:, ln:ln + 1], self.i0, self.mus, GN_ITERS) for ln in range(g.shape[1])]
        outs = self.pool.map(_gn_job, jobs, chunksize=max(1, len(jobs) // (8 * self.cores))) if self.pool \
            else [_gn_job(j) for j in jobs]
        a = np.concatenate(outs, axis=0)
        a[mask] = 0
        return a

    def close(self):
        if self.pool:
            self.pool.close()
            self.pool.join()

    def describe(self, value):
        return {'value': value, 'unit': UNIT, 'cores': self.cores, 'threads_used': self.threads, 'kind': 'port',
                'sample': f'{self.n_views} of {self.full_views} views of the {self.name} scan, evenly spaced, all rows and channels '
                          f'({self.rays} rays): oracle/siddon_oracle.c float64 Siddon 1985 + Beer-Lambert (OpenMP, {self.threads} '
                          f'threads) then oracle/gn_oracle.py optimize_sino_cpu float64 numpy ({GN_ITERS} iters, {self.cores} worker '
                          f'processes, one detector line per job)'}


def auto_cpu_views(name):
    """A sample that keeps the host cores busy for roughly 10-20 s: the numpy Newton loop costs ~80 ms per 900 pixels x 120
    bins x 10 iterations and core, the C tracer a third of that."""
    from dexct import synthetic as S
    c = S.CONFIGS[name]
    cores = os.cpu_count() or 1
    per_view_core_s = 0.080 * (c['n_rows'] * c['n_ch'] / 900.0) * (c['n_bins'] / 120.0)
    return int(min(c['n_views'], max(1, round(10.0 * cores / per_view_core_s))))


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    ws = int(os.environ.get('WORLD_SIZE', '1'))
    arm = CpuArm(args.config, args.cpu_views or auto_cpu_views(args.config))
    n_warm = min(args.warmup, 1)
    for _ in range(n_warm):
        arm.step()
    t = []
    for _ in range(max(1, min(args.steps, 3))):
        t0 = time.perf_counter()
        arm.step()
        t.append(time.perf_counter() - t0)
    arm.close()
    steps = len(t)
    ms = 1e3 * sum(t) / steps
    value = arm.rays / (ms * 1e-3)
    line = {'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': steps,
            'warmup': n_warm, 'ms_per_step': ms, 'higher_is_better': True, 'scaling': SCALING[args.config],
            'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': config_dict(args.config, ws),
            'step': 'bounded sample of the config: see cpu_baseline.sample',
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

class ScanBench:
    """One configuration on this rank's view range: resident inputs, output buffers (symmetric peer allocations at N > 1),
    and `step()` = projection of both spectra + decomposition + assembly of the sinograms on every rank."""

    def __init__(self, name, dev, rank, ws, dtype='f32', collective='peer', with_fbp=False, assemble=True):
        import torch
        import torch.distributed as dist
        from dexct import dist as ddist
        from dexct import synthetic as S
        from dexct.ops import Projector, dtype_code, make_scan
        from xtomosim import forward_project as FP
        import matdecomp
        self.torch, self.dist = torch, dist
        self.name, self.dev, self.rank, self.ws = name, dev, rank, ws
        self.scaling = SCALING[name]
        self.code = dtype_code(dtype)
        self.td = td = torch.float32 if dtype == 'f32' else torch.float64
        self.esz = esz = 4 if dtype == 'f32' else 8
        c = S.CONFIGS[name]
        self.n_views = c['n_views'] * (ws if self.scaling == 'weak' else 1)
        on_dev = c['nz'] > 1                                     # large volumes are rasterised on the device
        ct, ph, lo, hi = S.make_config(name, n_views=self.n_views, device=dev if on_dev else None)
        self.ct, self.ph, self.lo, self.hi = ct, ph, lo, hi
        self.v0, self.nv = ddist.shard_range(self.n_views, rank, ws)
        self.R, self.C = ct.N_rows, ct.N_channels
        self.scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, n_rows=ct.N_rows, h_iso=ct.h_iso,
                              rotation=ct.rotation)
        self.pj = Projector(self.scan, ph.M, (ph.dx, ph.dy, ph.dz), device=dev)
        self.mu = torch.as_tensor(ph.mu_table(lo.E), dtype=td, device=dev)
        self.W = torch.as_tensor(np.stack([FP.spectral_weights(ct, lo), FP.spectral_weights(ct, hi)]), dtype=td, device=dev)
        _, i0_h, mus_h = matdecomp.spectral_tables(ct, lo, hi)
        self.i0 = torch.as_tensor(i0_h, dtype=td, device=dev)
        self.mus = torch.as_tensor(mus_h, dtype=td, device=dev)
        self.nE = int(self.i0.shape[1])
        self.rays_local = self.nv * self.R * self.C
        self.rays_total = self.n_views * self.R * self.C
        nv, R, Cn = self.nv, self.R, self.C
        even = self.n_views % ws == 0
        self.pg = None
        self.collective = 'none'
        self.push_mode = None
        self.launches = 0
        slot_s, slot_b = 4 * nv * R * Cn * esz, nv * R * Cn * 2 * esz
        self.slot_s, self.slot_b = slot_s, slot_b
        if ws > 1 and collective == 'peer' and even:
            from dexct.dist import all_agree
            from dexct.peer import PeerGroup
            sbuf = None
            try:                                                          # (a failure on one rank raises on all: peer.py)
                self.pg = PeerGroup()
                sbuf, bbuf = self.pg.alloc(ws * slot_s), self.pg.alloc(ws * slot_b)
            except Exception as exc:                                      # noqa: BLE001
                print(f'bench.py: rank {rank}: peer memory unavailable ({exc})', file=sys.stderr)
                sbuf = None
            if not all_agree(self.pg is not None and sbuf is not None):   # belt and braces: all ranks take the same path
                self.pg = None        # (not closed: closing is collective, and the ranks may not agree on what exists)
                if rank == 0:
                    print('bench.py: using NCCL all-gathers on every rank', file=sys.stderr)
            else:
                self.collective = 'peer'
                self.sbuf, self.bbuf = sbuf, bbuf
        if ws > 1 and self.pg is None:
            self.collective = 'nccl'
        if self.pg is not None:
            self.g_sino = self.sbuf.view(td, (ws, 4, nv, R, Cn))             # [rank][counts lo, hi, logs lo, hi]
            self.g_basis = self.bbuf.view(td, (ws, nv, R, Cn, 2))
            self.counts, self.logs, self.basis = self.g_sino[rank, 0:2], self.g_sino[rank, 2:4], self.g_basis[rank]
            self.mirrors = self.bbuf.table(rank * slot_b, exclude_self=True)
            self.push_table = self.sbuf.table(rank * slot_s, exclude_self=True)
            mode = PUSH_MODE
            if mode == 'auto':
                # small shards: with the tensor-core decomposition kernel a store kernel on a side stream costs less than the
                # copy riding inside the kernel (its compute warps have no slack to hide the peer stores in: C2 at N = 8
                # 0.588 ms against 0.639, copy engines 1.09); the FP32-pipe kernel hid the rider better (round 1)
                small = 'sm' if os.environ.get('DEXCT_GN_TC', '1') != '0' else 'fused'
                mode = 'dma' if slot_s > DMA_THRESHOLD else small
            self.push_mode = mode
            self.side = [torch.cuda.Stream(device=dev) for _ in range(4)]
        else:
            self.counts = torch.empty((2, nv, R, Cn), dtype=td, device=dev)
            self.logs = torch.empty_like(self.counts)
            self.basis = torch.empty((nv, R, Cn, 2), dtype=td, device=dev)
            if ws > 1 and even and assemble:
                self.g_counts = torch.empty((ws, 2, nv, R, Cn), dtype=td, device=dev)
                self.g_logs = torch.empty_like(self.g_counts)
                self.g_basis = torch.empty((ws, nv, R, Cn, 2), dtype=td, device=dev)
            self.even = even and assemble
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if name == 'C2' else None
        self.with_fbp = with_fbp
        if with_fbp:
            self.n_matrix, self.fov, self.ramp = ph.Nx, ph.Nx * ph.dx, 0.8
            self.images = [torch.empty((R, self.n_matrix, self.n_matrix), dtype=td, device=dev) for _ in range(4)]
            if ws > 1 and R % ws == 0:
                self.slabs = [torch.empty((R // ws, self.n_matrix, self.n_matrix), dtype=td, device=dev) for _ in range(4)]

    # ---- peer-memory pushes on the copy engines: one DMA per destination, destinations taken in ring order from this
    # rank's neighbour (so that the ranks do not all write to rank 0 first), spread over four side streams
    def _dma_push(self, buf, offset_bytes, src):
        torch = self.torch
        main = torch.cuda.current_stream()
        nb = src.numel() * src.element_size()
        import ctypes as C
        from dexct import _native as N
        for k in range(1, self.ws):
            r = (self.rank + k) % self.ws
            st = self.side[(k - 1) % len(self.side)]
            st.wait_stream(main)
            with torch.cuda.stream(st):
                tab = (C.c_void_p * 1)(buf.ptrs[r] + int(offset_bytes))
                N.check(N.lib().dexct_peer_push_dma(N.ptr(src), nb, tab, 1, -1, N.stream_ptr()))

    def _join_sides(self):
        main = self.torch.cuda.current_stream()
        for st in self.side:
            main.wait_stream(st)

    def step(self, t=None):
        """t: four CUDA events (start, after projection, after decomposition, end) or None."""
        torch, dist = self.torch, self.dist
        from dexct.ops import gn_air_key, gn_decompose
        rank, ws, pg, code = self.rank, self.ws, self.pg, self.code
        counts, logs, basis = self.counts, self.logs, self.basis
        if t is not None:
            t[0].record()
        self.pj.project(self.mu, self.W, view_begin=self.v0, view_count=self.nv, dtype=code, want_log=True, counts=counts, logs=logs)
        n = 1                                                             # siddon kernel
        if t is not None:
            t[1].record()
        if pg is not None and self.push_mode == 'dma':
            # count/log shard: seven DMAs beside the decomposition.  Decomposition: GN_CHUNKS view ranges against the air
            # threshold of the whole sinogram; each finished range of basis pixels follows on the copy engines.
            self._dma_push(self.sbuf, rank * self.slot_s, self.g_sino[rank])
            key = gn_air_key(counts[0], dtype=code, reduce_max_over_ranks=True, peers=pg)
            n += 2                                                        # gn_max + peer_max
            nch = max(1, min(GN_CHUNKS, self.nv))
            bounds = [self.nv * k // nch for k in range(nch + 1)]
            per_view = self.R * self.C * 2 * self.esz
            for a, b in zip(bounds[:-1], bounds[1:]):
                gn_decompose(counts[:, a:b].reshape(2, -1), self.i0, self.mus, GN_ITERS, mask_thresh=MASK_THRESH, dtype=code,
                             out=basis[a:b].view(-1, 2), air_key=key)
                n += 4 + GN_EXTRA                                         # strided gather copy, gn_table, gn_kernel, gn_fixup (+ GN_EXTRA)
                self._dma_push(self.bbuf, rank * self.slot_b + a * per_view, basis[a:b])
            if t is not None:
                t[2].record()
            self._join_sides()
            pg.barrier()                                                  # every rank's stores have landed in every copy
            n += 1
        elif pg is not None and self.push_mode == 'fused':
            # one kernel: Newton iterations + the copy of the count/log shard into the peers, a slice per Newton iteration
            gn_decompose(counts.reshape(2, -1), self.i0, self.mus, GN_ITERS, mask_thresh=MASK_THRESH, dtype=code, out=basis.view(-1, 2),
                         reduce_max_over_ranks=True, peers=pg, mirrors=self.mirrors, push=(self.g_sino[rank], self.push_table))
            if t is not None:
                t[2].record()
            pg.barrier()
            n += 6 + GN_EXTRA                                             # gn_max, peer_max, gn_table, gn_kernel, gn_fixup, barrier
        elif pg is not None:                                              # 'sm': store kernel on a side stream
            main = torch.cuda.current_stream()
            self.side[0].wait_stream(main)
            with torch.cuda.stream(self.side[0]):
                pg.push(self.g_sino[rank], self.sbuf, rank * self.slot_s, dma=False, max_ctas=PUSH_CTAS)
            gn_decompose(counts.reshape(2, -1), self.i0, self.mus, GN_ITERS, mask_thresh=MASK_THRESH, dtype=code, out=basis.view(-1, 2),
                         reduce_max_over_ranks=True, peers=pg, mirrors=self.mirrors)
            if t is not None:
                t[2].record()
            main.wait_stream(self.side[0])
            pg.barrier()
            n += 7 + GN_EXTRA
        else:
            pending = []
            if ws > 1 and self.even:
                # the sinogram gathers ride NCCL's stream behind the projection and overlap the decomposition, which only
                # needs the local view shard
                pending.append(dist.all_gather_into_tensor(self.g_counts, counts, async_op=True))
                pending.append(dist.all_gather_into_tensor(self.g_logs, logs, async_op=True))
            gn_decompose(counts.view(2, -1), self.i0, self.mus, GN_ITERS, mask_thresh=MASK_THRESH, dtype=code, out=basis.view(-1, 2),
                         reduce_max_over_ranks=ws > 1)
            n += 4 + GN_EXTRA                                             # gn_max, gn_table, gn_kernel, gn_fixup
            if t is not None:
                t[2].record()
            if ws > 1 and self.even:
                pending.append(dist.all_gather_into_tensor(self.g_basis, basis, async_op=True))
                for w in pending:
                    w.wait()
        if t is not None:
            t[3].record()
        self.launches += n

    def pipeline_step(self, marks):
        """configs[4]'s whole pipeline (/root/reference/main.py:109-168) on this rank's view range: projection, FBP of both
        log sinograms, decomposition, FBP of both basis sinograms.  View-sharded FBP needs no assembled sinogram, so nothing
        is replicated; the partial volumes are summed into row slabs (one reduce-scatter per volume, each rank ends with
        R / N detector rows of every image).  Slice-by-slice fan-beam FBP: a throughput figure for multi-row scans, not a
        cone-beam (FDK) reconstruction off the mid-plane."""
        torch, dist = self.torch, self.dist
        from dexct.ops import fbp_fan, gn_decompose
        ev = lambda: torch.cuda.Event(enable_timing=True)                 # noqa: E731
        m = [ev() for _ in range(6)]
        m[0].record()
        self.pj.project(self.mu, self.W, view_begin=self.v0, view_count=self.nv, dtype=self.code, counts=self.counts, logs=self.logs)
        m[1].record()
        for s in range(2):
            fbp_fan(self.scan, self.logs[s], self.n_matrix, self.fov, self.ramp, view_begin=self.v0, dtype=self.code, image=self.images[s])
        m[2].record()
        gn_decompose(self.counts.reshape(2, -1), self.i0, self.mus, GN_ITERS, mask_thresh=MASK_THRESH, dtype=self.code,
                     out=self.basis.view(-1, 2), reduce_max_over_ranks=self.ws > 1, peers=self.pg)
        m[3].record()
        for q in range(2):
            fbp_fan(self.scan, self.basis[..., q].contiguous(), self.n_matrix, self.fov, self.ramp, view_begin=self.v0, dtype=self.code,
                    image=self.images[2 + q])
        m[4].record()
        if self.ws > 1:
            for k, im in enumerate(self.images):
                if hasattr(self, 'slabs'):
                    dist.reduce_scatter_tensor(self.slabs[k], im)
                else:
                    dist.all_reduce(im)
        m[5].record()
        marks.append(m)

    def barrier(self):
        if self.ws > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def run(self, steps, warmup):
        """Timed steps; returns {'total', 'siddon', 'gn'} in ms per step, max over ranks."""
        torch, dist = self.torch, self.dist
        for _ in range(warmup):
            if self.flush is not None:
                self.flush.zero_()
            self.step()
        self.barrier()
        self.launches = 0
        marks = []
        self.barrier()
        for _ in range(steps):
            if self.flush is not None:
                self.flush.zero_()
            t = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            self.step(t)
            marks.append(t)
        self.barrier()
        tot = sum(t[0].elapsed_time(t[3]) for t in marks) / steps
        t_sid = sum(t[0].elapsed_time(t[1]) for t in marks) / steps
        t_gn = sum(t[1].elapsed_time(t[2]) for t in marks) / steps
        tt = torch.tensor([tot, t_sid, t_gn], dtype=torch.float64, device=self.dev)
        if self.ws > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        tot, t_sid, t_gn = [float(x) for x in tt.cpu()]
        return {'total': tot, 'siddon': t_sid, 'gn': t_gn}

    def run_pipeline(self, reps):
        torch, dist = self.torch, self.dist
        marks = []
        self.pipeline_step(marks)
        marks.clear()
        self.barrier()
        for _ in range(reps):
            self.pipeline_step(marks)
        self.barrier()
        names = ['siddon_poly', 'fbp_2_log_sinograms', 'gn', 'fbp_2_basis_sinograms', 'sum_partial_volumes']
        vals = [sum(m[k].elapsed_time(m[k + 1]) for m in marks) / reps for k in range(5)]
        vals.append(sum(m[0].elapsed_time(m[5]) for m in marks) / reps)
        tt = torch.tensor(vals, dtype=torch.float64, device=self.dev)
        if self.ws > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        out = dict(zip(names + ['total'], [float(x) for x in tt.cpu()]))
        return out

    def verify_peer(self):
        """The peer-assembled buffers against NCCL's all-gather of the same shards (outside the timed region)."""
        if self.pg is None:
            return None
        torch, dist = self.torch, self.dist
        ok = True
        for whole, mine in ((self.g_sino, self.g_sino[self.rank]), (self.g_basis, self.g_basis[self.rank])):
            chk = torch.empty_like(whole)
            dist.all_gather_into_tensor(chk, mine.contiguous())
            ok = ok and bool(torch.equal(chk, whole))
            del chk
        flag = torch.tensor([int(ok)], device=self.dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        return bool(int(flag[0]))

    def segments_per_ray(self, n_sample=4):
        """Mean voxel segments per ray on a sample of this rank's views, averaged over ranks (the unit of the roofline)."""
        from dexct.ops import count_segments
        torch, dist = self.torch, self.dist
        sample = list(range(self.v0, self.v0 + self.nv, max(1, self.nv // n_sample)))[:n_sample]
        seg = sum(count_segments(self.pj, v, 1) for v in sample) / (len(sample) * self.R * self.C)
        st = torch.tensor([seg], dtype=torch.float64, device=self.dev)
        if self.ws > 1:
            dist.all_reduce(st)
            st /= self.ws
        return float(st[0])

    def rooflines(self, t, clocks, peaks):
        """Bytes-per-ray roofline of the projection kernel and FMA-pipe roofline of the decomposition kernel (whole job:
        peaks x N)."""
        ws = self.ws
        hbm = float(peaks.get('hbm_gbs', 6650.0))
        seg = self.segments_per_ray()
        bytes_per_ray = seg * 1 + 2 * self.esz
        ach = self.rays_total * bytes_per_ray / (t['siddon'] * 1e-3) / 1e9
        kern = 'siddon_skip3_kernel' if self.name != 'C2' else 'siddon_skip_kernel'
        traffic, traffic_src = None, None
        try:                                                              # measured under ncu, committed with its source
            tr = json.load(open(os.path.join(ROOT, 'profiles', 'ncu_traffic.json')))[f'{kern}:{self.name}']
            traffic, traffic_src = tr['dram_bytes_read'] + tr['dram_bytes_write'], tr['source']
        except Exception:                                                 # noqa: BLE001
            pass
        sm_mhz = (clocks or {}).get('sm_mhz') or peaks.get('sm_max_mhz', 1965.0)
        fma_peak = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12 * ws               # TFLOP/s fp32 at the clock seen during the run
        gn_flops = 28.0 * self.nE * GN_ITERS * self.rays_total
        gn_ach = gn_flops / (t['gn'] * 1e-3) / 1e12
        roof = {'bound': 'hbm', 'achieved': ach, 'peak': hbm * ws, 'unit': 'GB/s', 'frac': ach / (hbm * ws), 'traffic': traffic,
                'traffic_source': traffic_src, 'algorithmic_bytes_per_launch': self.rays_local * bytes_per_ray,
                'kernel': kern, 'bytes_per_ray': bytes_per_ray, 'segments_per_ray': seg, 'launch_ms': t['siddon'],
                'peak_source': ('MEASURED_PEAKS.json hbm_gbs' if peaks else 'fallback') + (f' x {ws} GPUs' if ws > 1 else '')}
        if os.environ.get('DEXCT_GN_TC', '1') != '0' and self.nE <= 448:
            # tensor-core kernel: the twelve moment sums run on tcgen05 (3xTF32), what is left on the SM is one MUFU.EX2 per
            # (pixel, bin, iteration) -- the unit that bounds it: 15.9 ex2 per clock and SM measured (tools/microbench/xu_rate.cu,
            # profiles/r02m_xu_rate.txt).  The FP32-FMA figure of the earlier rounds is kept beside it for continuity: it can
            # exceed 1 now, the FMAs are no longer on that pipe.
            ex2 = float(self.nE) * GN_ITERS * self.rays_total
            # share of the 128-pixel tiles of this rank's shard that the kernel does not decompose because all their pixels are
            # air (first measurement >= MASK_THRESH x its maximum: get_basismat_sinos zeroes them, matdecomp.py:195-196,204-205)
            air_tiles = None
            try:
                torch = self.torch
                g1 = self.counts[0].reshape(-1)
                mx = g1.max()
                if ws > 1:
                    self.dist.all_reduce(mx, op=self.dist.ReduceOp.MAX)
                pad = (-g1.numel()) % 128
                air = g1 >= MASK_THRESH * mx
                if pad:
                    air = torch.cat([air, torch.ones(pad, dtype=torch.bool, device=air.device)])
                air_tiles = float(air.view(-1, 128).all(dim=1).float().mean()) if MASK_THRESH > 0 else 0.0
                air_px = float(air[:g1.numel()].float().mean()) if MASK_THRESH > 0 else 0.0
            except Exception:                                             # noqa: BLE001
                air_px = None
            compact = os.environ.get('DEXCT_GN_COMPACT', '1') != '0'     # air pixels zeroed up front, the others listed
            evaluated = (max(GN_ITERS - 1, 0) / GN_ITERS) * (1.0 - ((air_px if compact else air_tiles) or 0.0))
            xu_peak = 15.9 * 148 * sm_mhz * 1e6 / 1e9 * ws                # Gex2/s
            xu_ach = ex2 / (t['gn'] * 1e-3) / 1e9
            roof_gn = {'bound': 'mufu', 'achieved': xu_ach, 'peak': xu_peak, 'unit': 'Gex2/s', 'frac': xu_ach / xu_peak,
                       'kernel': 'gn_tc_kernel', 'ex2_per_pixel': float(self.nE) * GN_ITERS, 'launch_ms': t['gn'], 'sm_mhz': sm_mhz,
                       'air_pixel_fraction': air_px, 'air_tile_fraction': air_tiles, 'air_pixels_compacted_away': compact,
                       'evaluated_share_of_algorithmic': evaluated,
                       'frac_of_evaluated': xu_ach / xu_peak * evaluated,
                       'note': 'achieved counts the exponentials of the algorithm (nE x iterations for every pixel: what the reference '
                               'evaluates).  The kernel evaluates fewer for the same output: the sums of the first iteration are the '
                               'same for every pixel and are computed once, and air pixels -- zeroed by the air mask whatever the iterations '
                               'give -- are zeroed up front and left out of the list of pixels the kernel iterates (without the list: '
                               'tiles of 128 pixels that are all air are skipped).  frac_of_evaluated = the MUFU work actually done '
                               'over the peak',
                       'peak_source': '15.9 MUFU.EX2 per clock and SM (measured, profiles/r02m_xu_rate.txt) x 148 SMs x sm_mhz',
                       'fp32_fma_equivalent': {'achieved': gn_ach, 'peak': fma_peak, 'unit': 'TFLOP/s', 'frac': gn_ach / fma_peak,
                                               'flop_per_pixel': 28.0 * self.nE * GN_ITERS}}
        else:
            roof_gn = {'bound': 'fp32_fma', 'achieved': gn_ach, 'peak': fma_peak, 'unit': 'TFLOP/s', 'frac': gn_ach / fma_peak,
                       'kernel': 'gn_kernel', 'flop_per_pixel': 28.0 * self.nE * GN_ITERS, 'launch_ms': t['gn'], 'sm_mhz': sm_mhz}
        return roof, roof_gn

    def collective_text(self):
        if self.collective == 'none':
            return 'none'
        if self.collective == 'nccl':
            return 'NCCL: all-gather of counts/logs (overlapping the decomposition) and of the basis sinograms + 8-byte MAX all-reduce'
        common = '; air maximum and end-of-step barrier through flags in peer memory (CUDA IPC over NVLink); no library collective in the step'
        if self.push_mode == 'dma':
            return (f'peer memory, copy engines: count/log shard ({self.slot_s >> 20} MiB per rank) sent by {self.ws - 1} DMAs beside the '
                    f'decomposition, which runs in {GN_CHUNKS} view chunks whose basis pixels follow by DMA as they finish' + common)
        if self.push_mode == 'fused':
            return ('peer memory, fused: the decomposition kernel copies the count/log shard into the peers\' copies (a slice per '
                    'Newton iteration) and its epilogue stores each basis pixel into all copies' + common)
        return 'peer memory, store kernel on a side stream + mirrored GN epilogue' + common

    def close(self):
        torch = self.torch
        status = 0
        if self.pg is not None:
            status = self.pg.status()
            for k in ('counts', 'logs', 'basis', 'g_sino', 'g_basis'):
                if hasattr(self, k):
                    delattr(self, k)
            self.pg.close()
            self.pg = None
        for k in list(self.__dict__):
            if k not in ('torch', 'dist', 'rank', 'ws', 'dev', 'name'):
                delattr(self, k)
        torch.cuda.empty_cache()
        if status:
            raise SystemExit(f'bench.py: rank {self.rank} waited in vain for rank {status - 1} (peer flags)')


def e2e_leg(name, dev, rank, ws, dtype, steps):
    """The same scan through the reference-facing entry points, host numpy arrays in and out (float64, like the
    reference's).  Returns the `e2e` object, or a note when the host cannot hold the arrays."""
    import torch
    import torch.distributed as dist
    from dexct import config as cfg
    from dexct import synthetic as S
    from xtomosim import forward_project as FP
    import matdecomp
    c = S.CONFIGS[name]
    n_views = c['n_views'] * (ws if SCALING[name] == 'weak' else 1)
    rays = n_views * c['n_rows'] * c['n_ch']
    # float64 host arrays alive at once on this node: 4 sinograms + 2 basis sinograms (+ the same again while the next step
    # runs before the old arrays are dropped), phantom
    need = 2 * 6 * rays * 8 + c['n'] * c['n'] * c['nz'] * ws
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:                                                     # noqa: BLE001
        avail = None
    if avail is not None and need > 0.8 * avail:
        return None, f'host has {avail / 1e9:.0f} GB available, the float64 arrays of {name} need {need / 1e9:.0f} GB'
    cfg.dtype = dtype
    # N > 1: every rank keeps the host arrays of ITS OWN view range (config.gather_views off) -- the data-parallel form in
    # which each rank writes its part of the output; with the default (every rank returns the whole arrays, like the
    # single-process reference) each rank would move the whole scan over its PCIe link and the host side could not scale.
    # The air-mask maximum is still reduced over all ranks.
    cfg.gather_views = ws == 1
    on_dev = c['nz'] > 1
    ct, ph, lo, hi = S.make_config(name, n_views=n_views, device=dev if on_dev else None)
    if on_dev:                                   # the numpy-facing entry points take the phantom as a HOST array (page-locked)
        host = torch.empty(ph.M.shape, dtype=torch.uint8, pin_memory=True)
        host.copy_(ph.M)
        torch.cuda.synchronize()
        ph.M = host.numpy()
    label_bytes = int(np.prod(ph.M.shape))

    def step():
        FP.clear_cache()                                                  # the phantom is uploaded and re-laid-out every step
        (raw_lo, log_lo), (raw_hi, log_hi) = FP.get_sinos(ct, ph, [lo, hi])
        m1, m2 = matdecomp.get_basismat_sinos(ct, raw_lo, raw_hi, lo, hi, n_iters=GN_ITERS, mask_thresh=MASK_THRESH)
        return raw_lo, m1

    def barrier():
        if ws > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(2):
        step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    barrier()
    dt = torch.tensor([(time.perf_counter() - t0) / steps], dtype=torch.float64, device=dev)
    if ws > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    FP.clear_cache()

    # the call sequence of /root/reference/main.py:109-120,153: one get_sino per spectrum (the second is served from the path
    # lengths the first one stored), then get_basismat_sinos
    def step_main_py():
        FP.clear_cache()
        raw_lo, log_lo = FP.get_sino(ct, ph, lo)
        raw_hi, log_hi = FP.get_sino(ct, ph, hi)
        m1, m2 = matdecomp.get_basismat_sinos(ct, raw_lo, raw_hi, lo, hi, n_iters=GN_ITERS, mask_thresh=MASK_THRESH)
        return raw_lo, m1

    two = None
    try:
        step_main_py()
        barrier()
        t0 = time.perf_counter()
        for _ in range(max(1, steps - 1)):
            step_main_py()
        barrier()
        dt2 = torch.tensor([(time.perf_counter() - t0) / max(1, steps - 1)], dtype=torch.float64, device=dev)
        if ws > 1:
            dist.all_reduce(dt2, op=dist.ReduceOp.MAX)
        two = {'value': rays / float(dt2[0]), 'unit': UNIT, 'ms_per_step': 1e3 * float(dt2[0]), 'steps': max(1, steps - 1),
               'api': 'get_sino(spec1), get_sino(spec2) [second call from the stored path lengths], get_basismat_sinos -- the '
                      'call sequence of the reference main.py:109-120,153'}
    except Exception as ex:                                               # noqa: BLE001
        two = {'unavailable': f'{type(ex).__name__}: {ex}'[:200]}
    FP.clear_cache()
    # up: labels + the two float64 count sinograms handed to get_basismat_sinos (they sit in page-locked memory: get_sinos
    # returned them); down: 2 x (raw, log) + 2 basis sinograms, float64 like the reference's arrays; summed over ranks; the
    # KB-sized tables are uploaded once per content
    h2d = ws * label_bytes + 2 * rays * 8
    d2h = 6 * rays * 8
    return {'value': rays / float(dt[0]), 'unit': UNIT, 'h2d_bytes_per_step': int(h2d), 'd2h_bytes_per_step': int(d2h),
            'ms_per_step': 1e3 * float(dt[0]), 'steps': steps, 'config': name, 'main_py_call_sequence': two,
            'api': 'xtomosim.forward_project.get_sinos + matdecomp.get_basismat_sinos (float64 numpy in/out through page-locked '
                   'buffers, in view chunks whose host copies overlap the kernels of the next ones; phantom re-uploaded and '
                   're-laid-out every step'
                   + (')' if ws == 1 else '; every rank keeps the host arrays of its own view range, bytes summed over ranks)')}, None


def run_b200(args):
    import torch
    import torch.distributed as dist

    from dexct import _native as N

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
    peaks = load_peaks()
    name = args.config
    warmup = max(args.warmup, 3)

    # ---- headline: device-resident steps
    sb = ScanBench(name, dev, rank, ws, dtype=args.dtype, collective=args.collective)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    for _ in range(warmup):                       # (warm-up outside the clock sampler's window)
        sb.step()
    sb.barrier()
    if sampler:
        sampler.start()
    t = sb.run(args.steps, 0)
    clocks = sampler.finish() if sampler else None
    launches = sb.launches
    peer_ok = sb.verify_peer()
    roof, roof_gn = sb.rooflines(t, clocks, peaks)
    value = sb.rays_total / (t['total'] * 1e-3)
    line = None
    if rank == 0:
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': ws, 'steps': args.steps, 'warmup': warmup,
            'ms_per_step': t['total'], 'higher_is_better': True, 'scaling': sb.scaling, 'vs_baseline': None, 'dtype': args.dtype,
            'data': 'synthetic', 'config': config_dict(name, ws),
            'assembly': sb.collective_text(),
            'kernel_ms': {'siddon_poly': t['siddon'], 'gn': t['gn']},
            **({'peer_buffers_equal_nccl_all_gather': peer_ok} if peer_ok is not None else {}),
            'siddon_rays_per_s': sb.rays_total / (t['siddon'] * 1e-3), 'gn_pixels_per_s': sb.rays_total / (t['gn'] * 1e-3),
            'roofline': roof, 'roofline_gn': roof_gn,
            'gpu_launches': launches, 'gpu_launches_note': 'kernels and device copies enqueued by step() on this rank, counted where they are launched',
            'clocks': clocks,
        }

    # ---- the whole pipeline of configs[4] (with FBP), on the same resident scan
    extra = {}
    if not args.no_extra and name == 'C5':
        sb.close()
        sb = ScanBench(name, dev, rank, ws, dtype=args.dtype, collective='nccl', with_fbp=True, assemble=False)
        tp = sb.run_pipeline(2)
        if rank == 0:
            extra['c5_pipeline'] = {
                'workload': WORKLOADS['C5'] + ', 4 x FBP -> 128 x 1024^2 (ramp 0.8)', 'ms': tp,
                'rays_per_s_pipeline': sb.rays_total / (tp['total'] * 1e-3),
                'fbp_pixel_views_per_s': 2.0 * sb.R * sb.n_matrix ** 2 * sb.n_views / (tp['fbp_2_log_sinograms'] * 1e-3),
                'fbp_note': 'slice-by-slice fan-beam FBP of every detector row (no cone-angle / FDK weighting): a throughput figure, '
                            'not a valid cone-beam reconstruction off the mid-plane',
                'collective': 'none' if ws == 1 else 'no sinogram assembly (view-sharded FBP reads local views only); partial volumes '
                              'summed by one NCCL reduce-scatter per volume into row slabs + 8-byte MAX all-reduce',
                'device_mem_gb': torch.cuda.max_memory_allocated(dev) / 1e9}
    sb.close()
    del sb

    # ---- end to end through the reference-facing entry points, host buffers in and out
    if not args.no_e2e:
        e2e, note = e2e_leg(name, dev, rank, ws, args.dtype, steps=max(2, min(args.steps, 3)) if name == 'C5' else max(3, min(args.steps, 10)))
        if e2e is None and name != 'C4':
            e2e, note2 = e2e_leg('C4', dev, rank, ws, args.dtype, steps=3)
            if e2e is not None:
                e2e['note'] = f'{note}; measured on C4 instead'
        if rank == 0:
            line['e2e'] = e2e if e2e is not None else {'value': None, 'unit': UNIT, 'note': note}
        torch.cuda.empty_cache()

    # ---- secondary configurations (continuity with round 1)
    if not args.no_extra:
        for other, steps in (('C4', 3), ('C2', 20)):
            if other == name:
                continue
            ob = ScanBench(other, dev, rank, ws, dtype=args.dtype, collective=args.collective)
            to = ob.run(steps, 3)
            ok = ob.verify_peer()
            ro, rg = ob.rooflines(to, clocks, peaks)
            if rank == 0:
                extra[other.lower()] = {'workload': WORKLOADS[other], 'scaling': ob.scaling, 'steps': steps, 'ms_per_step': to['total'],
                                        'kernel_ms': {'siddon_poly': to['siddon'], 'gn': to['gn']},
                                        'value': ob.rays_total / (to['total'] * 1e-3), 'unit': UNIT, 'rays_per_step': ob.rays_total,
                                        'assembly': ob.collective_text(), 'roofline': ro, 'roofline_gn': rg,
                                        **({'peer_buffers_equal_nccl_all_gather': ok} if ok is not None else {})}
            ob.close()
            del ob
        if ws == 1 and rank == 0:
            extra.update(extra_single_gpu(dev, args.dtype, peaks))

    if rank == 0:
        if ws == 1 and not args.no_cpu_baseline:
            arm = CpuArm(name, args.cpu_views or auto_cpu_views(name))
            t0 = time.perf_counter()
            arm.step()
            dtc = time.perf_counter() - t0
            arm.close()
            line['cpu_baseline'] = arm.describe(arm.rays / dtc)
            line['cpu_baseline']['segments_per_ray'] = arm.seg_per_ray
        if extra:
            line['extra'] = extra
        print(json.dumps(line), flush=True)
    if ws > 1:
        dist.barrier()
        dist.destroy_process_group()


def extra_single_gpu(dev, dtype, peaks):
    """Secondary device timings at N = 1: the 3-D projector on two 48-view ranges of C4 (the slowest and a typical one),
    workspace preparation, and fan-beam FBP of one C2 sinogram."""
    import torch
    from dexct import synthetic as S
    from dexct.ops import Projector, count_segments, fbp_fan, make_scan
    from xtomosim import forward_project as FP
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
    ct, ph, lo, hi = S.make_config('C4', device=dev)
    scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, n_rows=ct.N_rows, h_iso=ct.h_iso, rotation=ct.rotation)
    pj = Projector(scan, ph.M, (ph.dx, ph.dy, ph.dz), device=dev)
    out['prepare_c4_ms'] = timed(pj.prepare, 3)
    mu = torch.as_tensor(ph.mu_table(lo.E), dtype=td, device=dev)
    W = torch.as_tensor(np.stack([FP.spectral_weights(ct, lo), FP.spectral_weights(ct, hi)]), dtype=td, device=dev)
    counts = torch.empty((2, n_views, ct.N_rows, ct.N_channels), dtype=td, device=dev)
    logs = torch.empty_like(counts)
    rays = n_views * ct.N_rows * ct.N_channels
    out['siddon_3d_c4'] = []
    for v0 in (0, 150, 300, 450):
        t = timed(lambda: pj.project(mu, W, view_begin=v0, view_count=n_views, dtype=dtype, counts=counts, logs=logs), 5)
        seg = count_segments(pj, v0, 4) / (4 * ct.N_rows * ct.N_channels)
        ach = rays * (seg + 2 * esz) / (t * 1e-3) / 1e9
        out['siddon_3d_c4'].append({'views': [v0, v0 + n_views], 'ms': t, 'rays_per_s': rays / (t * 1e-3), 'segments_per_ray': seg,
                                    'roofline': {'bound': 'hbm', 'achieved': ach, 'peak': hbm, 'unit': 'GB/s', 'frac': ach / hbm}})
    out['workload_c4_ranges'] = '512^3 phantom, 64 rows x 900 ch, 48 of 1200 views per range, 120 bins, 2 spectra fused'
    return out


def main():
    args = parse()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_b200(args)


if __name__ == '__main__':
    main()
