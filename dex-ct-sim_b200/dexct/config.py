"""Run-time switches of the drop-in modules (the reference has none: its constants are hard-coded in main.py)."""
import os

# Arithmetic of the kernels: 'f32' (default, the throughput path) or 'f64' (matches the reference's float64 numpy
# results to ~1e-10).  Environment override: DEXCT_DTYPE.
dtype = os.environ.get('DEXCT_DTYPE', 'f32')

# Keep the per-material path lengths of the last (ct, phantom) pair on the device so that the second get_sino call
# of a dual-energy pair (/root/reference/main.py:109-120) only re-runs the detection epilogue.
cache_lengths = os.environ.get('DEXCT_CACHE_LENGTHS', '1') != '0'

# Shard projection views (and decomposition pixels) over torch.distributed ranks when a process group is initialised.
shard_views = os.environ.get('DEXCT_SHARD_VIEWS', '1') != '0'

# With sharded views: True (default) = every rank returns the WHOLE arrays, like the single-process reference (all-gather +
# a full device-to-host copy on every rank, so the host side does not scale); False = every rank returns, and accepts, only
# ITS OWN contiguous view range [dexct.dist.shard_range(N_proj, rank, world)] -- the data-parallel form in which each rank
# writes its part of the output files.  The air-mask maximum is still taken over all ranks.
gather_views = os.environ.get('DEXCT_GATHER_VIEWS', '1') != '0'

# Print progress the way the reference does (matdecomp.py:111-112)
verbose = os.environ.get('DEXCT_VERBOSE', '0') != '0'

# Poisson counting statistics on the simulated counts.  Off by default so that results are
# deterministic functions of the inputs; a scanner object can override it (FanBeamGeometry(noise=...)).  The generator is
# counter-based (Philox4x32-10 keyed by noise_seed, addressed by the element's index in the whole sinogram), so the
# noise realisation does not depend on how views are sharded over GPUs.
noise = os.environ.get('DEXCT_NOISE', '0') != '0'
noise_seed = int(os.environ.get('DEXCT_NOISE_SEED', '20240'))
# sino_log of a noisy scan is -ln(max(counts, noise_floor) / air counts): a pixel that detected nothing stays finite
noise_floor = 0.5
