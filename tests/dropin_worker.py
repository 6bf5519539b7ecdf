"""Worker of tests/test_gpu_sharded.py: one NCCL rank of a view-sharded run through the drop-in entry points
(get_sinos / get_basismat_sinos / get_recon), checked against the same calls without sharding."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'dex-ct-sim_b200')):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import matdecomp  # noqa: E402
from dexct import config  # noqa: E402
from dexct import dist as ddist  # noqa: E402
from dexct import synthetic as S  # noqa: E402
from xtomosim import back_project as BP  # noqa: E402
from xtomosim import forward_project as FP  # noqa: E402


def main():
    rank, ws = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', rank=rank, world_size=ws, device_id=torch.device('cuda', rank))
    ct, ph, lo, hi = S.make_config('C2', n_views=37, n=96)          # 37 views: ragged shards
    ct.noise, ct.noise_seed = True, 5                                # noise must not depend on the sharding either

    def run():
        FP.clear_cache()
        ct.noise_stream = 0
        (r1, l1), (r2, l2) = FP.get_sinos(ct, ph, [lo, hi])
        m1, m2 = matdecomp.get_basismat_sinos(ct, r1, r2, lo, hi, n_iters=12, mask_thresh=0.95)
        img, hu = BP.get_recon(l1, ct, lo, 64, 51.2, 0.8)
        return r1, l1, r2, l2, m1, m2, img, hu

    config.shard_views = False
    whole = run()
    config.shard_views, config.gather_views = True, True
    gathered = run()
    for a, b in zip(whole[:6], gathered[:6]):
        assert a.shape == b.shape and np.array_equal(a, b), 'gathered arrays differ from the unsharded run'
    for a, b in zip(whole[6:], gathered[6:]):                       # partial images are summed in another order
        assert np.max(np.abs(a - b)) <= 2e-6 * np.max(np.abs(a))
    config.gather_views = False
    local = run()
    v0, nv = ddist.shard_range(ct.N_proj, rank, ws)
    for a, b in zip(whole[:6], local[:6]):
        assert b.shape[0] == nv and np.array_equal(a[v0:v0 + nv], b), 'local view range differs from the unsharded run'
    for a, b in zip(whole[6:], local[6:]):                          # the image is whole on every rank (all-reduce)
        assert a.shape == b.shape and np.max(np.abs(a - b)) <= 2e-6 * np.max(np.abs(a))
    try:
        matdecomp.get_basismat_sinos(ct, whole[0][:5], whole[2][:5], lo, hi, n_iters=2)
        raise SystemExit('a sinogram that is neither whole nor this rank\'s range must be refused')
    except ValueError:
        pass
    dist.barrier()
    dist.destroy_process_group()
    print(f'dropin worker {rank}/{ws} ok', flush=True)


if __name__ == '__main__':
    main()
