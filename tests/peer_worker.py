"""Worker of tests/test_gpu_peer.py: one rank of a peer-memory group (run as a subprocess; gloo carries the handles).

Ranks use distinct GPUs when the box has enough, otherwise they share cuda:0 (CUDA IPC works between processes on one
device; their kernels are then time-sliced)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, 'dex-ct-sim_b200')):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from dexct import dist as ddist  # noqa: E402
from dexct import synthetic as S  # noqa: E402
from dexct.ops import F32, F64, gn_decompose  # noqa: E402
from dexct.peer import PeerGroup  # noqa: E402


def main():
    rank, ws = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
    dev = torch.device('cuda', rank % torch.cuda.device_count())
    torch.cuda.set_device(dev)
    dist.init_process_group('gloo', rank=rank, world_size=ws)
    # ---- a set-up failure on ONE rank (injected: allocation on rank 0, mapping on the last rank) raises on ALL ranks, and
    # what had succeeded is undone: the next group works
    from dexct._native import DexctError
    for fault in ('alloc:0', f'open:{ws - 1}'):
        os.environ['DEXCT_PEER_FAIL'] = fault
        try:
            PeerGroup(timeout_s=10.0)
        except DexctError as exc:
            assert 'rank' in str(exc) or 'mapped' in str(exc), str(exc)
        else:
            raise AssertionError(f'rank {rank}: PeerGroup came up despite the injected fault {fault}')
        finally:
            del os.environ['DEXCT_PEER_FAIL']
    assert ddist.all_agree(True) and not ddist.all_agree(rank != 0)
    pg = PeerGroup(timeout_s=10.0)
    assert pg.rank == rank and pg.world_size == ws

    # ---- push + barrier: every rank's slot arrives in every copy, over several epochs
    n = 4096
    buf = pg.alloc(ws * n * 4)
    full = buf.view(torch.float32, (ws, n))
    for step in range(1, 6):
        mine = torch.full((n,), float(100 * step + rank), device=dev)
        full[rank].copy_(mine)
        pg.push(full[rank], buf, rank * n * 4, max_ctas=8, dma=(step % 2 == 0))
        pg.barrier()
        torch.cuda.synchronize()
        want = torch.tensor([100.0 * step + r for r in range(ws)], device=dev)[:, None].expand(ws, n)
        assert torch.equal(full, want), f'rank {rank} step {step}: {full[:, 0].tolist()}'
        pg.barrier()                                       # nobody overwrites a slot a peer is still checking
    # ---- integer MAX all-reduce of the air-maximum key
    for code, td, kd in ((F32, torch.float32, torch.int32), (F64, torch.float64, torch.int64)):
        for step in range(3):
            vals = [3.5 + ((r + step) % ws) for r in range(ws)]
            scratch = torch.zeros(16, dtype=torch.uint8, device=dev)
            scratch.view(kd)[0] = torch.tensor(vals[rank], dtype=td).view(kd)
            pg.max_key(scratch, code)
            torch.cuda.synchronize()
            got = float(scratch.view(kd)[:1].view(td)[0])
            assert got == max(vals), (rank, code, got, vals)
    # ---- decomposition mirrored into every rank's copy of the basis sinogram == decomposition of the whole
    g, i0, mus, _ = S.gn_problem(8 * ws, 96, 40)
    g[0, 0, :7] *= 50.0                                     # an air-like maximum that only rank 0 owns
    for code, td in ((F32, torch.float32), (F64, torch.float64)):
        whole = gn_decompose(torch.as_tensor(g, dtype=td, device=dev), i0, mus, 12, mask_thresh=0.9, dtype=code)
        esz = 4 if code == F32 else 8
        bbuf = pg.alloc(8 * ws * 96 * 2 * esz)
        basis = bbuf.view(td, (8 * ws, 96, 2))
        v0, nv = ddist.shard_range(8 * ws, rank, ws)
        local = torch.as_tensor(g[:, v0:v0 + nv], dtype=td, device=dev).contiguous()
        off = v0 * 96 * 2 * esz
        # ... and the same kernel carries this rank's count shard into the peers' copies of the count sinogram
        cbuf = pg.alloc(2 * 8 * ws * 96 * esz)
        csino = cbuf.view(td, (ws, 2, 8, 96))
        csino[rank].copy_(local)
        coff = rank * 2 * 8 * 96 * esz
        gn_decompose(csino[rank], i0, mus, 12, mask_thresh=0.9, dtype=code, out=basis[v0:v0 + nv], reduce_max_over_ranks=True,
                     peers=pg, mirrors=bbuf.table(off, exclude_self=True), push=(csino[rank], cbuf.table(coff, exclude_self=True)))
        pg.barrier()
        torch.cuda.synchronize()
        assert torch.equal(basis, whole), f'rank {rank} dtype {code}: max diff {(basis - whole).abs().max()}'
        want = torch.as_tensor(g, dtype=td, device=dev).view(2, ws, 8, 96).permute(1, 0, 2, 3)
        assert torch.equal(csino, want), f'rank {rank} dtype {code}: pushed counts differ'
        assert bool((basis[0, :7] == 0).all(dim=-1).any())    # masked on every rank, whoever owns the maximum
    assert pg.status() == 0
    dist.barrier()
    pg.close()
    dist.destroy_process_group()
    print(f'peer worker {rank}/{ws} ok', flush=True)


if __name__ == '__main__':
    main()
