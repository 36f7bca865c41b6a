"""The N>1 path on CPU: contiguous frame sharding over ranks, host-side gather of per-frame results and the
max-over-ranks timing reduction the benchmark uses — world_size 2 over gloo (no GPU, no data-path collective)."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from librectify_b200.runner import shard_bounds


def test_shard_bounds_partition():
    for n in (0, 1, 7, 64, 4096, 4097):
        for g in (1, 2, 3, 4, 8):
            blocks = [shard_bounds(n, g, r) for r in range(g)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(g - 1))
            sizes = [hi - lo for lo, hi in blocks]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, n_frames, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_bounds(n_frames, world, rank)
    # stand-in for the per-rank engine: each frame yields an 80-byte transform record keyed by its global index
    local = torch.arange(lo, hi, dtype=torch.float64).repeat_interleave(10).reshape(-1, 10)
    counts = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([hi - lo]))
    gathered = [torch.zeros((int(c.item()), 10), dtype=torch.float64) for c in counts]
    dist.all_gather_object(obj := [None] * world, local.numpy())
    t = torch.tensor([0.25 + rank])         # per-rank elapsed seconds; the job's time is the slowest rank's
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        allr = np.concatenate(obj)
        np.save(out, np.concatenate([allr[:, 0], [t.item()]]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gather(tmp_path):
    out = str(tmp_path / "g.npy")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, 37, out), nprocs=2, join=True)
    r = np.load(out)
    assert np.array_equal(r[:-1], np.arange(37, dtype=np.float64))      # every frame exactly once, in order
    assert r[-1] == 1.25                                                 # max over ranks
