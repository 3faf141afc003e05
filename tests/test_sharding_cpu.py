"""CPU, world size 2, gloo: the N>1 host path (unit sharding + result gather)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from visual_slam_b200.sharding import gather_results, shard_range, shard_sequences


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 8, 8000, 8001):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    assert shard_sequences(8, 1, 4) == [1, 5] and sum(len(shard_sequences(8, r, 8)) for r in range(8)) == 8


def _worker(rank, world, port, n_units):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_range(n_units, rank, world)
    local = torch.arange(lo, hi, dtype=torch.float64)[:, None] * torch.tensor([[1.0, 10.0, 100.0]])
    out = gather_results(local, n_units, world)
    want = torch.arange(n_units, dtype=torch.float64)[:, None] * torch.tensor([[1.0, 10.0, 100.0]])
    assert torch.equal(out, want)
    dist.barrier()
    dist.destroy_process_group()


def test_gather_results_world_size_2_gloo():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_worker, args=(2, port, 11), nprocs=2, join=True)
