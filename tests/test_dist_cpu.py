"""N>1 host logic on CPU: spp shards tile [0,spp) and the reduce step sums per-rank accumulators (gloo, world 2)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from raytracingweekend_b200.dist import reduce_accumulators, shard_range


def test_shard_ranges_tile_the_sample_range():
    for spp in (1, 7, 64, 1024, 4096):
        for world in (1, 2, 3, 4, 8):
            ranges = [shard_range(spp, r, world) for r in range(world)]
            assert ranges[0][0] == 0 and sum(c for _, c in ranges) == spp
            for (b0, c0), (b1, _) in zip(ranges, ranges[1:]):
                assert b0 + c0 == b1
    assert shard_range(1024, 3, 8) == (384, 128)
    with pytest.raises(ValueError):
        shard_range(8, 2, 2)


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    nx, ny, spp = 6, 4, 10
    begin, count = shard_range(spp, rank, world)
    # stand-in for the device accumulator of this shard: sample s contributes (s+1) to every channel
    acc = torch.zeros((ny, nx, 4))
    for s in range(begin, begin + count):
        acc[..., :3] += float(s + 1)
    acc[..., 3] = count
    reduce_accumulators(acc, dst=0)
    if rank == 0:
        np.save(os.path.join(out_dir, "acc.npy"), acc.numpy())
    dist.destroy_process_group()


def test_two_rank_reduce_gloo(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    acc = np.load(tmp_path / "acc.npy")
    assert np.all(acc[..., :3] == sum(range(1, 11))) and np.all(acc[..., 3] == 10)
