"""Multi-GPU plumbing: samples-per-pixel sharding + one reduce of the radiance accumulators.

One process per GPU (torchrun).  GPU g of G renders global sample indices [g*spp/G, (g+1)*spp/G) of EVERY pixel;
Philox streams are keyed by the global sample index, so shards are disjoint by construction and the scene is
simply replicated.  The only exchange step is `torch.distributed.reduce(accum, dst=0, SUM)` — ncclReduce of
4*nx*ny floats over NVLink; rank 0 then tonemaps and writes the PPM.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(spp: int, rank: int, world: int) -> Tuple[int, int]:
    """(spp_begin, spp_count) of `rank`: [rank*spp/world, (rank+1)*spp/world). Ranges tile [0, spp) exactly."""
    if world < 1 or not (0 <= rank < world) or spp < 0:
        raise ValueError("bad shard arguments")
    lo = rank * spp // world
    hi = (rank + 1) * spp // world
    return lo, hi - lo


def reduce_accumulators(accum: torch.Tensor, dst: int = 0, group=None) -> torch.Tensor:
    """Sum the per-rank [ny,nx,4] accumulators onto rank `dst` (ncclReduce on CUDA tensors, gloo on CPU).
    The w channel carries sample counts, so the reduced buffer is again (sum rgb, samples)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return accum


def render_sharded(scene, nx: int, ny: int, spp: int, max_depth: int = 100, seed: int = 0, flags: int = 0,
                   accum: Optional[torch.Tensor] = None, reduce: bool = True):
    """Render this rank's sample shard on its GPU and (optionally) reduce to rank 0.

    Returns (accum tensor on the scene's device, stats of the local shard).  `accum` is complete only on rank 0.
    """
    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    begin, count = shard_range(spp, rank, world)
    dev = torch.device("cuda", scene.device)
    if accum is None:
        accum = torch.empty((ny, nx, 4), dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream(dev)
    if count > 0:
        scene.render_device(accum.data_ptr(), nx, ny, count, spp_begin=begin, max_depth=max_depth, seed=seed, flags=flags,
                            stream=stream.cuda_stream)
        stats = scene.finish()
    else:
        accum.zero_()
        stats = {"paths": 0, "rays": 0, "kernel_ms": 0.0, "kernel_launches": 0}
    if reduce:
        reduce_accumulators(accum, dst=0)
    return accum, stats
