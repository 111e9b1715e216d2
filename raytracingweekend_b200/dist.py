"""Multi-GPU plumbing: samples-per-pixel sharding + one reduce of the radiance accumulators.

One process per GPU (torchrun).  GPU g of G renders global sample indices [g*spp/G, (g+1)*spp/G) of EVERY pixel;
Philox streams are keyed by the global sample index, so shards are disjoint by construction and the scene is
simply replicated.  The only exchange step is one ncclReduce of 4*nx*ny floats over NVLink; rank 0 then tonemaps and
writes the PPM.  The data plane is the C ABI (`Communicator` = rtw_comm: ncclCommInitRank / ncclReduce inside
librtw_b200.so; `Scene.render_sharded` = rtw_render_sharded); torch.distributed only carries the 128-byte NCCL id at
start-up.  `reduce_accumulators` / `render_sharded` below are the torch.distributed twins used on CPU (gloo tests).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch
import torch.distributed as dist

from . import _lib


class Communicator:
    """rtw_comm: an NCCL communicator owned by librtw_b200.so (include/rtw_b200.h, section 4)."""

    def __init__(self, comm_id: bytes, rank: int, world: int, device: int):
        self._l = _lib.load()
        buf = (C.c_uint8 * _lib.RTW_COMM_ID_BYTES).from_buffer_copy(comm_id)
        self.handle = self._l.rtw_comm_create(buf, rank, world, device)
        if not self.handle:
            raise _lib.RtwError("rtw_comm_create: " + self._l.rtw_last_error().decode())
        self.rank, self.world, self.device = rank, world, device

    @staticmethod
    def unique_id() -> bytes:
        l = _lib.load()
        buf = (C.c_uint8 * _lib.RTW_COMM_ID_BYTES)()
        _lib.check(l.rtw_comm_unique_id(buf), "rtw_comm_unique_id")
        return bytes(buf)

    @classmethod
    def from_torch(cls, rank: int, world: int, device: int, group=None) -> "Communicator":
        """Rank 0 draws the NCCL id, torch.distributed broadcasts its 128 bytes, every rank joins."""
        ids = [cls.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0, group=group)
        return cls(ids[0], rank, world, device)

    def reduce(self, accum_ptr: int, nx: int, ny: int, root: int = 0, stream: int = 0) -> None:
        """In-place ncclReduce(SUM) of the [ny, nx, 4] float32 device buffer onto `root`, asynchronous on `stream`."""
        _lib.check(self._l.rtw_reduce_accumulators(self.handle, C.c_void_p(accum_ptr), nx, ny, root, C.c_void_p(stream)),
                   "rtw_reduce_accumulators")

    def close(self):
        if self.handle:
            self._l.rtw_comm_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def shard_range(spp: int, rank: int, world: int) -> Tuple[int, int]:
    """(spp_begin, spp_count) of `rank`: [rank*spp/world, (rank+1)*spp/world). Ranges tile [0, spp) exactly."""
    if world < 1 or not (0 <= rank < world) or spp < 0:
        raise ValueError("bad shard arguments")
    b, n = C.c_int(), C.c_int()
    _lib.check(_lib.load().rtw_shard_range(spp, rank, world, C.byref(b), C.byref(n)), "rtw_shard_range")
    return b.value, n.value


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
