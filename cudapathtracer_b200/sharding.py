"""Host-side work partitioning for the multi-GPU path.

Every (pixel, sample) is independent given a replicated read-only scene, and the Philox stream is keyed by the
GLOBAL sample index, so splitting a frame by sample ranges needs no data-path collective except one sum of the
f32 accumulators at the end.  The reference has no multi-GPU code (single device 0, kernel.cu:569).

  sample_range(rank, world, spp)      strong split of `spp` samples: contiguous, disjoint, covering, sizes differ by <= 1
  weak_sample_range(rank, spp)        weak scaling (bench.py): every rank renders `spp` samples, rank r starts at r*spp
  tile_range(rank, world, n_tiles)    the tile-sharding alternative (gather instead of reduce)
  reduce_accumulators(t, dst)         torch.distributed sum of the per-rank f32 sums onto rank `dst` (NCCL on GPUs, gloo in tests)
  normalise(sum_image, total_samples) mean image = sum / samples (kernel.cu:552 keeps a running mean instead)
"""
from __future__ import annotations

from typing import Tuple


def sample_range(rank: int, world: int, spp: int) -> Tuple[int, int]:
    """(first_sample, num_samples) of `rank`"""
    if not (0 <= rank < world) or spp < 0:
        raise ValueError("bad rank/world/spp")
    base, extra = divmod(spp, world)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def weak_sample_range(rank: int, spp: int) -> Tuple[int, int]:
    if rank < 0 or spp < 0:
        raise ValueError("bad rank/spp")
    return rank * spp, spp


def tile_range(rank: int, world: int, n_tiles: int) -> Tuple[int, int]:
    return sample_range(rank, world, n_tiles)


def reduce_accumulators(accum, dst: int = 0):
    """in-place sum onto rank `dst`; a no-op without an initialised process group"""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum


def normalise(sum_image, total_samples: int):
    return sum_image / float(max(1, total_samples))
