"""Sample-per-pixel sharding over the GPUs of one box (the reference has no multi-GPU path).

One process per GPU (torchrun). Rank g of G renders the sample indices {g, g+G, g+2G, ...} of EVERY pixel into its own
accumulation buffer; Philox is keyed by (seed; pixel, sample, bounce), so a sample is the same whichever rank computes it.
The buffers are int64 x {R, G, B, samples} in Q31.32 fixed point (include/crt.h), so ONE sum-reduce over int64 is exact and
order-free: the image of G ranks is bit-identical to the image one GPU renders for the same sample set. torch.distributed
is only the plumbing (NCCL over NVLink on the GPU box, gloo in the CPU tests); the path has no other exchange step.
"""
from __future__ import annotations

from typing import Tuple


def sample_shard(rank: int, world: int, spp_total: int) -> Tuple[int, int, int]:
    """(spp_begin, spp_stride, spp_count) of `rank`: sample indices rank, rank+world, ... below spp_total.

    The shards are disjoint, their union is range(spp_total), and their sizes differ by at most one."""
    if world < 1 or not (0 <= rank < world) or spp_total < 0:
        raise ValueError("sample_shard: need 0 <= rank < world and spp_total >= 0")
    count = (spp_total - rank + world - 1) // world if spp_total > rank else 0
    return rank, world, count


def tile_shard(rank: int, world: int, height: int, rows_per_tile: int = 16) -> list:
    """Alternative partition (image tiles): interleaved row slabs [(y0, y1), ...] of `rank`. Each pixel lives on one GPU, so
    no reduction is needed, only a gather; kept for comparison, spp sharding is the default."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("tile_shard: need 0 <= rank < world")
    slabs = []
    for i, y0 in enumerate(range(0, height, rows_per_tile)):
        if i % world == rank:
            slabs.append((y0, min(y0 + rows_per_tile, height)))
    return slabs


def reduce_accum(accum, dst: int = 0, group=None):
    """Sum-reduce the int64 accumulation tensors of all ranks onto `dst` (in place). Exact for any rank count and order."""
    import torch
    import torch.distributed as dist

    if accum.dtype != torch.int64:
        raise TypeError("reduce_accum: the accumulation buffer is int64 x {R, G, B, samples} (Q31.32 fixed point)")
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return accum


def render_sharded(renderer, accum, spp_total: int, max_depth: int, rank: int, world: int, seed: int = 1984, variant: int = 0,
                   dst: int = 0, group=None):
    """This rank's share of an spp_total-sample frame into `accum` (a zeroed int64 (H, W, 4) CUDA tensor registered with
    renderer.set_accum_device_buffer), followed by the one reduce. Returns the number of samples this rank rendered."""
    begin, stride, count = sample_shard(rank, world, spp_total)
    if count:
        renderer.render(spp=count, max_depth=max_depth, spp_begin=begin, spp_stride=stride, seed=seed, variant=variant)
    reduce_accum(accum, dst=dst, group=group)
    return count
