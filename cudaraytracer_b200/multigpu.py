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


def row_shard(rank: int, world: int, height: int) -> Tuple[int, int]:
    """Rows [begin, end) whose reduction + resolve `rank` owns in the peer-memory close-out: contiguous slabs that cover
    range(height) exactly once, sizes differing by at most one row (a rank may own no row when world > height)."""
    if world < 1 or not (0 <= rank < world) or height < 0:
        raise ValueError("row_shard: need 0 <= rank < world and height >= 0")
    return rank * height // world, (rank + 1) * height // world


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


class PeerFrame:
    """The frame close-out over peer memory (include/crt.h: crt_resolve_peers) instead of reduce-to-root + resolve.

    Every rank keeps accumulating into its renderer's OWN buffer; at construction the ranks exchange CUDA IPC handles of those
    buffers (and of the root's RGBA8 frame) through torch.distributed and map them. close() then runs, on every rank, one kernel
    that sums the rows this rank owns over all ranks' buffers (int64: the same bits as the NCCL sum), resolves them and stores
    the packed pixels into the root's frame over NVLink. Two one-element all-reduces on the launching stream order it after
    every rank's render and before anybody's next clear; nothing synchronises with the host. torch.distributed stays the
    plumbing (handle exchange, the two stream-ordered barriers); the data moves inside our kernel.

    PRECONDITION: the renderer runs on torch's CURRENT stream - renderer.set_stream(torch.cuda.current_stream().cuda_stream)
    before constructing this, and the same stream current at every close_frame*(). The fences are all-reduces on torch's current
    stream; with the render / resolve kernels on another stream they would order nothing, and a rank could read a peer's buffer
    while that peer is still rendering into it (or clearing it for the next frame). Checked at construction and at every fence."""

    def __init__(self, renderer, rank: int, world: int, dst: int = 0, group=None):
        import torch
        import torch.distributed as dist
        from . import capi

        self.r, self.rank, self.world, self.dst, self.group = renderer, rank, world, dst, group
        self._check_stream()
        self.rows = row_shard(rank, world, renderer.height)
        mine = (renderer.ipc_export(capi.BUFFER_ACCUM), renderer.ipc_export(capi.BUFFER_RGBA) if rank == dst else b"")
        handles = [None] * world
        if world > 1:
            dist.all_gather_object(handles, mine, group=group)
        else:
            handles[0] = mine
        self._opened = []

        def mapped(handle, local_ptr, owner):
            if owner == rank:
                return local_ptr
            p = renderer.ipc_open(handle)
            self._opened.append(p)
            return p

        self.peer_ptrs = [mapped(handles[g][0], renderer.accum_ptr(), g) for g in range(world)]
        self.out_ptr = mapped(handles[dst][1], renderer.rgba_ptr(), dst)
        self._token = torch.zeros(1, dtype=torch.int32, device=torch.device("cuda", torch.cuda.current_device()))

    def _check_stream(self):
        import torch
        cur = int(torch.cuda.current_stream().cuda_stream)
        if self.r.stream_handle is None or int(self.r.stream_handle) != cur:
            raise RuntimeError("PeerFrame: the renderer must run on torch's current stream (renderer.set_stream(torch.cuda.current_stream()"
                               ".cuda_stream)): the stream-ordered fences are issued there and would not order kernels on another stream")

    def _fence(self):
        """Stream-ordered barrier: completes on this rank's stream only after every rank's earlier work on its stream."""
        import torch.distributed as dist
        self._check_stream()
        if self.world > 1:
            dist.all_reduce(self._token, group=self.group)

    def close_frame(self):
        """After every rank's render of the frame: the root's RGBA8 frame (renderer.read_rgba() on the root) is complete when
        the root's stream reaches the end of this call."""
        self._fence()
        b, e = self.rows
        self.r.resolve_peers(self.peer_ptrs, b, e, self.out_ptr)
        self._fence()

    def close_frame_tiles(self, rows_per_tile: int = 16):
        """Close-out of an image-TILE sharded frame (every rank rendered all samples of capi.row_tiles(rank, world, rows_per_tile)
        with its own buffer): nothing to sum - each rank resolves its rows straight into the root's frame (4 bytes per pixel
        over NVLink) between two fences: after the root's reads of the previous frame, and before the root reads this one."""
        from . import capi
        self._fence()  # the root's reads of the previous frame (on its stream) come before anybody's stores into it
        self.r.resolve_rows(capi.row_tiles(self.rank, self.world, rows_per_tile), self.out_ptr)
        self._fence()

    def release(self):
        for p in self._opened:
            self.r.ipc_close(p)
        self._opened = []
