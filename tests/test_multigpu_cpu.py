"""CPU (gloo, world_size 2 and 3) tests of the multi-GPU host logic: sample sharding and the exact int64 reduce.
The render itself needs a GPU (tests/test_gpu_multigpu.py); here every rank fills its buffer with a deterministic stand-in for
the fixed-point sample values, which is all the reduce sees."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cudaraytracer_b200.multigpu import reduce_accum, row_shard, sample_shard, tile_shard


def test_sample_shards_partition_the_sample_range():
    for world in (1, 2, 3, 4, 8):
        for spp in (0, 1, 5, 8, 64, 1023, 1024):
            seen = []
            sizes = []
            for rank in range(world):
                b, s, c = sample_shard(rank, world, spp)
                assert (b, s) == (rank, world)
                seen += [b + i * s for i in range(c)]
                sizes.append(c)
            assert sorted(seen) == list(range(spp))
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sample_shard(2, 2, 8)


def test_tile_shards_cover_every_row_once():
    for world in (1, 2, 3, 8):
        for H in (1, 15, 16, 17, 1080, 2160):
            rows = []
            for rank in range(world):
                for y0, y1 in tile_shard(rank, world, H):
                    rows += list(range(y0, y1))
            assert sorted(rows) == list(range(H))


def test_row_tiles_of_the_c_abi_agree_with_tile_shard():
    """crt_render_params::row_tiles (what the kernels map rows with) and multigpu.tile_shard describe the same slabs."""
    from cudaraytracer_b200 import capi
    lib = capi.load()
    for world in (1, 2, 3, 8, 255):
        for rpt in (1, 5, 16, 64):
            for H in (1, 15, 16, 17, 83, 1080, 4320):
                counts = []
                for rank in range(world):
                    want = sum(y1 - y0 for y0, y1 in tile_shard(rank, world, H, rpt))
                    got = lib.crt_row_tiles_rows(capi.row_tiles(rank, world, rpt), H)
                    assert got == want, (world, rpt, H, rank)
                    counts.append(got)
                assert sum(counts) == H
    assert lib.crt_row_tiles_rows(0, 123) == 123
    assert lib.crt_row_tiles_rows((16 << 16) | (2 << 8) | 2, 100) == 0  # rank == world
    assert lib.crt_row_tiles_rows((0 << 16) | (2 << 8) | 1, 100) == 0   # zero rows per slab
    with pytest.raises(ValueError):
        capi.row_tiles(3, 3)


def test_row_shards_are_contiguous_slabs_that_cover_the_frame():
    for world in (1, 2, 3, 8, 16):
        for H in (0, 1, 7, 16, 1080, 2160, 4320):
            prev, sizes = 0, []
            for rank in range(world):
                b, e = row_shard(rank, world, H)
                assert b == prev and b <= e
                prev = e
                sizes.append(e - b)
            assert prev == H and max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        row_shard(3, 3, 10)


def _fixed_sample(pixel, sample):
    """Stand-in for crt_to_fixed(radiance): any int64 function of (pixel, sample), including negative and large values."""
    with np.errstate(over="ignore"):
        k = np.array([sample], dtype=np.uint64) * np.uint64(0xD1B54A32D192ED03)
        x = (pixel.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15) + k) >> np.uint64(20)
    return x.astype(np.int64) - (1 << 42)


def _full(H, W, spp):
    px = np.arange(H * W, dtype=np.int64)
    acc = np.zeros((H * W, 4), dtype=np.int64)
    for s in range(spp):
        v = _fixed_sample(px, s)
        acc[:, 0] += v; acc[:, 1] += v >> 1; acc[:, 2] -= v
    acc[:, 3] = spp
    return acc.reshape(H, W, 4)


def _worker(rank, world, port, H, W, spp, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        b, st, c = sample_shard(rank, world, spp)
        px = np.arange(H * W, dtype=np.int64)
        acc = np.zeros((H * W, 4), dtype=np.int64)
        for i in range(c):
            v = _fixed_sample(px, b + i * st)
            acc[:, 0] += v; acc[:, 1] += v >> 1; acc[:, 2] -= v
        acc[:, 3] = c
        t = torch.from_numpy(acc.reshape(H, W, 4))
        reduce_accum(t, dst=0)
        if rank == 0:
            np.save(out, t.numpy())
        with pytest.raises(TypeError):
            reduce_accum(torch.zeros(4, dtype=torch.float32))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,spp", [(2, 8), (2, 7), (3, 10)])
def test_reduce_is_exact_and_independent_of_the_rank_count(tmp_path, world, spp):
    H, W = 9, 13
    out = str(tmp_path / "acc.npy")
    mp.spawn(_worker, args=(world, _free_port(), H, W, spp, out), nprocs=world, join=True)
    got = np.load(out)
    assert np.array_equal(got, _full(H, W, spp))  # bit-identical to the single-process sum
