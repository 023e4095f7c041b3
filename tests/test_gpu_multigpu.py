"""GPU test of the spp-sharded path on ONE GPU: G 'ranks' (contexts) render their sample shards into separate int64
buffers, the buffers are summed like the NCCL reduce does, and the result must be bit-identical to the single render."""
import os

import numpy as np
import pytest

from cudaraytracer_b200 import Renderer, Scene, capi
from cudaraytracer_b200.multigpu import sample_shard

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world", [2, 3, 8])
def test_sharded_sum_is_bit_identical(world):
    import torch
    W, H, spp, depth = 128, 72, 13, 50
    s = Scene.generate(capi.SCENE_RTIOW, 1984)
    cam = s.default_camera(W, H)
    ref_r = Renderer(0)
    ref_r.upload(s); ref_r.set_camera(cam); ref_r.resize(W, H)
    ref = torch.zeros((H, W, 4), dtype=torch.int64, device="cuda")
    ref_r.set_accum_device_buffer(ref.data_ptr())
    ref_r.render(spp=spp, max_depth=depth)
    ref_r.synchronize()
    total = torch.zeros_like(ref)
    for rank in range(world):
        r = Renderer(0)
        r.upload(s); r.set_camera(cam); r.resize(W, H)
        acc = torch.zeros((H, W, 4), dtype=torch.int64, device="cuda")
        r.set_accum_device_buffer(acc.data_ptr())
        b, st, c = sample_shard(rank, world, spp)
        if c:
            r.render(spp=c, max_depth=depth, spp_begin=b, spp_stride=st,
                     variant=[capi.VARIANT_MEGAKERNEL, capi.VARIANT_PIPELINE, capi.VARIANT_WAVEFRONT][rank % 3])
        r.synchronize()
        total += acc  # what dist.reduce(SUM) over int64 computes
        r.close()
    assert torch.equal(total, ref)
    assert int(ref[..., 3].min()) == spp == int(ref[..., 3].max())
    ref_r.close()


@pytest.mark.parametrize("world", [1, 2, 5, 16])
def test_resolve_peers_equals_sum_then_resolve(world):
    """crt_resolve_peers (sum over the ranks' buffers + resolve, by row slabs) on ONE GPU, the G buffers being plain local
    pointers: the frame and the summed buffer must be bit-identical to 'add the buffers, then crt_resolve'."""
    import torch
    from cudaraytracer_b200.multigpu import row_shard
    W, H, spp, depth = 200, 37, 9, 50
    s = Scene.generate(capi.SCENE_RTIOW, 1984)
    cam = s.default_camera(W, H)
    ranks = []
    for rank in range(world):
        r = Renderer(0)
        r.upload(s); r.set_camera(cam); r.resize(W, H)
        b, st, c = sample_shard(rank, world, spp)
        if c:
            r.render(spp=c, max_depth=depth, spp_begin=b, spp_stride=st)
        r.synchronize()
        ranks.append(r)
    # the reference result: one context renders all samples, resolves
    ref_r = Renderer(0)
    ref_r.upload(s); ref_r.set_camera(cam); ref_r.resize(W, H)
    ref = torch.zeros((H, W, 4), dtype=torch.int64, device="cuda")
    ref_r.set_accum_device_buffer(ref.data_ptr())
    ref_r.render(spp=spp, max_depth=depth)
    want = ref_r.resolve()
    # every 'rank' closes its own row slab into rank 0's frame
    root = ranks[0]
    ptrs = [r.accum_ptr() for r in ranks]
    total = torch.full((H, W, 4), -1, dtype=torch.int64, device="cuda")
    covered = 0
    for rank, r in enumerate(ranks):
        b, e = row_shard(rank, world, H)
        covered += e - b
        r.resolve_peers(ptrs, b, e, root.rgba_ptr(), total.data_ptr())
        r.synchronize()
    assert covered == H
    got = root.read_rgba()
    assert np.array_equal(got, want)
    assert torch.equal(total, ref)
    # argument checks
    with pytest.raises(Exception):
        root.resolve_peers(ptrs, 3, H + 1, root.rgba_ptr())
    with pytest.raises(Exception):
        root.resolve_peers(ptrs * 17, 0, H, root.rgba_ptr())
    with pytest.raises(Exception):
        root.resolve_peers([ptrs[0] + 8], 0, H, root.rgba_ptr())
    for r in ranks:
        r.close()
    ref_r.close()


@pytest.mark.parametrize("world,rows_per_tile,variant", [(1, 16, capi.VARIANT_PIPELINE), (2, 16, capi.VARIANT_PIPELINE),
                                                        (3, 5, capi.VARIANT_MEGAKERNEL), (8, 16, capi.VARIANT_WAVEFRONT),
                                                        (7, 1, capi.VARIANT_PIPELINE), (4, 64, capi.VARIANT_AUTO)])
def test_row_tile_shards_tile_the_full_render(world, rows_per_tile, variant):
    """Image-tile sharding (crt_render_params::row_tiles): each 'rank' renders ALL samples of its row slabs; the shards touch
    disjoint rows (the others stay zero), agree with multigpu.tile_shard, and together are bit-identical to the full render -
    buffers and resolved frame (crt_resolve_rows of every rank into one frame)."""
    import torch
    from cudaraytracer_b200.multigpu import tile_shard
    W, H, spp, depth = 150, 83, 5, 50
    s = Scene.generate(capi.SCENE_RTIOW, 1984)
    cam = s.default_camera(W, H)
    ref_r = Renderer(0)
    ref_r.upload(s); ref_r.set_camera(cam); ref_r.resize(W, H)
    ref = torch.zeros((H, W, 4), dtype=torch.int64, device="cuda")
    ref_r.set_accum_device_buffer(ref.data_ptr())
    ref_r.render(spp=spp, max_depth=depth)
    want = ref_r.resolve()
    frame = torch.zeros((H, W), dtype=torch.int32, device="cuda")
    r = Renderer(0)
    r.upload(s); r.set_camera(cam); r.resize(W, H)
    acc = torch.zeros((H, W, 4), dtype=torch.int64, device="cuda")
    r.set_accum_device_buffer(acc.data_ptr())
    total = torch.zeros_like(ref)
    paths = 0
    for rank in range(world):
        acc.zero_()
        r.reset_counters()
        rt = capi.row_tiles(rank, world, rows_per_tile)
        r.render(spp=spp, max_depth=depth, variant=variant, row_tiles=rt)
        r.resolve_rows(rt, frame.data_ptr())
        r.synchronize()
        mine = np.zeros(H, dtype=bool)
        for y0, y1 in tile_shard(rank, world, H, rows_per_tile):
            mine[y0:y1] = True
        touched = (acc[..., 3] != 0).any(dim=1).cpu().numpy()
        assert np.array_equal(touched, mine)
        assert torch.equal(acc[torch.from_numpy(mine).cuda()], ref[torch.from_numpy(mine).cuda()])
        paths += r.counters().paths
        total += acc
    assert torch.equal(total, ref)
    assert paths == W * H * spp
    assert np.array_equal(frame.cpu().numpy().view(np.uint32), want)
    with pytest.raises(Exception):
        r.render(spp=1, max_depth=depth, row_tiles=(16 << 16) | (2 << 8) | 2)  # rank == world
    with pytest.raises(Exception):
        r.render(spp=1, max_depth=depth, row_tiles=(0 << 16) | (2 << 8) | 1)  # zero rows per tile
    r.close()
    ref_r.close()


_IPC_CHILD = r"""
import sys, hashlib
import numpy as np
sys.path.insert(0, sys.argv[1])
from cudaraytracer_b200 import Renderer, Scene, capi
handle, W, H = bytes.fromhex(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
r = Renderer(0)
r.resize(W, H)
p = r.ipc_open(handle)                       # the parent's accumulation buffer, mapped into this process
r.resolve_peers([p], 0, H, r.rgba_ptr())     # sum of one buffer = the buffer: resolve it into OUR frame
img = r.read_rgba()
r.ipc_close(p)
r.close()
print("sha1", hashlib.sha1(img.tobytes()).hexdigest())
"""


def test_ipc_export_open_close_across_processes():
    """crt_ipc_export in this process, crt_ipc_open + crt_resolve_peers in a child process on the same GPU."""
    import hashlib
    import subprocess
    import sys
    W, H = 96, 54
    s = Scene.generate(capi.SCENE_RTIOW, 1984)
    r = Renderer(0)
    r.upload(s); r.set_camera(s.default_camera(W, H)); r.resize(W, H)
    r.render(spp=4, max_depth=50)
    want = r.resolve()
    handle = r.ipc_export(capi.BUFFER_ACCUM)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", _IPC_CHILD, root, handle.hex(), str(W), str(H)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert out.stdout.split()[-1] == hashlib.sha1(want.tobytes()).hexdigest()
    r.close()


def test_native_multigpu_host_example(tmp_path):
    """examples/multigpu_render.cpp (C++ over the C ABI, one thread per GPU) on one GPU and on every GPU of the box, with each of
    its close-outs - ONE ncclReduce(ncclInt64, ncclSum) to GPU 0 issued from C++ (north_star), the peer-memory sum + resolve
    (crt_peer_enable + crt_resolve_peers), image tiles (crt_resolve_rows into GPU 0's frame): every PNG byte-identical to the
    one-GPU frame."""
    import subprocess
    import torch
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "examples", "multigpu_render")
    G = torch.cuda.device_count()
    files = []
    for g in sorted({1, G}):
        for mode in ("nccl", "spp", "tiles"):
            out = str(tmp_path / f"f_{g}_{mode}.png")
            r = subprocess.run([exe, str(g), mode, "1", "200", "113", "6", "50", out], capture_output=True, text=True, timeout=300)
            assert r.returncode == 0, (g, mode, r.stdout, r.stderr)
            files.append(((g, mode), open(out, "rb").read()))
    assert files[0][1].startswith(b"\x89PNG")
    for key, f in files[1:]:
        assert f == files[0][1], f"{key} differs from {files[0][0]}"


def test_close_outs_across_real_gpus_under_torchrun():
    """tools/peer_check.py as a gated test: one process per GPU (torchrun, NCCL), the three close-outs of bench.py - NCCL reduce
    + resolve, the peer-memory kernel, image tiles - must give one frame, identical to what one GPU renders alone. Needs two
    GPUs; the one-GPU box of the default run skips it (tests/test_multigpu_cpu.py covers the shard arithmetic on gloo)."""
    import subprocess
    import sys
    import torch
    G = torch.cuda.device_count()
    if G < 2:
        pytest.skip("needs two GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for scene, W, H, spp in (("rtiow", 640, 360, 2 * G + 1), ("stress", 320, 200, G)):
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={G}", "--master-addr", "127.0.0.1",
               "--master-port", "29541", os.path.join(root, "tools", "peer_check.py"), str(W), str(H), str(spp), scene]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
        assert "frames identical (peer == nccl == tiles == one GPU): True" in r.stdout


def test_c_host_example_renders_a_ppm(tmp_path):
    """examples/headless_render.c (plain C over the C ABI): progressive passes, resolve, PPM on disk."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = str(tmp_path / "f.ppm")
    r = subprocess.run([os.path.join(root, "examples", "headless_render"), "1", "96", "54", "4", "50", "3", out], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    data = open(out, "rb").read()
    head = b"P6\n96 54\n255\n"
    assert data.startswith(head) and len(data) == len(head) + 96 * 54 * 3
    img = np.frombuffer(data[len(head):], dtype=np.uint8).reshape(54, 96, 3)
    assert img[0].mean() > img[-1].mean() * 0.5 and img.std() > 10  # an actual picture: sky on top (rows flipped on write)


def test_two_devices_in_one_process():
    """Contexts on different GPUs of one process (the per-device launch configuration must not be shared): both render the
    same bits. Skipped on a one-GPU box; the torchrun path (one process per GPU) is what bench.py uses."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    W, H = 160, 90
    s = Scene.generate(capi.SCENE_RTIOW, 1984)
    cam = s.default_camera(W, H)
    out = []
    for dev in (0, 1):
        r = Renderer(dev)
        r.upload(s); r.set_camera(cam); r.resize(W, H)
        for variant in (capi.VARIANT_PIPELINE, capi.VARIANT_WAVEFRONT, capi.VARIANT_MEGAKERNEL):
            r.clear()
            r.render(spp=3, max_depth=50, variant=variant)
            out.append(r.read_accum().view(np.uint32).copy())
        r.close()
    for a in out[1:]:
        assert np.array_equal(out[0], a)
