"""GPU test of the spp-sharded path on ONE GPU: G 'ranks' (contexts) render their sample shards into separate int64
buffers, the buffers are summed like the NCCL reduce does, and the result must be bit-identical to the single render."""
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
