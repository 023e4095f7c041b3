"""Peer-memory frame close-out against the NCCL reduce, under torchrun (one rank per GPU):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29513 tools/peer_check.py [W H spp scene]
Every rank renders its sample shard into its renderer's own buffer; (1) PeerFrame.close_frame() -> the root's frame;
(2) copies of the buffers are NCCL-reduced to the root and resolved there. The two frames must be identical, and identical to
what the root renders alone. Prints the device time of each close-out."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cudaraytracer_b200 import Renderer, Scene, capi
from cudaraytracer_b200.multigpu import PeerFrame, sample_shard

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
W = int(sys.argv[1]) if len(sys.argv) > 1 else 1920
H = int(sys.argv[2]) if len(sys.argv) > 2 else 1080
spp = int(sys.argv[3]) if len(sys.argv) > 3 else 8
kind = {"rtiow": capi.SCENE_RTIOW, "stress": capi.SCENE_STRESS_10K}[sys.argv[4] if len(sys.argv) > 4 else "rtiow"]
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
stream = torch.cuda.Stream(dev)
torch.cuda.set_stream(stream)
s = Scene.generate(kind, 1984)
r = Renderer(local)
r.set_stream(stream.cuda_stream)
r.upload(s); r.set_camera(s.default_camera(W, H)); r.resize(W, H)
pf = PeerFrame(r, rank, world)
b, st, c = sample_shard(rank, world, spp)


def ev():
    return torch.cuda.Event(enable_timing=True)


t_peer, t_nccl = [], []
mine = torch.zeros((H, W, 4), dtype=torch.int64, device=dev)
rgba = torch.zeros((H, W), dtype=torch.int32, device=dev)
for it in range(4):
    r.clear()
    if c:
        r.render(spp=c, max_depth=50, spp_begin=b, spp_stride=st)
    e0, e1, e2, e3 = ev(), ev(), ev(), ev()
    # NCCL path on a copy of this rank's buffer (the copy is outside the timed region)
    torch.cuda.synchronize(dev)
    import ctypes
    ctypes.CDLL("libcudart.so").cudaMemcpy(ctypes.c_void_p(mine.data_ptr()), ctypes.c_void_p(r.accum_ptr()), ctypes.c_size_t(mine.numel() * 8), 3)
    dist.barrier()
    e0.record(stream)
    pf.close_frame()
    e1.record(stream)
    dist.barrier()
    e2.record(stream)
    dist.reduce(mine, dst=0, op=dist.ReduceOp.SUM)
    if rank == 0:
        r.set_accum_device_buffer(mine.data_ptr())
        r.resolve_device(rgba.data_ptr())
        r.set_accum_device_buffer(0)
    e3.record(stream)
    torch.cuda.synchronize(dev)
    dist.barrier()
    tp = torch.tensor([e0.elapsed_time(e1), e2.elapsed_time(e3)], dtype=torch.float64, device=dev)
    dist.all_reduce(tp, op=dist.ReduceOp.MAX)
    t_peer.append(float(tp[0])); t_nccl.append(float(tp[1]))
ok = True
got = r.read_rgba() if rank == 0 else None
# image-tile sharding: every rank renders ALL samples of its row slabs and resolves them into the root's frame
dist.barrier()
t_tiles = []
for it in range(3):
    r.clear()
    r.render(spp=spp, max_depth=50, row_tiles=capi.row_tiles(rank, world, 16))
    torch.cuda.synchronize(dev)
    dist.barrier()
    e0, e1 = ev(), ev()
    e0.record(stream)
    pf.close_frame_tiles(16)
    e1.record(stream)
    torch.cuda.synchronize(dev)
    dist.barrier()
    tp = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(tp, op=dist.ReduceOp.MAX)
    t_tiles.append(float(tp[0]))
if rank == 0:
    got_tiles = r.read_rgba()
    want = rgba.cpu().numpy().view(np.uint32)
    solo = Renderer(local)
    solo.upload(s); solo.set_camera(s.default_camera(W, H)); solo.resize(W, H)
    solo.render(spp=spp, max_depth=50)
    alone = solo.resolve()
    ok = bool(np.array_equal(got, want) and np.array_equal(got, alone) and np.array_equal(got_tiles, alone) and ((got >> 24) == 255).all())
    print(f"peer_check world {world} {W}x{H} spp {spp}: frames identical (peer == nccl == tiles == one GPU): {ok}; "
          f"close-out ms peer {min(t_peer[1:]):.3f} nccl+resolve {min(t_nccl[1:]):.3f} tiles {min(t_tiles[1:]):.3f} "
          f"(all: {t_peer} / {t_nccl} / {t_tiles})", flush=True)
    solo.close()
dist.barrier()
pf.release()
r.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
