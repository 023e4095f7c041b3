#!/usr/bin/env python
"""bench.py — headline benchmark of the render path (BASELINE.json: Mrays/s and 1080p spp/s).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload C2|C1|C3|C4|C5]

One "step" = one full render of the workload's frame (all samples per pixel) into the accumulation buffer plus
the RGBA8 resolve. N = 1: workload C2 of BASELINE.json (RTIOW final scene, 1920x1080, 1024 spp, depth 50).
N > 1 (launched by torchrun, one rank per GPU): samples are sharded over ranks (rank g renders sample indices
g, g+N, ...; every rank renders `spp` samples, so the frame carries N x spp samples: weak scaling), the per-rank
fixed-point (int64) accumulation buffers are summed with ONE NCCL reduce to rank 0 inside the timed region (integer
addition is order-free, so the N-GPU image is bit-identical to the 1-GPU image of the same samples), rank 0 resolves.

With the default workload the line also carries `c4_strong`: BASELINE.json's scaling config (C4: 1 M spheres, 3840x2160, 256 spp
IN TOTAL sharded over the N ranks), device-timed and end to end, with its efficiency against ONE GPU measured in the same run,
and `parity`: primary-hit IDs of the timed scene at the timed resolution against the oracle (and, for N > 1, the N-GPU frame
against one GPU rendering the same samples), computed outside the timed regions; a mismatch fails the run.

Prints ONE JSON line (rank 0). `value` is device-timed (CUDA events on the stream the kernels run on) with the
scene resident in HBM; `e2e` goes through the C ABI with HOST buffers (scene upload + BVH build + render + resolve
+ D2H of the frame) every step. `--impl reference` times the reference's own CPU code for the same path
(oracle/_ref/libcrt_refhost.so: its headers compiled for the host; else the plain-C oracle port) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # name: (scene kind, size_hint, width, height, spp, depth, description)
    "C1": ("SCENE_RTIOW", 0, 400, 225, 16, 50, "RTIOW final scene 400x225, 16 spp, depth 50"),
    "C2": ("SCENE_RTIOW", 0, 1920, 1080, 1024, 50, "RTIOW final scene 1920x1080, 1024 spp, depth 50"),
    "C3": ("SCENE_TEXTURED", 0, 1920, 1080, 512, 50, "textured scene 1920x1080, 512 spp, depth 50"),
    "C4": ("SCENE_GRID_1M", 0, 3840, 2160, 256, 50, "1M procedural spheres 3840x2160, 256 spp, depth 50"),
    "C5": ("SCENE_STRESS_10K", 0, 7680, 4320, 8, 100, "10k dielectric/metal spheres 7680x4320, 8 spp per pass, depth 100"),
}
FP32_LANES = 148 * 128  # SMs x FP32 lanes per SM on B200


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


class ClockSampler:
    """Samples nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(power), "samples": len(sm),
                "reasons": sorted(reasons)}


# BVHNode::Hit keeps 16 stack entries with no overflow check (Hittable.cuh:399): the reference's own code, host or CUDA build,
# smashes its stack on scenes much above 60 k hittables (SURVEY.md appendix probes). Above this size the CPU baseline is the
# oracle port (same algorithm, 64 entries) and the reference CUDA kernel is reported as unavailable.
# Image::value casts u * width with u = NaN at sphere poles (Texture.cuh:88-89, acos in Hittable.cuh:119-125): defined on the
# GPU (-> 0), undefined on the host, where the reference's headers segfault on the textured scene from 4 spp up. Scenes with
# image textures therefore also use the port, which applies the GPU's conversion (oracle/oracle.c f2i_gpu).
REFERENCE_MAX_HITTABLES = 60000


def cpu_reference_lib(scene):
    import oracle_lib as O
    if O.have_refhost() and scene.n_hittables <= REFERENCE_MAX_HITTABLES and int(scene.desc.n_images) == 0:
        return "reference", O.refhost_render
    return "port", O.oracle_render


def time_cpu_baseline(scene, cam, W, H, depth, sample_spp, threads):
    """The reference's CPU implementation of the path on the host cores, on a bounded sample of the workload."""
    kind, fn = cpu_reference_lib(scene)
    t0 = time.perf_counter()
    out = fn(scene, cam, W, H, sample_spp, depth, threads=threads, want_rgba=True, want_mean=False)
    dt = time.perf_counter() - t0
    return kind, out["rays"] / dt / 1e6, dt, out["rays"]


def cpu_sample_spp(scene, cam, W, H, depth, spp, cap, threads, budget_s=10.0):
    """spp of the bounded CPU sample: one calibration pass at 1 spp (also the warm-up of the host code), then as many spp as
    fit into about budget_s seconds, at most `cap` and at most the workload's own spp."""
    _, _, t1, _ = time_cpu_baseline(scene, cam, W, H, depth, 1, threads)
    return max(1, min(spp, cap, int(budget_s / max(t1, 1e-3))))


def run_reference_arm(args):
    """--impl reference: the reference's own CPU code, all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from cudaraytracer_b200 import Scene, capi  # ctypes declarations only: the scene comes from the HOST-ONLY generator library
    kind_name, hint, W, H, spp, depth, desc = WORKLOADS[args.workload]
    scene = Scene.generate(getattr(capi, kind_name), 1984, hint, host_only=True)  # libcrt_scenegen.so; libcrt.so stays unmapped
    cam = scene.default_camera(W, H)
    threads = os.cpu_count() or 1
    # bounded sample: the full frame at a reduced spp (the rate in Mrays/s does not depend on spp)
    sample_spp = cpu_sample_spp(scene, cam, W, H, depth, spp, args.cpu_spp, threads)
    rates, times = [], []
    kind = "port"
    for i in range(args.warmup + args.steps):
        kind, mr, dt, rays = time_cpu_baseline(scene, cam, W, H, depth, sample_spp, threads)
        if i >= args.warmup:
            rates.append(mr); times.append(dt)
    value = sum(rates) / len(rates)
    sample = f"{W}x{H} at {sample_spp} spp of {spp} (depth {depth}), one full-frame pass per step"
    line = {"impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {desc}", "sample": sample},
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "spp_per_s_1080p": value * 1e6 / (rays / sample_spp) * (W * H) / (1920 * 1080), "gpu_launches": 0}
    print(json.dumps(line))
    return 0


class Leg:
    """One workload on this rank's GPU: scene resident, the `device_step` / `e2e_step` pair of the contract, and their timing.

    scaling "weak": every rank renders the workload's spp (the frame carries N x spp samples); "strong": the workload's spp in
    total, sample g, g + N, ... on rank g. Close-outs: "nccl" = ONE NCCL int64 sum-reduce of the accumulation buffer to rank 0 +
    resolve there; "peer" = sum + resolve of a rank's rows in one kernel over peer memory; "tiles" = image-row-tile sharding
    (each rank renders all samples of its slabs and resolves them into the root's frame)."""

    def __init__(self, args, torch, dist, dev, stream, rank, world, local, workload, scaling, closeout, solo=False):
        import numpy as np
        from cudaraytracer_b200 import Renderer, Scene, capi
        self.np, self.torch, self.dist, self.capi = np, torch, dist, capi
        self.args, self.dev, self.stream, self.rank, self.local = args, dev, stream, rank, local
        self.world = 1 if solo else world          # ranks that share THIS leg's frame (solo: rank 0 renders everything alone)
        self.group_world = world                   # ranks of the job (barriers, max over ranks)
        self.workload, self.scaling, self.closeout = workload, scaling, closeout if self.world > 1 else "nccl"
        kind_name, hint, W, H, spp, depth, desc = WORKLOADS[workload]
        if args.spp and workload == args.workload:
            spp = args.spp
        self.W, self.H, self.depth, self.desc, self.spp_workload = W, H, depth, desc, spp
        self.scene = Scene.generate(getattr(capi, kind_name), 1984, hint)
        self.cam = self.scene.default_camera(W, H)
        self.textures = None
        if int(self.scene.desc.n_images) > 0:
            self.textures = "asset files (CRT_ASSET_DIR)" if os.environ.get("CRT_ASSET_DIR") else "procedural texels of the assets' sizes (no CRT_ASSET_DIR on this box)"
        r = self.r = Renderer(local)
        r.set_stream(stream.cuda_stream)
        r.upload(self.scene)
        r.set_camera(self.cam)
        self.lens = LENS.get(workload)
        if self.lens:
            r.set_lens(*self.lens)
        r.resize(W, H)
        self.stats = r.bvh_stats()
        if scaling == "strong":
            self.spp_rank = len(range(rank if not solo else 0, spp, self.world))
            self.spp_total = spp
        else:
            self.spp_rank = spp
            self.spp_total = spp * self.world
        self.rank_in_leg = 0 if solo else rank
        self.accum = torch.zeros((H, W, 4), dtype=torch.int64, device=dev)
        self.peer = None
        self.tiles = self.closeout == "tiles"
        if self.closeout in ("peer", "tiles"):
            from cudaraytracer_b200.multigpu import PeerFrame
            self.peer = PeerFrame(r, rank, world)
        else:
            r.set_accum_device_buffer(self.accum.data_ptr())
        self.rgba = torch.zeros((H, W), dtype=torch.int32, device=dev)
        self.host_rgba = torch.zeros((H, W), dtype=torch.int32).pin_memory()
        self.variant = {"auto": capi.VARIANT_AUTO, "mega": capi.VARIANT_MEGAKERNEL, "wavefront": capi.VARIANT_WAVEFRONT,
                        "pipeline": capi.VARIANT_PIPELINE}[args.variant]
        self.resolved = args.variant if args.variant != "auto" else ("pipeline" if self.scene.n_hittables >= 256 else "mega")
        self.active = (not solo) or rank == 0

    # -- one frame: this rank's samples (+ close-out) + resolve on rank 0 --
    def _render_and_close(self):
        r, capi = self.r, self.capi
        if self.tiles:
            r.clear()
            r.render(spp=self.spp_total, max_depth=self.depth, variant=self.variant, row_tiles=capi.row_tiles(self.rank, self.world, self.args.tile_rows))
            self.peer.close_frame_tiles(self.args.tile_rows)
            return
        if self.peer is not None:
            r.clear()
            if self.spp_rank:
                r.render(spp=self.spp_rank, max_depth=self.depth, spp_begin=self.rank_in_leg, spp_stride=self.world, variant=self.variant)
            self.peer.close_frame()
            return
        self.accum.zero_()
        if self.spp_rank:
            r.render(spp=self.spp_rank, max_depth=self.depth, spp_begin=self.rank_in_leg, spp_stride=self.world, variant=self.variant)
        if self.world > 1:
            self.dist.reduce(self.accum, dst=0, op=self.dist.ReduceOp.SUM)
        if self.rank_in_leg == 0:
            r.resolve_device(self.rgba.data_ptr())

    def device_step(self):
        if self.active:
            self._render_and_close()

    def e2e_step(self):
        """Through the C ABI with HOST buffers: the scene description and the camera go in (crt_scene_upload: BVH build + H2D,
        or a content-hash check when it is the scene already on the device), the RGBA8 frame comes back (D2H on rank 0)."""
        if not self.active:
            return
        self.r.upload(self.scene)
        self.r.set_camera(self.cam)
        self._render_and_close()
        if self.rank_in_leg == 0:
            if self.peer is not None:
                self.r.read_rgba(self.host_rgba.numpy().view(self.np.uint32))
            else:
                self.host_rgba.copy_(self.rgba, non_blocking=True)

    def barrier(self):
        if self.group_world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def timed(self, step_fn, steps, warmup, flush):
        torch, dist = self.torch, self.dist
        for _ in range(warmup):
            step_fn()
        self.barrier()
        self.r.reset_counters()
        times = []
        for _ in range(steps):
            flush.fill_(1)  # L2 flush between timed iterations (not timed)
            self.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(self.stream)
            step_fn()
            e1.record(self.stream)
            self.barrier()
            ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=self.dev)
            if self.group_world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)  # max over ranks
            times.append(float(ms.item()))
        c = self.r.counters()
        cnt = torch.tensor([c.rays, c.paths, c.kernel_launches], dtype=torch.int64, device=self.dev)
        if self.group_world > 1:
            dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        return times, [int(x) for x in cnt.tolist()]

    def frame_sha1(self):
        import hashlib
        self.torch.cuda.synchronize(self.dev)
        return hashlib.sha1(self.host_rgba.numpy().tobytes()).hexdigest()[:16]

    def parity_ids(self):
        """Primary-hit IDs of THIS scene at THIS resolution against the oracle (outside every timed region; rank 0 only)."""
        import oracle_lib as O
        np = self.np
        n_px = self.W * self.H
        big = self.scene.n_hittables > REFERENCE_MAX_HITTABLES
        if big:      # the oracle's brute-force closest hit over ALL hittables, on every 64th pixel of every 60th row
            stride, accel, what = (64, 60), O.BRUTE_FORCE, "oracle brute force over all hittables"
        elif n_px > 4 * 1920 * 1080:
            stride, accel, what = (16, 16), O.BVH_REFERENCE, "oracle, reference BVH (Hittable.cuh:303-439)"
        else:
            stride, accel, what = (1, 1), O.BVH_REFERENCE, "oracle, reference BVH (Hittable.cuh:303-439)"
        t0 = time.perf_counter()
        ids = self.r.primary_ids(self.capi.JITTER_CENTER)
        ref = O.oracle_render(self.scene, self.cam, self.W, self.H, 0, 5, want_ids=True, want_rgba=False, want_mean=False, accel=accel, stride=stride)["ids"]
        got, want = ids[::stride[1], ::stride[0]], ref[::stride[1], ::stride[0]]
        return {"checker": what, "resolution": [self.W, self.H], "pixel_stride": list(stride), "pixels": int(got.size),
                "ids_mismatch": int((got != want).sum()), "distinct_hittables": int(len(np.unique(got))),
                "seconds": round(time.perf_counter() - t0, 2)}

    def close(self):
        if self.peer is not None:
            self.peer.release()
        self.r.close()


# thin lens of the workloads that ask for defocus (BASELINE.json config 3: "defocus blur"; SURVEY §8d C3: aperture 0.1, focus 10)
LENS = {"C3": (0.1, 10.0)}


def run_gpu_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    from cudaraytracer_b200 import capi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    # a non-default torch stream: our kernels, torch's fills and NCCL all run on it, and torch events time it
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2: written between timed steps

    leg = Leg(args, torch, dist, dev, stream, rank, world, local, args.workload, args.scaling, args.closeout)
    r, scene, cam, W, H, depth, spp = leg.r, leg.scene, leg.cam, leg.W, leg.H, leg.depth, leg.spp_rank
    stats, resolved, peer = leg.stats, leg.resolved, leg.peer
    kernel_name = {"mega": "k_render_mega", "wavefront": "k_wf_extend + k_wf_shade", "pipeline": "k_render_pipe"}[resolved]

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    times, (rays, paths, launches) = leg.timed(leg.device_step, args.steps, args.warmup, flush)
    if args.variant == "auto":  # what AUTO ran: measured by the library where both kernels can stage the scene (crt_last_variant)
        resolved = {capi.VARIANT_MEGAKERNEL: "mega", capi.VARIANT_WAVEFRONT: "wavefront", capi.VARIANT_PIPELINE: "pipeline"}.get(r.last_variant(), resolved)
        kernel_name = {"mega": "k_render_mega", "wavefront": "k_wf_extend + k_wf_shade", "pipeline": "k_render_pipe"}[resolved]
    clocks = sampler.stop() if rank == 0 else None
    total_s = sum(times) / 1e3
    value = rays / total_s / 1e6
    ms_per_step = sum(times) / len(times)
    L = rays / paths

    e_times, (e_rays, _, _) = leg.timed(leg.e2e_step, args.steps, 1, flush)
    e2e_value = e_rays / (sum(e_times) / 1e3) / 1e6
    e2e_stats = r.bvh_stats()  # of the last e2e step: h2d_bytes == 0 and cache_hits > 0 when the unchanged scene was recognised
    e2e_full = None
    if True:  # the same step with the unchanged-scene cache off: every step validates, builds the BVH and copies the whole description
        r.set_scene_cache(False)
        f_times, (f_rays, _, _) = leg.timed(leg.e2e_step, max(2, args.steps // 2), 1, flush)
        r.set_scene_cache(True)
        e2e_full = {"value": f_rays / (sum(f_times) / 1e3) / 1e6, "unit": "Mrays/s", "ms_per_step": sum(f_times) / len(f_times),
                    "h2d_bytes_per_step": int(r.bvh_stats().h2d_bytes) * world,
                    "note": "unchanged-scene cache off (crt_set_scene_cache(0)): validation + BVH build + H2D of the whole description every step"}
        r.upload(scene)

    line = None
    parity = None
    # ---- parity, outside every timed region, on the workload that was timed ----
    if rank == 0:
        img = leg.host_rgba.numpy().view(np.uint32)
        if not ((img >> 24) == 255).all():
            raise SystemExit("bench.py: resolved frame is incomplete")
        parity = {"primary_ids": leg.parity_ids()}
        if parity["primary_ids"]["ids_mismatch"] != 0:
            raise SystemExit(f"bench.py: primary-hit IDs differ from the oracle: {parity['primary_ids']}")
    if world > 1:
        # the N-GPU frame against ONE GPU rendering the same sample set alone (small spp): the same bytes
        small = Leg(args, torch, dist, dev, stream, rank, world, local, args.workload, "strong", args.closeout)
        small.spp_total = 2 * world
        small.spp_rank = 2
        small.e2e_step()
        small.barrier()
        sha_n = small.frame_sha1() if rank == 0 else None
        small.close()
        solo = Leg(args, torch, dist, dev, stream, rank, world, local, args.workload, "strong", "nccl", solo=True)
        solo.spp_total = solo.spp_rank = 2 * world
        solo.e2e_step()
        solo.barrier()
        sha_1 = solo.frame_sha1() if rank == 0 else None
        solo.close()
        if rank == 0:
            parity["multi_gpu_frame"] = {"what": f"RGBA8 frame of {2 * world} spp: {world} GPUs (2 spp each, close-out {args.closeout}) vs rank 0 alone",
                                         "sha1_n_gpus": sha_n, "sha1_one_gpu": sha_1, "identical": sha_n == sha_1}
            if sha_n != sha_1:
                raise SystemExit(f"bench.py: the {world}-GPU frame differs from the 1-GPU frame of the same samples: {parity['multi_gpu_frame']}")

    if rank == 0:
        accum = leg.accum
        # roofline of the dominant kernel: algorithmic flops per ray measured on our own BVH by a counting launch of the
        # MEGAKERNEL build (F_ray = 21 N_box + 26 N_sph + 50). The timed kernel is `kernel_name`; it walks the same
        # tree in the same per-ray order, so the counts are its counts.
        r.reset_counters()
        r.clear() if peer is not None else accum.zero_()
        r.render(spp=min(max(spp, 1), 4), max_depth=depth, flags=capi.FLAG_COUNT_TESTS, variant=capi.VARIANT_MEGAKERNEL)
        cc = r.counters()
        n_box, n_sph = cc.box_tests / cc.rays, cc.prim_tests / cc.rays
        f_ray = 21.0 * n_box + 26.0 * n_sph + 50.0
        peaks, peaks_src = measured_peaks()
        sm_mhz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
        peak_tflops = FP32_LANES * 2 * sm_mhz * 1e6 / 1e12
        # kernel-only duration: the render launch of one more step, measured by the library's own events on its stream
        r.reset_counters()
        r.clear() if peer is not None else accum.zero_()
        r.render(spp=max(spp, 1), max_depth=depth, spp_begin=rank, spp_stride=world, variant=leg.variant)
        k_ms = r.last_render_ms()
        kc = r.counters()
        achieved = kc.rays * f_ray / (k_ms / 1e3) / 1e12
        # algorithmic DRAM bytes per ray: mega = one 32-byte pixel read + write per launch; pipeline/wavefront = three 8-byte
        # reductions per finished path into the accumulation buffer (+ wavefront: ~160 B of path state per ray)
        rpp = kc.rays / kc.paths
        b_ray = {"mega": 64.0 / rpp / max(spp, 1), "pipeline": 24.0 / rpp, "wavefront": 24.0 / rpp + 160.0}[resolved]
        hbm_gbs = kc.rays * b_ray / (k_ms / 1e3) / 1e9
        traffic = None  # dram bytes per launch and issue / lane figures of this kernel, from the committed ncu capture of THIS build
        issue = None
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                tj = json.load(f)
            traffic = tj.get(f"{args.workload}:{resolved}")
            issue = tj.get(f"{args.workload}:{resolved}_issue")
        except Exception:
            pass
        roofline = {"bound": "fp32", "achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s", "frac": achieved / peak_tflops,
                    "traffic": traffic, "kernel": kernel_name, "kernel_ms": k_ms, "kernel_share_of_step": k_ms / ms_per_step,
                    "bound_note": "no dense contraction and ~0.1 B/ray of DRAM traffic: neither tensor nor HBM bound; the kernel is "
                                  "instruction-issue / shared-memory-port bound (SURVEY.md 8d, DESIGN.md 3), the HBM fraction is given beside it",
                    "issue": issue, "flop_per_ray": f_ray, "box_tests_per_ray": n_box, "prim_tests_per_ray": n_sph,
                    "test_counts_from": "a counting launch of k_render_mega on the same tree (same per-ray visit order as the timed kernel)",
                    "peak_source": f"148 SMs x 128 lanes x 2 x {sm_mhz:.0f} MHz (SM clock sampled during the timed region)",
                    "hbm": {"achieved": hbm_gbs, "peak": peaks.get("hbm_gbs"), "unit": "GB/s", "frac": hbm_gbs / peaks.get("hbm_gbs", 6650.0),
                            "bytes_per_ray": b_ray, "peak_source": peaks_src}}
        # CPU baseline: the reference's own CPU code for the path, bounded sample, rank 0 at N = 1 only
        cpu = None
        if world == 1 and not args.no_cpu:
            threads = os.cpu_count() or 1
            sample_spp = cpu_sample_spp(scene, cam, W, H, depth, leg.spp_workload, args.cpu_spp, threads)
            kind, mr, dt, _ = time_cpu_baseline(scene, cam, W, H, depth, sample_spp, threads)
            cpu = {"value": mr, "unit": "Mrays/s", "cores": threads, "kind": kind, "seconds": dt,
                   "sample": f"{W}x{H} at {sample_spp} spp of {leg.spp_workload} (depth {depth})" + (", pinhole (the reference has no lens)" if leg.lens else "")}
        # The reference's own CUDA kernel on this GPU, bounded sample; reported beside, not timed inside
        refk = None
        if world == 1 and not args.no_refkernel:
            import oracle_lib as O
            if scene.n_hittables > REFERENCE_MAX_HITTABLES:
                refk = {"unavailable": "the reference traversal's stack[16] overflows at this scene size (Hittable.cuh:399)"}
            elif O.have_refkernel():
                rk_spp = max(1, min(leg.spp_workload, args.refkernel_spp))
                _, ms = O.refkernel_render(scene, cam, W, H, rk_spp, depth, reps=3)
                covered = (W // 16 * 16) * (H // 16 * 16)
                best = min(ms)
                refk = {"value": covered * rk_spp * L / (best / 1e3) / 1e6, "unit": "Mrays/s", "ms": ms, "spp": rk_spp,
                        "pixels_covered": covered, "rays_per_path_assumed": L,
                        "how": "Kernel.cu unmodified, nvcc -O3 -use_fast_math sm_100a, wall clock around LaunchKernel, best of 3"}
                refk["ratio_new_over_reference_kernel"] = value / refk["value"]
            else:
                refk = {"unavailable": "oracle/_ref/libcrt_refkernel.so not built"}
        tiles = leg.tiles
        config = {"workload": f"{args.workload}: {leg.desc}", "scene_hittables": scene.n_hittables, "bvh_nodes": int(stats.n_nodes),
                  "bvh_builder": {1: "host binned SAH (threaded)", 2: "GPU LBVH"}.get(int(stats.builder), "?"),
                  "variant": args.variant, "variant_resolved": resolved, "spp_per_rank": spp, "spp_total": leg.spp_total,
                  "parallelism": f"spp-sharded x{world}" + ("" if world == 1 else ", one NCCL reduce (int64 sum) of the fixed-point accumulation buffer" if peer is None
                                                            else ", close-out in one kernel per rank over peer memory (sum + resolve of its rows, no reduce-to-root)" if not tiles
                                                            else f", image-tile sharded instead (slabs of {args.tile_rows} rows, every rank resolves its rows into the root's frame over peer memory)"),
                  "closeout": leg.closeout,
                  "l2": "256 MB buffer written between timed steps (L2 flush)", "rays_per_path": L}
        if leg.lens:
            config["lens"] = {"aperture": leg.lens[0], "focus_dist": leg.lens[1]}
        if leg.textures:
            config["textures"] = leg.textures
        line = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": leg.scaling, "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": config,
                "spp_per_s_1080p": leg.spp_total / (ms_per_step / 1e3) * (W * H) / (1920 * 1080),
                "roofline": roofline, "cpu_baseline": cpu,
                "e2e": {"value": e2e_value, "unit": "Mrays/s",
                        "h2d_bytes_per_step": (int(e2e_stats.h2d_bytes) + 72) * world, "d2h_bytes_per_step": W * H * 4,
                        "ms_per_step": sum(e_times) / len(e_times),
                        "scene_upload": ("content hash of the description matched the scene on the device: no rebuild, no copy "
                                         f"(crt_bvh_stats.cache_hits = {int(e2e_stats.cache_hits)}); the camera (72 B) goes in every step"
                                         if int(e2e_stats.cache_hits) > 0 else "validated, BVH built and copied every step"),
                        "first_upload_h2d_bytes": int(stats.h2d_bytes)},
                "e2e_full_upload": e2e_full,
                "gpu_launches": launches, "clocks": clocks, "reference_kernel": refk, "parity": parity}
    leg.barrier()
    leg.close()

    # ---- north_star's scaling claim: C4 (1 M spheres, 3840x2160), 256 spp IN TOTAL sharded over the N ranks (strong scaling) ----
    c4 = None
    if args.workload == "C2" and not args.no_c4:
        c4 = c4_strong_leg(args, torch, dist, dev, stream, rank, world, local, flush)
    if world > 1:
        dist.destroy_process_group()
    if line is not None:
        line["c4_strong"] = c4
        print(json.dumps(line))
    return 0


def c4_strong_leg(args, torch, dist, dev, stream, rank, world, local, flush):
    steps = max(5, args.steps)
    closeout = args.c4_closeout
    leg = Leg(args, torch, dist, dev, stream, rank, world, local, "C4", "strong", closeout)
    times, (rays, paths, _) = leg.timed(leg.device_step, steps, 3, flush)
    e_times, (e_rays, _, _) = leg.timed(leg.e2e_step, steps, 1, flush)
    e_stats = leg.r.bvh_stats()
    leg.r.set_scene_cache(False)
    f_times, (f_rays, _, _) = leg.timed(leg.e2e_step, 3, 1, flush)
    leg.r.set_scene_cache(True)
    leg.r.upload(leg.scene)
    parity = None
    if rank == 0:
        parity = leg.parity_ids()
        if parity["ids_mismatch"] != 0:
            raise SystemExit(f"bench.py: C4 primary-hit IDs differ from the oracle's brute force: {parity}")
    ms, e_ms, f_ms = sum(times) / len(times), sum(e_times) / len(e_times), sum(f_times) / len(f_times)
    out = {"workload": WORKLOADS["C4"][6] + f", {leg.spp_total} spp in total = {leg.spp_total // world} per GPU on {world} GPU(s)", "scaling": "strong",
           "steps": steps, "ms_per_step": ms, "value": rays / (sum(times) / 1e3) / 1e6, "unit": "Mrays/s",
           "e2e": {"value": e_rays / (sum(e_times) / 1e3) / 1e6, "ms_per_step": e_ms,
                   "h2d_bytes_per_step": (int(e_stats.h2d_bytes) + 72) * world, "d2h_bytes_per_step": leg.W * leg.H * 4,
                   "scene_upload": "content-hash hit, camera only" if int(e_stats.cache_hits) > 0 else "full upload every step"},
           "e2e_full_upload": {"value": f_rays / (sum(f_times) / 1e3) / 1e6, "ms_per_step": f_ms,
                               "note": "unchanged-scene cache off: every rank validates, uploads and builds the LBVH of the 1 M spheres every step"},
           "closeout": leg.closeout, "bvh_builder": {1: "host binned SAH (threaded)", 2: "GPU LBVH"}.get(int(leg.stats.builder), "?"),
           "parity": parity}
    leg.barrier()
    leg.close()
    if world == 1:
        out["efficiency_vs_1gpu_same_run"] = {"device": 1.0, "e2e": 1.0, "e2e_full_upload": 1.0}
        out["limiter"] = "one GPU: nothing to share"
        return out
    # the 1-GPU reference of the SAME run: rank 0 renders all 256 spp alone, the other ranks wait at the barriers
    solo = Leg(args, torch, dist, dev, stream, rank, world, local, "C4", "strong", "nccl", solo=True)
    s_times, _ = solo.timed(solo.device_step, 2, 1, flush)
    se_times, _ = solo.timed(solo.e2e_step, 2, 1, flush)
    solo.r.set_scene_cache(False)
    sf_times, _ = solo.timed(solo.e2e_step, 2, 1, flush)
    solo.barrier()
    solo.close()
    s_ms, se_ms, sf_ms = sum(s_times) / len(s_times), sum(se_times) / len(se_times), sum(sf_times) / len(sf_times)
    out["one_gpu_same_run"] = {"ms_per_step": s_ms, "e2e_ms_per_step": se_ms, "e2e_full_upload_ms_per_step": sf_ms}
    out["efficiency_vs_1gpu_same_run"] = {"device": s_ms / (world * ms), "e2e": se_ms / (world * e_ms), "e2e_full_upload": sf_ms / (world * f_ms)}
    fixed = e_ms - ms
    out["limiter"] = (f"per-step fixed cost outside the render kernel: {fixed:.1f} ms of a {e_ms:.1f} ms e2e step "
                      f"(scene hash + camera + {leg.W}x{leg.H} RGBA8 D2H on rank 0; close-out {leg.closeout} is inside the {ms:.1f} ms device step)")
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="C2", choices=sorted(WORKLOADS))
    ap.add_argument("--variant", default="auto", choices=["auto", "mega", "wavefront", "pipeline"])
    ap.add_argument("--closeout", default="nccl", choices=["nccl", "peer", "tiles"],
                    help="multi-GPU frame close-out: NCCL reduce to rank 0 + resolve; sum + resolve over peer memory in one kernel per "
                         "rank; or image-tile sharding (each rank renders all samples of its row slabs and resolves them into the root's frame)")
    ap.add_argument("--tile-rows", type=int, default=16, help="rows per slab of --closeout tiles")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="N > 1: weak = every rank renders the workload's spp (default, what `value` is quoted on); strong = the workload's spp in total")
    ap.add_argument("--c4-closeout", default="nccl", choices=["nccl", "peer", "tiles"], help="close-out of the strong-scaled C4 leg")
    ap.add_argument("--no-c4", action="store_true", help="skip the strong-scaled C4 leg that accompanies the default workload")
    ap.add_argument("--spp", type=int, default=0, help="override the workload's samples per pixel (debug only)")
    ap.add_argument("--cpu-spp", type=int, default=128, help="upper bound of the spp of the bounded CPU-baseline sample (sized to about 10 s of host work)")
    ap.add_argument("--refkernel-spp", type=int, default=64, help="spp of the bounded reference-CUDA-kernel sample")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-refkernel", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "native":
        args.warmup = max(args.warmup, 3)  # timing rule: W >= 3
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
