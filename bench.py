#!/usr/bin/env python
"""bench.py — headline benchmark of the render path (BASELINE.json: Mrays/s and 1080p spp/s).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload C2|C1|C3|C4|C5]

One "step" = one full render of the workload's frame (all samples per pixel) into the accumulation buffer plus
the RGBA8 resolve. N = 1: workload C2 of BASELINE.json (RTIOW final scene, 1920x1080, 1024 spp, depth 50).
N > 1 (launched by torchrun, one rank per GPU): samples are sharded over ranks (rank g renders sample indices
g, g+N, ...; every rank renders `spp` samples, so the frame carries N x spp samples: weak scaling), the per-rank
fixed-point (int64) accumulation buffers are summed with ONE NCCL reduce to rank 0 inside the timed region (integer
addition is order-free, so the N-GPU image is bit-identical to the 1-GPU image of the same samples), rank 0 resolves.

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
    from cudaraytracer_b200 import Scene, capi
    kind_name, hint, W, H, spp, depth, desc = WORKLOADS[args.workload]
    scene = Scene.generate(getattr(capi, kind_name), 1984, hint)
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


def run_gpu_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    from cudaraytracer_b200 import Renderer, Scene, capi

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

    kind_name, hint, W, H, spp, depth, desc = WORKLOADS[args.workload]
    if args.spp:
        spp = args.spp
    scene = Scene.generate(getattr(capi, kind_name), 1984, hint)
    cam = scene.default_camera(W, H)
    # a non-default torch stream: our kernels, torch's fills and NCCL all run on it, and torch events time it
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    r = Renderer(local)
    r.set_stream(stream.cuda_stream)
    r.upload(scene)
    r.set_camera(cam)
    r.resize(W, H)
    # the progressive accumulation buffer: int64 x {R, G, B, samples}, Q31.32 sums (order-free, so the NCCL sum is exact)
    accum = torch.zeros((H, W, 4), dtype=torch.int64, device=dev)
    peer = None
    tiles = args.closeout == "tiles"
    if args.closeout in ("peer", "tiles"):
        # the ranks keep their library-owned buffers, mapped into each other (CUDA IPC): no reduce-to-root (multigpu.PeerFrame)
        from cudaraytracer_b200.multigpu import PeerFrame
        peer = PeerFrame(r, rank, world)
    else:
        r.set_accum_device_buffer(accum.data_ptr())
    rgba = torch.zeros((H, W), dtype=torch.int32, device=dev)
    host_rgba = torch.zeros((H, W), dtype=torch.int32).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2: written between timed steps
    stats = r.bvh_stats()
    variant = {"auto": capi.VARIANT_AUTO, "mega": capi.VARIANT_MEGAKERNEL, "wavefront": capi.VARIANT_WAVEFRONT,
               "pipeline": capi.VARIANT_PIPELINE}[args.variant]
    # what AUTO resolves to (crt_api.cu: pipelined persistent kernel from 256 primitives up, megakernel below)
    resolved = args.variant if args.variant != "auto" else ("pipeline" if scene.n_hittables >= 256 else "mega")
    kernel_name = {"mega": "k_render_mega", "wavefront": "k_wf_extend + k_wf_shade", "pipeline": "k_render_pipe"}[resolved]

    def device_step():
        """Scene resident in HBM: clear + render this rank's samples (+ reduce) + resolve on rank 0."""
        if tiles:  # image-tile sharding: ALL samples of this rank's row slabs; nothing to sum, 4 B/pixel to the root's frame
            r.clear()
            r.render(spp=spp * world, max_depth=depth, variant=variant, row_tiles=capi.row_tiles(rank, world, args.tile_rows))
            peer.close_frame_tiles(args.tile_rows)
            return
        if peer is not None:
            r.clear()
            r.render(spp=spp, max_depth=depth, spp_begin=rank, spp_stride=world, variant=variant)
            peer.close_frame()  # fence, sum + resolve of this rank's rows into the root's frame over NVLink, fence
            return
        accum.zero_()
        r.render(spp=spp, max_depth=depth, spp_begin=rank, spp_stride=world, variant=variant)
        if world > 1:
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
        if rank == 0:
            r.resolve_device(rgba.data_ptr())

    def e2e_step():
        """Through the C ABI with HOST buffers: scene upload (BVH build + H2D), camera, render, (reduce), resolve, D2H."""
        t0 = time.perf_counter()
        r.upload(scene)
        t1 = time.perf_counter()
        r.set_camera(cam)
        if peer is not None:
            r.clear()
            if tiles:
                r.render(spp=spp * world, max_depth=depth, variant=variant, row_tiles=capi.row_tiles(rank, world, args.tile_rows))
                peer.close_frame_tiles(args.tile_rows)
            else:
                r.render(spp=spp, max_depth=depth, spp_begin=rank, spp_stride=world, variant=variant)
                peer.close_frame()
            if rank == 0:
                r.read_rgba(host_rgba.numpy().view(np.uint32))  # D2H of the root's frame (synchronises the root's stream)
            return
        accum.zero_()
        r.render(spp=spp, max_depth=depth, spp_begin=rank, spp_stride=world, variant=variant)
        if os.environ.get("BENCH_DEBUG"):
            print(f"[e2e] upload {1e3 * (t1 - t0):.1f} ms, render {r.last_render_ms():.1f} ms (library events)", file=sys.stderr)
        if world > 1:
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
        if rank == 0:
            r.resolve_device(rgba.data_ptr())
            host_rgba.copy_(rgba, non_blocking=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(step_fn, steps, warmup):
        for _ in range(warmup):
            step_fn()
        barrier()
        r.reset_counters()
        times = []
        for _ in range(steps):
            flush.fill_(1)  # L2 flush between timed iterations (not timed)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            step_fn()
            e1.record(stream)
            barrier()
            ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)  # max over ranks
            times.append(float(ms.item()))
        c = r.counters()
        cnt = torch.tensor([c.rays, c.paths, c.kernel_launches], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        return times, [int(x) for x in cnt.tolist()]

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    times, (rays, paths, launches) = timed(device_step, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    total_s = sum(times) / 1e3
    value = rays / total_s / 1e6
    ms_per_step = sum(times) / len(times)
    L = rays / paths

    e_times, (e_rays, _, _) = timed(e2e_step, args.steps, 1)
    e2e_value = e_rays / (sum(e_times) / 1e3) / 1e6

    line = None
    if rank == 0:
        img = host_rgba.numpy().view(np.uint32)
        if not ((img >> 24) == 255).all():
            raise SystemExit("bench.py: resolved frame is incomplete")
        # roofline of the dominant kernel (k_render_mega: > 99 % of the step): algorithmic flops per ray measured
        # on our own BVH by a counting launch (F_ray = 21 N_box + 26 N_sph + 50)
        r.reset_counters()
        r.clear() if peer is not None else accum.zero_()
        r.render(spp=min(spp, 4), max_depth=depth, flags=capi.FLAG_COUNT_TESTS, variant=capi.VARIANT_MEGAKERNEL)
        cc = r.counters()
        n_box, n_sph = cc.box_tests / cc.rays, cc.prim_tests / cc.rays
        f_ray = 21.0 * n_box + 26.0 * n_sph + 50.0
        peaks, peaks_src = measured_peaks()
        sm_mhz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
        peak_tflops = FP32_LANES * 2 * sm_mhz * 1e6 / 1e12
        # kernel-only duration: the render launch of the last step, measured by the library's own events
        r.reset_counters()
        r.clear() if peer is not None else accum.zero_()
        r.render(spp=spp, max_depth=depth, spp_begin=rank, spp_stride=world, variant=variant)
        k_ms = r.last_render_ms()
        kc = r.counters()
        achieved = kc.rays * f_ray / (k_ms / 1e3) / 1e12
        # algorithmic DRAM bytes per ray: mega = one 32-byte pixel read + write per launch; pipeline/wavefront = three 8-byte
        # reductions per finished path into the accumulation buffer (+ wavefront: ~160 B of path state per ray)
        rpp = kc.rays / kc.paths
        b_ray = {"mega": 64.0 / rpp / spp, "pipeline": 24.0 / rpp, "wavefront": 24.0 / rpp + 160.0}[resolved]
        hbm_gbs = kc.rays * b_ray / (k_ms / 1e3) / 1e9
        traffic = None  # dram__bytes_read.sum + dram__bytes_write.sum of this kernel at this config, from the committed ncu capture
        issue = None    # what actually bounds the kernel: issue-slot and lane utilisation of the same capture
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
                                  "FP32/ALU issue bound (SURVEY.md 8d), the HBM fraction is given beside it",
                    "issue": issue, "flop_per_ray": f_ray, "box_tests_per_ray": n_box, "prim_tests_per_ray": n_sph,
                    "peak_source": f"148 SMs x 128 lanes x 2 x {sm_mhz:.0f} MHz (SM clock sampled during the timed region)",
                    "hbm": {"achieved": hbm_gbs, "peak": peaks.get("hbm_gbs"), "unit": "GB/s", "frac": hbm_gbs / peaks.get("hbm_gbs", 6650.0),
                            "bytes_per_ray": b_ray, "peak_source": peaks_src}}
        # CPU baseline: the reference's own CPU code for the path, bounded sample, rank 0 at N = 1 only
        cpu = None
        if world == 1 and not args.no_cpu:
            threads = os.cpu_count() or 1
            sample_spp = cpu_sample_spp(scene, cam, W, H, depth, spp, args.cpu_spp, threads)
            kind, mr, dt, _ = time_cpu_baseline(scene, cam, W, H, depth, sample_spp, threads)
            cpu = {"value": mr, "unit": "Mrays/s", "cores": threads, "kind": kind, "seconds": dt,
                   "sample": f"{W}x{H} at {sample_spp} spp of {spp} (depth {depth})"}
        # The reference's own CUDA kernel on this GPU, bounded sample; reported beside, not timed inside
        refk = None
        if world == 1 and not args.no_refkernel:
            import oracle_lib as O
            if scene.n_hittables > REFERENCE_MAX_HITTABLES:
                refk = {"unavailable": "the reference traversal's stack[16] overflows at this scene size (Hittable.cuh:399)"}
            elif O.have_refkernel():
                rk_spp = max(1, min(spp, args.refkernel_spp))
                _, ms = O.refkernel_render(scene, cam, W, H, rk_spp, depth, reps=3)
                covered = (W // 16 * 16) * (H // 16 * 16)
                best = min(ms)
                refk = {"value": covered * rk_spp * L / (best / 1e3) / 1e6, "unit": "Mrays/s", "ms": ms, "spp": rk_spp,
                        "pixels_covered": covered, "rays_per_path_assumed": L,
                        "how": "Kernel.cu unmodified, nvcc -O3 -use_fast_math sm_100a, wall clock around LaunchKernel, best of 3"}
            else:
                refk = {"unavailable": "oracle/_ref/libcrt_refkernel.so not built"}
        line = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic",
                "config": {"workload": f"{args.workload}: {desc}", "scene_hittables": scene.n_hittables, "bvh_nodes": int(stats.n_nodes),
                           "bvh_builder": {1: "host binned SAH (threaded)", 2: "GPU LBVH"}.get(int(stats.builder), "?"),
                           "variant": args.variant, "variant_resolved": resolved, "spp_per_rank": spp, "spp_total": spp * world,
                           "parallelism": f"spp-sharded x{world}" + ("" if world == 1 else ", one NCCL reduce (int64 sum) of the fixed-point accumulation buffer" if peer is None
                                                                     else ", close-out in one kernel per rank over peer memory (sum + resolve of its rows, no reduce-to-root)" if not tiles
                                                                     else f", image-tile sharded instead (slabs of {args.tile_rows} rows, every rank resolves its rows into the root's frame over peer memory)"),
                           "closeout": args.closeout,
                           "l2": "256 MB buffer written between timed steps (L2 flush)", "rays_per_path": L},
                "spp_per_s_1080p": (spp * world) / (ms_per_step / 1e3) * (W * H) / (1920 * 1080),
                "roofline": roofline, "cpu_baseline": cpu,
                "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": int(stats.h2d_bytes) * world,
                        "d2h_bytes_per_step": W * H * 4, "ms_per_step": sum(e_times) / len(e_times)},
                "gpu_launches": launches, "clocks": clocks, "reference_kernel": refk}
    barrier()
    if peer is not None:
        peer.release()
    r.close()
    if world > 1:
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line))
    return 0


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
