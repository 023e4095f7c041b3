// kernels.cuh — the __global__ entry points.
//   k_render_mega   persistent-path megakernel: one thread owns a pixel and walks all of its samples as ONE
//                   flattened bounce loop (a lane whose path ends starts its next sample at once instead of
//                   waiting for the longest path of the warp, which is what Kernel.cu:137-148 + color() do).
//   k_primary_ids   exact-arithmetic primary-ray closest hit -> hittable index (parity hook).
//   k_resolve       accumulation buffer -> packed RGBA8, bit-compatible with Kernel.cu:12-19,151-157.
//   k_clear         zero the accumulation buffer.
// Wavefront kernels live in wavefront.cuh.
#pragma once

#include "shade.cuh"

// 16x16 pixel tile per 256-thread CTA; each warp covers an 8x4 sub-tile so primary rays stay coherent.
__device__ __forceinline__ void tile_pixel(uint32_t tiles_x, int& x, int& y)
{
    const uint32_t tile = blockIdx.x;
    const int tx = tile % tiles_x, ty = tile / tiles_x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    x = tx * 16 + (warp & 1) * 8 + (lane & 7);
    y = ty * 16 + (warp >> 1) * 4 + (lane >> 3);
}

__device__ __forceinline__ void flush_counters(DevCounters* c, unsigned paths, unsigned rays, unsigned nbox, unsigned nprim, bool count)
{
    // warp-aggregated: one atomic per warp per counter
    for (int off = 16; off > 0; off >>= 1) {
        paths += __shfl_xor_sync(0xffffffffu, paths, off);
        rays += __shfl_xor_sync(0xffffffffu, rays, off);
        if (count) {
            nbox += __shfl_xor_sync(0xffffffffu, nbox, off);
            nprim += __shfl_xor_sync(0xffffffffu, nprim, off);
        }
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&c->paths, (unsigned long long)paths);
        atomicAdd(&c->rays, (unsigned long long)rays);
        if (count) {
            atomicAdd(&c->box_tests, (unsigned long long)nbox);
            atomicAdd(&c->prim_tests, (unsigned long long)nprim);
        }
    }
}

template <bool EXACT, bool COUNT, bool SMEM_ALL>
__global__ void __launch_bounds__(CRT_BLOCK_THREADS) k_render_mega(const __grid_constant__ RenderArgs A)
{
    extern __shared__ float4 smem[];
    const SceneView S = stage_scene(A.scene, smem);
    __syncthreads();

    int x, y;
    tile_pixel((A.cover_w + 15) / 16, x, y);
    const bool inside = (uint32_t)x < A.cover_w && (uint32_t)y < A.cover_h;
    const uint32_t pixel = (uint32_t)y * A.width + (uint32_t)x;

    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (inside) acc = A.accum[pixel];

    unsigned n_paths = 0, n_rays = 0, n_box = 0, n_prim = 0;
    uint32_t s = 0;
    uint32_t bounce = 0;
    bool need_gen = true;
    float3 o = f3(0.f, 0.f, 0.f), d = f3(0.f, 0.f, 1.f), thr = f3(1.f, 1.f, 1.f);
    uint32_t sample = 0;
    const uint32_t stride = A.spp_stride ? A.spp_stride : 1u;

    while (inside) {
        if (need_gen) {
            if (s >= A.spp_count) break;
            sample = A.spp_begin + s * stride;
            uint4 r = crt_philox4x32_10(make_uint4(pixel, sample, 0u, CRT_STREAM_CAMERA), A.seed_lo, A.seed_hi);
            float j1 = 0.5f, j2 = 0.5f;
            if (A.jitter == CRT_JITTER_RANDOM) { j1 = crt_uniform(r.x); j2 = crt_uniform(r.y); }
            camera_ray<EXACT>(A.cam, A.width, A.height, x, y, j1, j2, o, d);
            if (A.cam.lens_radius > 0.f) apply_lens(A.cam, crt_uniform(r.z), crt_uniform(r.w), o, d);
            thr = f3(1.f, 1.f, 1.f);
            bounce = 0;
            need_gen = false;
            n_paths++;
        }
        n_rays++;
        Hit h = closest_hit<EXACT, COUNT, SMEM_ALL>(S, o, d, 0.001f, 3.402823466e+38f, n_box, n_prim);
        bool finished;
        float3 add;
        if (h.leaf < 0) {
            add = thr * background(A.cam, d);
            finished = true;
        }
        else {
            uint4 r = crt_philox4x32_10(make_uint4(pixel, sample, bounce, CRT_STREAM_SCATTER), A.seed_lo, A.seed_hi);
            float4 xi = make_float4(crt_uniform(r.x), crt_uniform(r.y), crt_uniform(r.z), crt_uniform(r.w));
            bool alive = shade_hit<SMEM_ALL>(S, h, o, d, thr, add, xi);
            bounce++;
            finished = !alive;
            if (alive && bounce >= A.max_depth) { // exceeded recursion: black (Kernel.cu:79)
                finished = true;
                add = f3(0.f, 0.f, 0.f);
            }
        }
        if (finished) {
            acc.x += add.x; acc.y += add.y; acc.z += add.z; acc.w += 1.0f;
            s++;
            need_gen = true;
        }
    }
    if (inside) A.accum[pixel] = acc;
    flush_counters(A.counters, n_paths, n_rays, n_box, n_prim, COUNT);
}

struct IdArgs
{
    DevScene scene;
    DevCamera cam;
    int* out_ids;
    uint32_t width, height;
    uint32_t seed_lo, seed_hi;
    int jitter;
};

template <bool SMEM_ALL> __global__ void __launch_bounds__(CRT_BLOCK_THREADS) k_primary_ids(const __grid_constant__ IdArgs A)
{
    extern __shared__ float4 smem[];
    const SceneView S = stage_scene(A.scene, smem);
    __syncthreads();
    int x, y;
    tile_pixel((A.width + 15) / 16, x, y);
    if ((uint32_t)x >= A.width || (uint32_t)y >= A.height) return;
    const uint32_t pixel = (uint32_t)y * A.width + (uint32_t)x;
    float j1 = 0.5f, j2 = 0.5f;
    if (A.jitter == CRT_JITTER_RANDOM) {
        uint4 r = crt_philox4x32_10(make_uint4(pixel, 0u, 0u, CRT_STREAM_CAMERA), A.seed_lo, A.seed_hi);
        j1 = crt_uniform(r.x); j2 = crt_uniform(r.y);
    }
    float3 o, d;
    camera_ray<true>(A.cam, A.width, A.height, x, y, j1, j2, o, d);
    unsigned nb = 0, np = 0;
    Hit h = closest_hit<true, false, SMEM_ALL>(S, o, d, 0.001f, 3.402823466e+38f, nb, np);
    A.out_ids[pixel] = h.leaf < 0 ? -1 : __ldg(A.scene.prim_id + (h.leaf & CRT_LEAF_INDEX_MASK));
}

// Kernel.cu:151-157 + RgbToInt :12-19: col / spp -> 255 * sqrtf -> clamp [0,255] -> truncate -> A<<24|B<<16|G<<8|R.
// IEEE division and square root so the result is bit-identical to the host oracle for the same buffer.
__global__ void k_resolve(const float4* __restrict__ accum, uint32_t* __restrict__ out, uint32_t width, uint32_t cover_w, uint32_t cover_h)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cover_w * cover_h) return;
    uint32_t x = i % cover_w, y = i / cover_w;
    uint32_t pixel = y * width + x;
    float4 a = accum[pixel];
    uint32_t packed = 0xFF000000u;
    if (a.w > 0.f) {
        float r = mul_rn(255.0f, sqrt_rn(div_rn(a.x, a.w)));
        float g = mul_rn(255.0f, sqrt_rn(div_rn(a.y, a.w)));
        float b = mul_rn(255.0f, sqrt_rn(div_rn(a.z, a.w)));
        r = (r < 0.0f) ? 0.0f : ((r > 255.0f) ? 255.0f : r);
        g = (g < 0.0f) ? 0.0f : ((g > 255.0f) ? 255.0f : g);
        b = (b < 0.0f) ? 0.0f : ((b > 255.0f) ? 255.0f : b);
        packed = (255u << 24) | ((uint32_t)(int)b << 16) | ((uint32_t)(int)g << 8) | (uint32_t)(int)r;
    }
    out[pixel] = packed;
}

__global__ void k_clear(float4* __restrict__ accum, uint32_t n)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) accum[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// ------------------------------------------------------------------------------------------
// k_render_mega2 — megakernel with a FLAT per-lane state machine and thresholded lane refill.
//
// ncu of k_render_mega (profiles/r1) shows the traversal loop running with 17 of 32 lanes on average: every lane
// of a warp waits for the warp's longest traversal before any lane may shade. Here a lane is in one of three
// states — at a node (cur >= 0), at a leaf (cur < 0), or waiting to shade/regenerate (cur == IDLE) — and the warp
// runs shade/regenerate for the waiting lanes as soon as fewer than THRESH lanes are still traversing, instead of
// when none is. THRESH = 1 reproduces the old schedule; 32 refills as soon as any lane waits.
// Arithmetic, Philox keys and accumulation order are identical to k_render_mega, so the two are bit-identical.
// ------------------------------------------------------------------------------------------
#define CRT_IDLE ((int)0x80000000)

template <bool EXACT, bool COUNT, bool SMEM_ALL, int THRESH>
__global__ void __launch_bounds__(CRT_BLOCK_THREADS, 4) k_render_mega2(const __grid_constant__ RenderArgs A)
{
    extern __shared__ float4 smem[];
    const SceneView S = stage_scene(A.scene, smem);
    __syncthreads();

    int x, y;
    tile_pixel((A.cover_w + 15) / 16, x, y);
    const bool inside = (uint32_t)x < A.cover_w && (uint32_t)y < A.cover_h;
    const uint32_t pixel = (uint32_t)y * A.width + (uint32_t)x;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (inside) acc = A.accum[pixel];

    unsigned n_paths = 0, n_rays = 0, n_box = 0, n_prim = 0;
    uint32_t s = 0, bounce = 0, sample = 0;
    const uint32_t stride = A.spp_stride ? A.spp_stride : 1u;
    bool alive = inside && A.spp_count > 0;
    bool need_gen = true;
    float3 o = f3(0.f, 0.f, 0.f), d = f3(0.f, 0.f, 1.f), thr = f3(1.f, 1.f, 1.f);
    float3 id = f3(0.f, 0.f, 0.f), ood = f3(0.f, 0.f, 0.f);
    float a = 1.f, inv_a = 1.f;
    Hit h;
    h.t = 3.402823466e+38f;
    h.leaf = -1;
    int cur = CRT_IDLE, sp = 0;
    int stack[CRT_STACK_SIZE];
    const float t_min = 0.001f;

    while (true) {
        const bool trav = cur != CRT_IDLE;
        const bool want_s = alive && !trav;
        const unsigned mt = __ballot_sync(0xffffffffu, trav);
        const unsigned ms = __ballot_sync(0xffffffffu, want_s);
        if ((mt | ms) == 0u) break;
        if (ms != 0u && (mt == 0u || __popc(mt) < THRESH)) {
            if (want_s) {
                if (!need_gen && h.leaf < 0) { // the ray escaped: background (Kernel.cu:41-44)
                    float3 add = thr * background(A.cam, d);
                    acc.x += add.x; acc.y += add.y; acc.z += add.z; acc.w += 1.0f;
                    s++;
                    need_gen = true;
                }
                if (need_gen && s >= A.spp_count) {
                    alive = false;
                }
                else {
                    const bool gen = need_gen;
                    if (gen) sample = A.spp_begin + s * stride;
                    // one Philox block per lane, whichever state it is in
                    const uint4 r = crt_philox4x32_10(make_uint4(pixel, sample, gen ? 0u : bounce, gen ? CRT_STREAM_CAMERA : CRT_STREAM_SCATTER),
                                                      A.seed_lo, A.seed_hi);
                    const float4 xi = make_float4(crt_uniform(r.x), crt_uniform(r.y), crt_uniform(r.z), crt_uniform(r.w));
                    if (gen) {
                        float j1 = 0.5f, j2 = 0.5f;
                        if (A.jitter == CRT_JITTER_RANDOM) { j1 = xi.x; j2 = xi.y; }
                        camera_ray<EXACT>(A.cam, A.width, A.height, x, y, j1, j2, o, d);
                        if (A.cam.lens_radius > 0.f) apply_lens(A.cam, xi.z, xi.w, o, d);
                        thr = f3(1.f, 1.f, 1.f);
                        bounce = 0;
                        need_gen = false;
                        n_paths++;
                    }
                    else {
                        float3 add;
                        bool cont = shade_hit<SMEM_ALL>(S, h, o, d, thr, add, xi);
                        bounce++;
                        if (cont && bounce >= A.max_depth) { // exceeded recursion: black (Kernel.cu:79)
                            cont = false;
                            add = f3(0.f, 0.f, 0.f);
                        }
                        if (!cont) {
                            acc.x += add.x; acc.y += add.y; acc.z += add.z; acc.w += 1.0f;
                            s++;
                            need_gen = true; // regenerates at the next refill
                        }
                    }
                    if (!need_gen) { // start the closest-hit query of the new ray
                        n_rays++;
                        const float tiny = 1e-18f;
                        float3 dc = f3(fabsf(d.x) < tiny ? copysignf(tiny, d.x) : d.x, fabsf(d.y) < tiny ? copysignf(tiny, d.y) : d.y,
                                       fabsf(d.z) < tiny ? copysignf(tiny, d.z) : d.z);
                        if (EXACT) {
                            a = dot3_rn(d, d); inv_a = 0.f;
                            id = f3(div_rn(1.0f, dc.x), div_rn(1.0f, dc.y), div_rn(1.0f, dc.z));
                            ood = f3(0.f, 0.f, 0.f);
                        }
                        else {
                            a = dot3(d, d); inv_a = fast_rcp(a);
                            id = f3(fast_rcp(dc.x), fast_rcp(dc.y), fast_rcp(dc.z));
                            ood = f3(o.x * id.x, o.y * id.y, o.z * id.z);
                        }
                        h.t = 3.402823466e+38f;
                        h.leaf = -1;
                        cur = 0;
                        sp = 0;
                    }
                }
            }
            continue;
        }

        if (cur >= 0) { // ---- node step: both child boxes, near child first ----
            float4 w0, w1, w2;
            int2 ch;
            load_node<SMEM_ALL>(S, cur, w0, w1, w2, ch);
            if (COUNT) n_box += 2;
            float lx0, ly0, lz0, hx0, hy0, hz0, lx1, ly1, lz1, hx1, hy1, hz1;
            if (EXACT) {
                lx0 = (w0.x - o.x) * id.x; ly0 = (w0.y - o.y) * id.y; lz0 = (w0.z - o.z) * id.z;
                hx0 = (w0.w - o.x) * id.x; hy0 = (w1.x - o.y) * id.y; hz0 = (w1.y - o.z) * id.z;
                lx1 = (w1.z - o.x) * id.x; ly1 = (w1.w - o.y) * id.y; lz1 = (w2.x - o.z) * id.z;
                hx1 = (w2.y - o.x) * id.x; hy1 = (w2.z - o.y) * id.y; hz1 = (w2.w - o.z) * id.z;
            }
            else {
                lx0 = fmaf(w0.x, id.x, -ood.x); ly0 = fmaf(w0.y, id.y, -ood.y); lz0 = fmaf(w0.z, id.z, -ood.z);
                hx0 = fmaf(w0.w, id.x, -ood.x); hy0 = fmaf(w1.x, id.y, -ood.y); hz0 = fmaf(w1.y, id.z, -ood.z);
                lx1 = fmaf(w1.z, id.x, -ood.x); ly1 = fmaf(w1.w, id.y, -ood.y); lz1 = fmaf(w2.x, id.z, -ood.z);
                hx1 = fmaf(w2.y, id.x, -ood.x); hy1 = fmaf(w2.z, id.y, -ood.y); hz1 = fmaf(w2.w, id.z, -ood.z);
            }
            const float tn0 = fmaxf(fmaxf(fminf(lx0, hx0), fminf(ly0, hy0)), fmaxf(fminf(lz0, hz0), t_min));
            const float tf0 = fminf(fminf(fmaxf(lx0, hx0), fmaxf(ly0, hy0)), fminf(fmaxf(lz0, hz0), h.t));
            const float tn1 = fmaxf(fmaxf(fminf(lx1, hx1), fminf(ly1, hy1)), fmaxf(fminf(lz1, hz1), t_min));
            const float tf1 = fminf(fminf(fmaxf(lx1, hx1), fmaxf(ly1, hy1)), fminf(fmaxf(lz1, hz1), h.t));
            const bool h0 = tn0 <= tf0;
            const bool h1 = tn1 <= tf1;
            if (h0 && h1) {
                const bool swap = tn1 < tn0;
                cur = swap ? ch.y : ch.x;
                stack[sp++] = swap ? ch.x : ch.y;
            }
            else if (h0) cur = ch.x;
            else if (h1) cur = ch.y;
            else cur = sp ? stack[--sp] : CRT_IDLE;
        }
        if (cur < 0 && cur != CRT_IDLE) { // ---- leaf step ----
            const int code = ~cur;
            if (COUNT) n_prim++;
            float t;
            if (leaf_test<EXACT, SMEM_ALL>(S, code, o, d, a, inv_a, t_min, h.t, t)) {
                h.t = t;
                h.leaf = code;
            }
            cur = sp ? stack[--sp] : CRT_IDLE;
        }
    }
    if (inside) A.accum[pixel] = acc;
    flush_counters(A.counters, n_paths, n_rays, n_box, n_prim, COUNT);
}
