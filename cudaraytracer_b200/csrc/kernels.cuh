// kernels.cuh — the __global__ entry points.
//   k_render_mega   persistent-path megakernel: one thread owns a pixel and walks all of its samples as ONE
//                   flattened bounce loop (a lane whose path ends starts its next sample at once instead of
//                   waiting for the longest path of the warp, which is what Kernel.cu:137-148 + color() do).
//   k_primary_ids   exact-arithmetic primary-ray closest hit -> hittable index (parity hook).
//   k_resolve       accumulation buffer -> packed RGBA8, bit-compatible with Kernel.cu:12-19,151-157.
//   k_clear         zero the accumulation buffer.      k_accum_to_float  fixed-point buffer -> float4 {sum rgb, samples}.
// Wavefront kernels live in wavefront.cuh. Both variants share gen_primary() / scatter_step() below, so that every
// (pixel, sample) yields the same FP32 radiance whichever kernel computes it, and the integer accumulation makes the
// buffers bit-identical.
#pragma once

#include "path.cuh"

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
    // warp-aggregated: one atomic per warp per counter, spread over the counter shards
    for (int off = 16; off > 0; off >>= 1) {
        paths += __shfl_xor_sync(0xffffffffu, paths, off);
        rays += __shfl_xor_sync(0xffffffffu, rays, off);
        if (count) {
            nbox += __shfl_xor_sync(0xffffffffu, nbox, off);
            nprim += __shfl_xor_sync(0xffffffffu, nprim, off);
        }
    }
    if ((threadIdx.x & 31) == 0) {
        c += (blockIdx.x + (threadIdx.x >> 5)) & (CRT_COUNTER_SHARDS - 1);
        atomicAdd(&c->paths, (unsigned long long)paths);
        atomicAdd(&c->rays, (unsigned long long)rays);
        if (count) {
            atomicAdd(&c->box_tests, (unsigned long long)nbox);
            atomicAdd(&c->prim_tests, (unsigned long long)nprim);
        }
    }
}

// PAIRM (whole scene staged, fast arithmetic, no counting): 0 = the word-major node table and closest_hit; 1 / 2 = the PAIR
// staging of trace.cuh and closest_hit_pair, 2 when the scene holds only spheres.
template <bool EXACT, bool COUNT, bool SMEM_ALL, int PAIRM = 0>
__global__ void __launch_bounds__(CRT_BLOCK_THREADS) k_render_mega(const __grid_constant__ RenderArgs A)
{
    extern __shared__ float4 smem[];
    const SceneView S = PAIRM ? stage_scene<true, true>(A.scene, smem) : stage_scene<false>(A.scene, smem);
    __syncthreads();

    int x, y;
    tile_pixel((A.cover_w + 15) / 16, x, y);
    const bool inside = (uint32_t)x < A.cover_w && (uint32_t)y < A.cover_h;
    y = crt_image_row(A.tile_rows, A.tile_rank, A.tile_world, y);
    const uint32_t pixel = (uint32_t)y * A.width + (uint32_t)x;

    long long acc_r = 0, acc_g = 0, acc_b = 0; // this launch's samples; added to the buffer once at the end

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
            gen_primary<EXACT>(A, x, y, pixel, sample, o, d);
            thr = f3(1.f, 1.f, 1.f);
            bounce = 0;
            need_gen = false;
            n_paths++;
        }
        n_rays++;
        Hit h = PAIRM ? closest_hit_pair<PAIRM == 2>(S, o, d, 0.001f, 3.402823466e+38f)
                      : closest_hit<EXACT, COUNT, SMEM_ALL>(S, o, d, 0.001f, 3.402823466e+38f, n_box, n_prim);
        float3 add;
        if (!scatter_step<SMEM_ALL>(A, S, h, pixel, sample, bounce, o, d, thr, add)) {
            acc_r += crt_to_fixed(add.x); acc_g += crt_to_fixed(add.y); acc_b += crt_to_fixed(add.z);
            s++;
            need_gen = true;
        }
    }
    if (inside) { // the pixel is owned by this thread for the whole launch: plain read-modify-write
        longlong2* px = reinterpret_cast<longlong2*>(A.accum + 4 * (size_t)pixel);
        longlong2 rg = px[0], bn = px[1];
        rg.x += acc_r; rg.y += acc_g; bn.x += acc_b; bn.y += (long long)A.spp_count;
        px[0] = rg; px[1] = bn;
    }
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
    const SceneView S = stage_scene<false>(A.scene, smem);
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
// The FP32 sum is the one crt_read_accum reports (fixed point -> binary32, one rounding); IEEE division and square root
// so the result is bit-identical to the host oracle for the same buffer.
__device__ __forceinline__ uint32_t resolve_pack(longlong2 rg, longlong2 bn)
{
    float4 a = make_float4(crt_from_fixed(rg.x), crt_from_fixed(rg.y), crt_from_fixed(bn.x), (float)bn.y);
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
    return packed;
}

// (tile_rows, tile_rank, tile_world): resolve only the rows of a row-tile shard (cover_h = its local row count); out may be a
// peer mapping of the root's frame.
__global__ void k_resolve(const long long* __restrict__ accum, uint32_t* __restrict__ out, uint32_t width, uint32_t cover_w, uint32_t cover_h,
                          uint32_t tile_rows, uint32_t tile_rank, uint32_t tile_world)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cover_w * cover_h) return;
    uint32_t x = i % cover_w, y = (uint32_t)crt_image_row(tile_rows, tile_rank, tile_world, (int)(i / cover_w));
    uint32_t pixel = y * width + x;
    const longlong2* px = reinterpret_cast<const longlong2*>(accum + 4 * (size_t)pixel);
    out[pixel] = resolve_pack(px[0], px[1]);
}

// Multi-GPU frame close-out in ONE kernel over peer memory (NVLink): sum of the ranks' accumulation buffers + resolve + store
// of the packed pixels into the root's frame, for the pixel range this rank owns. Replaces reduce-to-root (every byte of every
// buffer into ONE GPU) + k_resolve on the root: each GPU pulls 1/G of each peer's buffer (the ingress is spread over all
// NVLink ports of the switch) and only 4 bytes per pixel travel to the root. P.p[g] are peer mappings (cudaIpcOpenMemHandle)
// or local pointers; the sum is integer, hence exact and the same bits as the NCCL sum. The n loads of a pixel are
// independent and issued back to back (the loop is unrolled under a predicate), L2-only (ld.cg): the peers rewrite these
// addresses every frame and L1 is not coherent with them.
#define CRT_MAX_PEERS 16
struct PeerAccum
{
    const long long* p[CRT_MAX_PEERS];
    int n;
};

// NP = compile-time bound of P.n (2, 4, 8 or 16): the loads live in registers, and a 16-wide array for two peers left one
// 256-thread block per SM (172 registers) - 2.4 MB in flight on the whole chip, latency-bound at 0.6 TB/s (profiles/r01_m).
template <int NP>
__global__ void __launch_bounds__(256) k_resolve_peers(const __grid_constant__ PeerAccum P, uint32_t* __restrict__ out,
                                                       long long* __restrict__ sum_out, uint32_t first_pixel, uint32_t n_pixels)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pixels) return;
    const size_t pixel = (size_t)first_pixel + i;
    longlong2 rg[NP], bn[NP];
#pragma unroll
    for (int g = 0; g < NP; g++) {
        if (g < P.n) {
            const longlong2* px = reinterpret_cast<const longlong2*>(P.p[g] + 4 * pixel);
            rg[g] = __ldcg(px);
            bn[g] = __ldcg(px + 1);
        }
    }
    longlong2 srg = make_longlong2(0, 0), sbn = make_longlong2(0, 0);
#pragma unroll
    for (int g = 0; g < NP; g++) {
        if (g < P.n) {
            srg.x += rg[g].x; srg.y += rg[g].y;
            sbn.x += bn[g].x; sbn.y += bn[g].y;
        }
    }
    if (sum_out) {
        longlong2* so = reinterpret_cast<longlong2*>(sum_out + 4 * pixel);
        so[0] = srg;
        so[1] = sbn;
    }
    out[pixel] = resolve_pack(srg, sbn);
}

__global__ void k_accum_to_float(const long long* __restrict__ accum, float4* __restrict__ out, uint32_t n)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const longlong2* px = reinterpret_cast<const longlong2*>(accum + 4 * (size_t)i);
    longlong2 rg = px[0], bn = px[1];
    out[i] = make_float4(crt_from_fixed(rg.x), crt_from_fixed(rg.y), crt_from_fixed(bn.x), (float)bn.y);
}

// Image texels arrive as tightly packed RGB8 (stb_image's layout, RawStbImage.h:11-22); the kernels fetch one aligned RGBA8
// word per lookup. One thread expands 4 texels: three 32-bit loads, one 128-bit store.
__global__ void k_rgb_to_rgba(const uint32_t* __restrict__ rgb, uchar4* __restrict__ rgba, size_t n_px)
{
    const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t p = 4 * g;
    if (p + 4 <= n_px) {
        const uint32_t a = rgb[3 * g], b = rgb[3 * g + 1], c = rgb[3 * g + 2];
        uint4 o;
        o.x = (a & 0x00ffffffu) | 0xff000000u;
        o.y = ((a >> 24) | (b << 8)) & 0x00ffffffu | 0xff000000u;
        o.z = ((b >> 16) | (c << 16)) & 0x00ffffffu | 0xff000000u;
        o.w = (c >> 8) | 0xff000000u;
        reinterpret_cast<uint4*>(rgba)[g] = o;
    }
    else if (p < n_px) {
        const unsigned char* src = reinterpret_cast<const unsigned char*>(rgb);
        for (size_t k = p; k < n_px; k++) rgba[k] = make_uchar4(src[3 * k], src[3 * k + 1], src[3 * k + 2], 255);
    }
}

__global__ void k_clear(longlong2* __restrict__ accum, uint32_t n2)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n2) accum[i] = make_longlong2(0, 0);
}
