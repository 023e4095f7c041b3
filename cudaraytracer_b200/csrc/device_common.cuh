// device_common.cuh — device-side scene view, Philox4x32-10, small vector helpers.
#pragma once

#include <cuda_runtime.h>
#ifndef CRT_HOST_SIM
#include <cuda_fp16.h>
#endif
#include <stdint.h>

#define CRT_STACK_SIZE 64
#define CRT_MAX_TOP 4
#define CRT_BLOCK_THREADS 256


struct DevImage
{
    const uchar4* texels; // RGB8 of the reference (Texture.cuh:76) padded to RGBA8 so one 32-bit load fetches a texel
    int width, height;
};

struct DevScene
{
    const float4* nodes;   // 4 x float4 per node
    const float4* nodes_pair; // the same nodes child-interleaved (trace.cuh crt_pair_node), or NULL: only k_render_pipe's global path reads it
    const float4* prims;   // 1 x float4 per primitive
    const float* prim_k;   // rect plane offsets
    const float4* mats;    // 2 x float4 per primitive
    const int* prim_id;    // leaf slot -> index in the uploaded hittable list
    const DevImage* images;
    int n_nodes, n_prims;
    int smem_nodes, smem_prims, smem_mats; // how many of each the kernels stage in shared memory
    // Primitives peeled off the top of the tree (trace.cuh trav_begin): a chain root = {huge leaf, rest} (the ground sphere of
    // every benchmark scene) costs every ray an inner-node visit whose outcome is known. Such leaves are tested when a ray
    // starts, and the traversal starts at `entry`, the first node below the chain.
    int entry, n_top;
    int top_code[CRT_MAX_TOP];
    int spheres_only; // no active rectangle: leaf visits skip the primitive-type dispatch
};

struct DevCamera
{
    // InputStruct (SharedStructs.h:3-24) + the derived rightV of Kernel.cu:133 computed on the host in IEEE binary32
    float ox, oy, oz;
    float fx, fy, fz; // forwardV = orientation
    float ux, uy, uz; // upV
    float rx, ry, rz; // rightV = Normalize(Cross(upV, forwardV))
    float far_plane, near_plane, fov, far_k; // far_k = (1.0f / fov * 10.0f), Kernel.cu:145
    float bg0[3], bg1[3];
    float lens_radius, focus_dist; // thin-lens extension; lens_radius == 0 is the reference camera
};

// Counters are SHARDED over CRT_COUNTER_SHARDS addresses 128 bytes apart (same-address atomics serialise in the L2, ~8 ns
// each on B200); crt_get_counters sums the shards.
#define CRT_COUNTER_SHARDS 64
struct alignas(128) DevCounters
{
    unsigned long long paths, rays, box_tests, prim_tests;
};

// Progressive accumulation buffer: four signed 64-bit integers per pixel {R, G, B, samples}. R/G/B hold the radiance SUM in
// Q31.32 fixed point (value * 2^32). Integer addition is associative, so the buffer is bit-identical for any schedule of the
// same sample set: any kernel variant, any order of atomic adds, any number of GPUs summed with NCCL (SURVEY.md §7 hard parts).
// Every binary32 sample >= 2^-32 in magnitude is represented exactly, i.e. the sum is the exact sum of the FP32 samples.
#define CRT_FIXED_ONE 4294967296.0f
#define CRT_FIXED_INV 2.3283064365386963e-10f
#define CRT_FIXED_MAX 1073741824.0f /* per-sample clamp (2^30): 2^33 samples of that size before overflow */

__device__ __forceinline__ long long crt_to_fixed(float x)
{
    if (!(x == x)) x = 0.0f; // a NaN sample (degenerate direction) contributes nothing
    x = fminf(fmaxf(x, -CRT_FIXED_MAX), CRT_FIXED_MAX);
    return __float2ll_rn(x * CRT_FIXED_ONE); // exact: power-of-two scaling, then an exact integer conversion
}
__host__ __device__ __forceinline__ float crt_from_fixed(long long v)
{
    return (float)v * CRT_FIXED_INV; // one rounding (int64 -> binary32), the scaling is exact
}

struct RenderArgs
{
    DevScene scene;
    DevCamera cam;
    long long* accum;    // width*height x {R, G, B, samples}, see above
    DevCounters* counters;
    uint32_t width, height;       // image size
    uint32_t cover_w, cover_h;    // pixels actually rendered (== width/height unless CRT_FLAG_TRUNCATED_GRID)
    uint32_t spp_begin, spp_count, spp_stride, max_depth;
    uint32_t seed_lo, seed_hi;
    // the ten Philox round keys of this seed, {k0 + r W0, k1 + r W1}: filled on the host (crt_philox_round_keys) so that the
    // device reads each of them as a constant-bank operand instead of adding a 32-bit immediate per round and word
    uint32_t philox_keys[20];
    int jitter;
    // row-tile sharding (crt_render_params::row_tiles): this launch renders only the rows of slabs tile_rank, tile_rank +
    // tile_world, ... (slabs of tile_rows rows); cover_h then counts the LOCAL rows. tile_world <= 1: the whole frame.
    uint32_t tile_rows, tile_rank, tile_world;
};

// local row of a launch -> image row (identity unless the launch is row-tile sharded)
__device__ __forceinline__ int crt_image_row(uint32_t tile_rows, uint32_t tile_rank, uint32_t tile_world, int lr)
{
    if (tile_world <= 1u) return lr;
    const uint32_t t = (uint32_t)lr / tile_rows;
    return (int)((t * tile_world + tile_rank) * tile_rows + ((uint32_t)lr - t * tile_rows));
}

// ------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11). Stateless: counter = (pixel, sample, bounce, stream), key = seed.
// ------------------------------------------------------------------------------------------
#define CRT_PHILOX_M0 0xD2511F53u
#define CRT_PHILOX_M1 0xCD9E8D57u
#define CRT_PHILOX_W0 0x9E3779B9u
#define CRT_PHILOX_W1 0xBB67AE85u

__host__ __device__ __forceinline__ void crt_mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo)
{
#ifdef __CUDA_ARCH__
    lo = a * b;
    hi = __umulhi(a, b);
#else
    uint64_t p = (uint64_t)a * b;
    lo = (uint32_t)p;
    hi = (uint32_t)(p >> 32);
#endif
}

__host__ __device__ __forceinline__ uint4 crt_philox4x32_10(uint4 c, uint32_t k0, uint32_t k1)
{
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t hi0, lo0, hi1, lo1;
        crt_mulhilo(CRT_PHILOX_M0, c.x, hi0, lo0);
        crt_mulhilo(CRT_PHILOX_M1, c.z, hi1, lo1);
        c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
        k0 += CRT_PHILOX_W0;
        k1 += CRT_PHILOX_W1;
    }
    return c;
}

__host__ __device__ __forceinline__ void crt_philox_round_keys(uint32_t k0, uint32_t k1, uint32_t* keys)
{
    for (int r = 0; r < 10; r++) { keys[2 * r] = k0 + (uint32_t)r * CRT_PHILOX_W0; keys[2 * r + 1] = k1 + (uint32_t)r * CRT_PHILOX_W1; }
}
// the same function with the round keys precomputed
__host__ __device__ __forceinline__ uint4 crt_philox4x32_10k(uint4 c, const uint32_t* keys)
{
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t hi0, lo0, hi1, lo1;
        crt_mulhilo(CRT_PHILOX_M0, c.x, hi0, lo0);
        crt_mulhilo(CRT_PHILOX_M1, c.z, hi1, lo1);
        c = make_uint4(hi1 ^ c.y ^ keys[2 * r], lo1, hi0 ^ c.w ^ keys[2 * r + 1], lo0);
    }
    return c;
}

#define CRT_STREAM_CAMERA 0u
#define CRT_STREAM_SCATTER 1u

// Same mapping as curand_uniform (curand_uniform.h:69-72): (0, 1].
__host__ __device__ __forceinline__ float crt_uniform(uint32_t x)
{
    return (float)x * 2.3283064e-10f + (2.3283064e-10f / 2.0f);
}

// ------------------------------------------------------------------------------------------
// float3 helpers. Device code is compiled with -fmad=false: the compiler never contracts a*b+c on its own, every fused
// multiply-add of the fast path is an explicit fmaf(). That makes the arithmetic of a function independent of the kernel it is
// inlined into, which is what lets the three render variants produce bit-identical samples.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float3 operator*(float s, float3 a) { return f3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float dot3(float3 a, float3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
// a + s * b
__device__ __forceinline__ float3 madd3(float s, float3 b, float3 a) { return f3(fmaf(s, b.x, a.x), fmaf(s, b.y, a.y), fmaf(s, b.z, a.z)); }

// Approximate single-instruction forms for the throughput path (MUFU.RCP / MUFU.SQRT; ~1-2 ulp).
#ifdef CRT_HOST_SIM // tests/sim/trace_sim.cpp compiles the traversal for the HOST (g++) to check its logic without a GPU
__device__ __forceinline__ float fast_rcp(float x) { return 1.0f / x; }
__device__ __forceinline__ float fast_sqrt(float x) { return sqrtf(x); }
#else
__device__ __forceinline__ float fast_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float fast_sqrt(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
#endif
__device__ __forceinline__ float fast_div(float a, float b) { return __fdividef(a, b); }
// float -> int as the GPU converts it (cvt.rzi.s32.f32: NaN -> 0, saturating); the host build of the same code (tests/sim) must
// not run into x86's undefined conversion of a NaN
__device__ __forceinline__ int crt_f2i(float x)
{
#ifdef CRT_HOST_SIM
    if (!(x == x)) return 0;
    if (x >= 2147483648.0f) return 2147483647;
    if (x <= -2147483648.0f) return -2147483647 - 1;
    return (int)x;
#else
    return __float2int_rz(x);
#endif
}

// IEEE, never-contracted versions used by the exact path (bit-for-bit the host oracle's arithmetic).
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub_rn(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float div_rn(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float sqrt_rn(float a) { return __fsqrt_rn(a); }
// Dot (Math.cuh:152-155): ((x*x' + y*y') + z*z'), left to right, unfused.
__device__ __forceinline__ float dot3_rn(float3 a, float3 b)
{
    return add_rn(add_rn(mul_rn(a.x, b.x), mul_rn(a.y, b.y)), mul_rn(a.z, b.z));
}
