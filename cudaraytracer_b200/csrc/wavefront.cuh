// wavefront.cuh — wavefront pipeline: generate -> extend -> shade (+ regenerate) -> accumulate, with SoA path queues in HBM.
//
// Replaces the same code as the megakernel (Kernel.cu:102-158 + color() :30-80) with a different schedule:
//   * a fixed pool of path SLOTS (not pixels) lives in HBM as SoA float4 arrays;
//   * k_wf_extend is a persistent kernel that only traverses: a lane whose ray finished its closest-hit query stores the hit
//     and fetches the next slot of its CTA's share of the pool at once (shared-memory batch cursor), so a warp never waits
//     for its longest traversal — the loss ncu shows in the megakernel (profiles/r01_a: 12 of 32 lanes active in the node loop);
//   * k_wf_shade runs one thread per active slot: miss/scatter/emit, adds the radiance of finished paths to the integer
//     accumulation buffer with atomics (order-free, see device_common.cuh) and REGENERATES the slot from the sharded work
//     counters ((pixel, sample) items). The work grant is aggregated with __ballot_sync/__popc inside a warp and ONE atomic
//     per block. In the steady state every slot is alive, so there is no index list to compact; a slot that finds no work
//     left is marked dead and skipped by extend while the pass drains.
// Work items are decoupled from slots, so pixels with long paths (glass) do not leave their lanes idle behind pixels with
// short ones (sky); Philox is keyed by (pixel, sample, bounce), so the schedule does not change any sample.
#pragma once

#include "kernels.cuh"

#define WF_EXT_THREADS 1024 /* one CTA per SM: one shared-memory copy of the scene per SM */
#define WF_SHADE_THREADS 256
#define WF_BATCH 64    /* slots a warp takes from its CTA's shared-memory cursor at a time */
#define WF_SHARDS 128  /* work-counter shards (power of two) */
#define WF_STAT_SHARDS 32
#define WF_DEAD ((int)0x80000001) /* leaf code of a slot that holds no path */

// Same-address global atomics serialise in the L2 (measured here: ~8 ns each, profiles/r01_b), so every counter that is
// hit once per block or more is SHARDED over addresses 128 bytes apart; nothing in the steady state funnels through one word.
struct alignas(128) WfShard
{
    unsigned int head, end; // unclaimed (pixel, sample) items of this pass: [head, end)
};
struct alignas(128) WfStat
{
    unsigned long long rays, paths;
    unsigned int alive[2]; // slots holding a live path after shade of an even / odd iteration
};
struct WfCounters
{
    WfShard shard[WF_SHARDS];
    WfStat stat[WF_STAT_SHARDS];
};

struct WfArgs
{
    RenderArgs R;
    float4* r0;   // {o.x, o.y, o.z, d.x}
    float4* r1;   // {d.y, d.z, t_hit, leaf code (int bits) or WF_DEAD}; extend overwrites .zw
    float4* thr;  // {throughput rgb, bounce (uint bits)}
    uint2* px;    // {pixel, sample}
    WfCounters* c;
    uint32_t n_pixels_cover; // work item w -> sample w / n_pixels_cover, pixel w % n_pixels_cover (row-major over the covered region)
    int iter;                // iteration index of the pass (rotates the shard a block asks first; parity of the alive counter)
    int capacity;            // slots in the pool
};

__device__ __forceinline__ void wf_item(const WfArgs& A, uint32_t w, int& x, int& y, uint32_t& pixel, uint32_t& sample)
{
    const uint32_t s = w / A.n_pixels_cover, q = w - s * A.n_pixels_cover;
    const int lr = (int)(q / A.R.cover_w);
    x = (int)(q - (uint32_t)lr * A.R.cover_w);
    y = crt_image_row(A.R.tile_rows, A.R.tile_rank, A.R.tile_world, lr);
    pixel = (uint32_t)y * A.R.width + (uint32_t)x;
    sample = A.R.spp_begin + s * A.R.spp_stride;
}

__device__ __forceinline__ void wf_store_path(const WfArgs& A, int slot, float3 o, float3 d, float3 thr, uint32_t bounce)
{
    A.r0[slot] = make_float4(o.x, o.y, o.z, d.x);
    A.r1[slot] = make_float4(d.y, d.z, 0.f, 0.f);
    A.thr[slot] = make_float4(thr.x, thr.y, thr.z, __uint_as_float(bounce));
}

// Fills the first min(capacity, work_total) slots with the first work items, marks the rest dead, and splits the
// remaining items over the work shards.
__global__ void __launch_bounds__(WF_SHADE_THREADS) k_wf_generate(const __grid_constant__ WfArgs A, uint32_t work_total)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n = min((uint32_t)A.capacity, work_total);
    if (i < WF_SHARDS) {
        const uint32_t rest = work_total - n, per = rest / WF_SHARDS, extra = rest % WF_SHARDS;
        const uint32_t b = n + i * per + min(i, extra);
        A.c->shard[i].head = b;
        A.c->shard[i].end = b + per + (i < extra ? 1u : 0u);
    }
    if (i < WF_STAT_SHARDS) A.c->stat[i].alive[0] = A.c->stat[i].alive[1] = 0u;
    if (i >= (uint32_t)A.capacity) return;
    if (i >= n) {
        A.r1[i] = make_float4(0.f, 0.f, 0.f, __int_as_float(WF_DEAD));
        return;
    }
    int x, y;
    uint32_t pixel, sample;
    wf_item(A, i, x, y, pixel, sample);
    float3 o, d;
    gen_primary<false>(A.R, x, y, pixel, sample, o, d);
    wf_store_path(A, (int)i, o, d, f3(1.f, 1.f, 1.f), 0u);
    A.px[i] = make_uint2(pixel, sample);
}

// ------------------------------------------------------------------------------------------
// extend: persistent traversal with per-lane refill.
// ------------------------------------------------------------------------------------------
template <int THRESH, bool SMEM_ALL>
__global__ void __launch_bounds__(WF_EXT_THREADS, 1) k_wf_extend(const __grid_constant__ WfArgs A)
{
    extern __shared__ float4 smem[];
    __shared__ int s_cursor;
    const SceneView S = stage_scene<true>(A.R.scene, smem);
    // This CTA's contiguous share of the slot pool; its warps take WF_BATCH slots at a time from a shared-memory cursor.
    const int per = (A.capacity + (int)gridDim.x - 1) / (int)gridDim.x;
    const int lo = min((int)blockIdx.x * per, A.capacity), hi = min(lo + per, A.capacity);
    if (threadIdx.x == 0) s_cursor = lo;
    // the alive counter this iteration's shade adds to (last written two iterations ago)
    if (blockIdx.x == 0 && threadIdx.x < WF_STAT_SHARDS) A.c->stat[threadIdx.x].alive[A.iter & 1] = 0u;
    __syncthreads();

    const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    const float t_min = 0.001f;

    int cursor = 0, cursor_end = 0; // warp-uniform: this warp's batch of slots
    bool more = hi > lo;            // warp-uniform: the CTA's range may still have unclaimed slots
    int slot = -1, cur = CRT_TRAV_DONE, sp = 0;
    TravRay r = trav_setup<false>(f3(0.f, 0.f, 0.f), f3(0.f, 0.f, 1.f));
    Hit h;
    h.t = 0.f;
    h.leaf = -1;
    int stack[CRT_STACK_SIZE];

    while (true) {
        // ---- lanes that finished: store the hit, then refill ----
        bool idle = cur == CRT_TRAV_DONE;
        if (idle && slot >= 0) {
            reinterpret_cast<float2*>(A.r1 + slot)[1] = make_float2(h.t, __int_as_float(h.leaf));
            slot = -1;
        }
        unsigned m = __ballot_sync(0xffffffffu, idle);
        while (m != 0u && more) {
            if (cursor == cursor_end) {
                int b = 0;
                if (lane == 0) b = atomicAdd(&s_cursor, WF_BATCH);
                b = __shfl_sync(0xffffffffu, b, 0);
                cursor = min(b, hi);
                cursor_end = min(b + WF_BATCH, hi);
                if (cursor == cursor_end) { more = false; break; }
            }
            const int avail = cursor_end - cursor;
            const int rank = __popc(m & lt);
            if (idle && rank < avail) {
                const int cand = cursor + rank;
                const float4 a1 = A.r1[cand];
                if (__float_as_int(a1.w) != WF_DEAD) { // dead slots (only while a pass drains) are skipped
                    const float4 a0 = A.r0[cand];
                    slot = cand;
                    r = trav_setup<false>(f3(a0.x, a0.y, a0.z), f3(a0.w, a1.x, a1.y));
                    h.t = 3.402823466e+38f;
                    h.leaf = -1;
                    cur = trav_begin<false, SMEM_ALL>(S, r, t_min, h);
                    sp = 0;
                    idle = false;
                }
            }
            cursor += min(__popc(m), avail);
            m = __ballot_sync(0xffffffffu, idle);
        }
        unsigned act = __ballot_sync(0xffffffffu, cur != CRT_TRAV_DONE);
        if (act == 0u) break;
        // ---- traverse until fewer than THRESH lanes are busy (or, with nothing left to fetch, until all are done) ----
        do {
            if (cur >= 0) node_step<false, SMEM_ALL, true>(S, r, t_min, h.t, cur, sp, stack);
            if (cur < 0 && cur != CRT_TRAV_DONE) leaf_step<false, SMEM_ALL>(S, r, t_min, h, cur, sp, stack);
            act = __ballot_sync(0xffffffffu, cur != CRT_TRAV_DONE);
        } while (more ? __popc(act) >= THRESH : act != 0u);
    }
}

// ------------------------------------------------------------------------------------------
// shade: one thread per slot (no indirection: in the steady state every slot holds a live path).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(WF_SHADE_THREADS) k_wf_shade(const __grid_constant__ WfArgs A)
{
    __shared__ int s_want[WF_SHADE_THREADS / 32];
    __shared__ unsigned int s_work_base, s_granted;

    SceneView S; // everything from global memory through the read-only path (this kernel makes ~3 small loads per ray)
    S.g_nodes = A.R.scene.nodes; S.g_prims = A.R.scene.prims; S.g_prim_k = A.R.scene.prim_k; S.g_mats = A.R.scene.mats;
    S.s_nodes = S.s_prims = S.s_mats = nullptr;
    S.images = A.R.scene.images;
    S.smem_nodes = S.smem_prims = S.smem_mats = 0;

    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, lt = (1u << lane) - 1u;
    unsigned n_rays = 0, n_paths = 0, n_alive = 0;

    for (int base = blockIdx.x * WF_SHADE_THREADS; base < A.capacity; base += gridDim.x * WF_SHADE_THREADS) {
        const int slot = base + (int)threadIdx.x;
        const bool valid = slot < A.capacity;
        float3 o = f3(0.f, 0.f, 0.f), d = f3(0.f, 0.f, 1.f), thr = f3(1.f, 1.f, 1.f);
        uint32_t bounce = 0, pixel = 0, sample = 0;
        bool alive = false, want = false;
        if (valid) {
            const float4 a1 = A.r1[slot];
            Hit h;
            h.t = a1.z;
            h.leaf = __float_as_int(a1.w);
            if (h.leaf == WF_DEAD) {
                want = true; // an empty slot keeps asking for work until the pass has none left
            }
            else {
                const float4 a0 = A.r0[slot], a2 = A.thr[slot];
                const uint2 ps = A.px[slot];
                o = f3(a0.x, a0.y, a0.z); d = f3(a0.w, a1.x, a1.y); thr = f3(a2.x, a2.y, a2.z);
                bounce = __float_as_uint(a2.w); pixel = ps.x; sample = ps.y;
                n_rays++;
                float3 add;
                alive = scatter_step<false>(A.R, S, h, pixel, sample, bounce, o, d, thr, add);
                if (!alive) { // accumulate: order-free integer atomics (fire-and-forget REDs to spread addresses)
                    unsigned long long* px = reinterpret_cast<unsigned long long*>(A.R.accum + 4 * (size_t)pixel);
                    atomicAdd(px + 0, (unsigned long long)crt_to_fixed(add.x));
                    atomicAdd(px + 1, (unsigned long long)crt_to_fixed(add.y));
                    atomicAdd(px + 2, (unsigned long long)crt_to_fixed(add.z));
                    want = true;
                }
            }
        }
        // ---- work grant: ballot/popc inside the warp, ONE atomic per block on one of WF_SHARDS counters ----
        const unsigned mw = __ballot_sync(0xffffffffu, want);
        if (lane == 0) s_want[warp] = __popc(mw);
        __syncthreads();
        if (threadIdx.x == 0) {
            int tw = 0;
#pragma unroll
            for (int k = 0; k < WF_SHADE_THREADS / 32; k++) tw += s_want[k];
            unsigned int wb = 0, granted = 0;
            for (int attempt = 0; attempt < 2 && tw > 0 && granted == 0u; attempt++) {
                WfShard* sh = &A.c->shard[(blockIdx.x + (unsigned)A.iter + (unsigned)attempt) & (WF_SHARDS - 1)];
                const unsigned int end = sh->end;
                if (*(volatile unsigned int*)&sh->head < end) { // skip exhausted shards without touching them
                    wb = atomicAdd(&sh->head, (unsigned int)tw);
                    granted = wb >= end ? 0u : min((unsigned int)tw, end - wb);
                }
            }
            s_work_base = wb;
            s_granted = granted;
        }
        __syncthreads();
        int pre_w = 0;
#pragma unroll
        for (int k = 0; k < WF_SHADE_THREADS / 32; k++) pre_w += k < (int)warp ? s_want[k] : 0;
        const unsigned int granted = s_granted, work_base = s_work_base;
        __syncthreads(); // the shared scratch is rewritten by the next trip
        if (want) {
            const unsigned int rank = (unsigned int)pre_w + __popc(mw & lt);
            if (rank < granted) { // (re)generate the slot with the next (pixel, sample) item
                int x, y;
                wf_item(A, work_base + rank, x, y, pixel, sample);
                gen_primary<false>(A.R, x, y, pixel, sample, o, d);
                A.px[slot] = make_uint2(pixel, sample);
                wf_store_path(A, slot, o, d, f3(1.f, 1.f, 1.f), 0u);
                n_paths++;
                n_alive++;
            }
            else if (alive == false && valid) {
                A.r1[slot] = make_float4(0.f, 0.f, 0.f, __int_as_float(WF_DEAD));
            }
        }
        else if (alive) {
            wf_store_path(A, slot, o, d, thr, bounce);
            n_alive++;
        }
    }
    // ray / path / alive counters: block-aggregated, sharded
    __shared__ unsigned int s_tot[3];
    if (threadIdx.x < 3) s_tot[threadIdx.x] = 0u;
    __syncthreads();
    for (int off = 16; off > 0; off >>= 1) {
        n_rays += __shfl_xor_sync(0xffffffffu, n_rays, off);
        n_paths += __shfl_xor_sync(0xffffffffu, n_paths, off);
        n_alive += __shfl_xor_sync(0xffffffffu, n_alive, off);
    }
    if (lane == 0) { atomicAdd(&s_tot[0], n_rays); atomicAdd(&s_tot[1], n_paths); atomicAdd(&s_tot[2], n_alive); }
    __syncthreads();
    if (threadIdx.x == 0) {
        WfStat* st = &A.c->stat[blockIdx.x & (WF_STAT_SHARDS - 1)];
        if (s_tot[0]) atomicAdd(&st->rays, (unsigned long long)s_tot[0]);
        if (s_tot[1]) atomicAdd(&st->paths, (unsigned long long)s_tot[1]);
        if (s_tot[2]) atomicAdd(&st->alive[A.iter & 1], s_tot[2]);
    }
}

// Adds `spp` to the sample count of every covered pixel (the wavefront's atomics only touch R, G, B).
__global__ void k_wf_add_count(long long* __restrict__ accum, uint32_t width, uint32_t cover_w, uint32_t cover_h, long long spp,
                               uint32_t tile_rows, uint32_t tile_rank, uint32_t tile_world)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cover_w * cover_h) return;
    uint32_t x = i % cover_w, y = (uint32_t)crt_image_row(tile_rows, tile_rank, tile_world, (int)(i / cover_w));
    accum[4 * ((size_t)y * width + x) + 3] += spp;
}

// Folds the pass counters into the context counters and publishes {live slots after iteration parity `p`, unclaimed work}.
__global__ void k_wf_fold_counters(WfCounters* c, DevCounters* out, unsigned long long first_paths, int p, unsigned int* alive_out)
{
    unsigned long long rays = 0, paths = 0;
    unsigned int alive = 0;
    for (int i = 0; i < WF_STAT_SHARDS; i++) {
        rays += c->stat[i].rays; paths += c->stat[i].paths; alive += c->stat[i].alive[p];
        c->stat[i].rays = 0; c->stat[i].paths = 0;
    }
    out->rays += rays;
    out->paths += paths + first_paths;
    unsigned int remaining = 0; // work items no slot has claimed yet
    for (int i = 0; i < WF_SHARDS; i++) {
        const unsigned int h = c->shard[i].head, e = c->shard[i].end;
        remaining += h < e ? e - h : 0u;
    }
    alive_out[0] = alive;
    alive_out[1] = remaining;
}

struct WavefrontState
{
    void* buffers = nullptr;
    size_t capacity = 0;
    float4 *r0 = nullptr, *r1 = nullptr, *thr = nullptr;
    uint2* px = nullptr;
    WfCounters* c = nullptr;
    unsigned int* d_alive = nullptr; // live slots after the last iteration (written by k_wf_fold_counters)
    unsigned int* h_alive = nullptr; // pinned copy, read once per batch of iterations
};

static inline void wavefront_free(WavefrontState& w)
{
    if (w.buffers) cudaFree(w.buffers);
    if (w.h_alive) cudaFreeHost(w.h_alive);
    w = WavefrontState();
}

static inline cudaError_t wavefront_reserve(WavefrontState& w, size_t capacity)
{
    if (w.buffers && w.capacity == capacity) return cudaSuccess;
    wavefront_free(w);
    const size_t per_slot = 3 * sizeof(float4) + sizeof(uint2);
    const size_t bytes = capacity * per_slot + sizeof(WfCounters) + 512;
    cudaError_t e = cudaMalloc(&w.buffers, bytes);
    if (e != cudaSuccess) return e;
    char* p = (char*)w.buffers;
    w.r0 = (float4*)p; p += capacity * sizeof(float4);
    w.r1 = (float4*)p; p += capacity * sizeof(float4);
    w.thr = (float4*)p; p += capacity * sizeof(float4);
    w.px = (uint2*)p; p += capacity * sizeof(uint2);
    p = (char*)(((uintptr_t)p + 127) & ~(uintptr_t)127);
    w.c = (WfCounters*)p; p += sizeof(WfCounters);
    w.d_alive = (unsigned int*)p;
    if ((e = cudaMemset(w.c, 0, sizeof(WfCounters) + 16)) != cudaSuccess) return e;
    if ((e = cudaMallocHost(&w.h_alive, 16)) != cudaSuccess) return e;
    w.capacity = capacity;
    return cudaSuccess;
}
