// pipeline.cuh — k_render_pipe: persistent-CTA megakernel with a shared-memory path pool and material-binned queues.
//
// Replaces Kernel.cu:102-158 + color() :30-80 like the other two variants, with the schedule that the ncu evidence asks for:
//   * k_render_mega (one thread = one pixel) runs its node loop with ~12-17 of 32 lanes: a warp waits for its longest
//     traversal and for its longest path (profiles/r01_a);
//   * the HBM wavefront (wavefront.cuh) fixes that with per-lane refill, but pays a kernel ramp/drain per bounce, L2 latency
//     on every refill and ~160 B of queue traffic per ray (profiles/r01_b).
// Here ONE CTA of 1024 threads per SM keeps N path slots in SHARED memory (SoA float4) together with the scene, and seven
// ring queues of slot ids: EXT (rays waiting for a closest-hit query), one queue per shading class (miss, Lambertian, metal,
// dielectric, light) and GEN (finished slots waiting for a new (pixel, sample) work item). Every warp alternates between
//   - extend: lanes pop rays from EXT individually as they go idle (ballot/popc + one CAS per warp), traverse, and push the
//     finished ray into the queue of the material it hit — rays are BINNED BY MATERIAL on the way out;
//   - shade: when a class queue holds a full warp of rays, the warp pops 32 of them and shades them with uniform control
//     flow, accumulates finished paths (order-free integer atomics), and pushes survivors back to EXT / finished slots to GEN;
//   - generate: pops 32 free slots, claims 32 work items with one atomic on a sharded global counter, starts camera paths.
// Nothing leaves the SM except the accumulation atomics and one work-counter atomic per 32 paths.
// The per-sample arithmetic is gen_primary()/scatter_step() of kernels.cuh, so the buffer is bit-identical to the other variants.
#pragma once

#include "wavefront.cuh"

#define PIPE_THREADS 1024
#define PIPE_WARPS (PIPE_THREADS / 32)
#define PIPE_EMPTY 0xffffu // ring entries are 16-bit slot ids

enum { PQ_EXT = 0, PQ_MISS = 1, PQ_LAMBERTIAN = 2, PQ_METAL = 3, PQ_DIELECTRIC = 4, PQ_LIGHT = 5, PQ_GEN = 6, PQ_COUNT = 7 };

struct PipeArgs
{
    RenderArgs R;
    WfShard* shards;         // WF_SHARDS global work counters over the (pixel, sample) items of this launch
    DevCounters* stat;       // CRT_COUNTER_SHARDS ray/path counters
    uint32_t n_pixels_cover; // work item w -> sample w / n_pixels_cover, pixel w % n_pixels_cover
    unsigned long long inv_pixels, inv_cover_w; // floor(2^64 / d) + 1: w / d == umul64hi(w, inv) for every 32-bit w (pipe_item)
    int n_slots;             // N, a power of two
    int thresh;              // traversal leaves its loop when fewer lanes than this are busy (a refill is due)
    int max_rounds;          // traversal rounds between two refills at most
};

// Control words of the queues (static shared memory). head/tail count reservations and never wrap in practice
// (32-bit, one increment per queue operation of one CTA).
struct PipeCtl
{
    unsigned int head[PQ_COUNT + 1];
    unsigned int tail[PQ_COUNT + 1];
    int live; // slots not yet retired; the CTA exits at 0
};

struct PipePool
{
    float4* a; // {o.x, o.y, o.z, d.x}
    float4* b; // {d.y, d.z, t_hit, leaf code (int bits)}
    float4* c; // {throughput rgb, bounce (uint bits)}
    uint2* px; // {pixel, sample}
    unsigned short* ring; // PQ_COUNT rings of n_slots entries (16-bit: the rings are a third of the pool otherwise)
    unsigned int mask;
    int n;
};

__device__ __forceinline__ volatile unsigned short* pipe_entry(const PipePool& P, int q, unsigned int pos)
{
    return (volatile unsigned short*)(P.ring + q * P.n + (int)(pos & P.mask));
}

// Hand-over of a slot between warps, in the terms of the PTX memory model: the producer fills the pool slot with plain
// shared-memory stores and then PUBLISHES the slot id with a release store at CTA scope (st.release.cta.shared: MEMBAR.ALL.CTA
// + STS); the consumer polls the ring entry with an acquire load (ld.acquire.cta.shared: a plain LDS, the SM keeps later loads
// behind it) and only then reads the pool slot. Round 1 relied on a compiler-only barrier and the SM's in-order shared stores.
__device__ __forceinline__ void pipe_entry_publish(volatile unsigned short* e, int slot)
{
    const unsigned a = (unsigned)__cvta_generic_to_shared(const_cast<unsigned short*>(e));
    asm volatile("st.release.cta.shared.u16 [%0], %1;" ::"r"(a), "h"((unsigned short)slot) : "memory");
}
__device__ __forceinline__ unsigned int pipe_entry_acquire(volatile unsigned short* e)
{
    const unsigned a = (unsigned)__cvta_generic_to_shared(const_cast<unsigned short*>(e));
    unsigned short v;
    asm volatile("ld.acquire.cta.shared.u16 %0, [%1];" : "=h"(v) : "r"(a) : "memory");
    return v;
}
// A ring has as many entries as there are slots and tail counts reservations without bound, so a producer can only meet an
// entry that a consumer has claimed (head CAS) but not yet read and cleared if n further pushes went through the same queue in
// those few instructions: impossible with the 1024 / 2048 slots the planner picks, conceivable with the tiny pools CRT_PIPE_SLOTS
// allows for tests. There the producer waits for the entry to be empty first.
__device__ __forceinline__ void pipe_entry_wait_empty(const PipePool& P, volatile unsigned short* e)
{
    if (P.n < 1024)
        while (*e != PIPE_EMPTY) { }
}

// Warp-aggregated push of `slot` (where pred) into queue q. The pool writes of the pushing lanes come before in program order.
__device__ __forceinline__ void pipe_push(const PipePool& P, PipeCtl* ctl, int q, bool pred, int slot, unsigned lane_lt)
{
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m == 0u) return;
    unsigned int base = 0;
    const int leader = __ffs(m) - 1;
    if ((int)(threadIdx.x & 31u) == leader) base = atomicAdd(&ctl->tail[q], (unsigned int)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (pred) {
        volatile unsigned short* e = pipe_entry(P, q, base + __popc(m & lane_lt));
        pipe_entry_wait_empty(P, e);
        pipe_entry_publish(e, slot);
    }
}

// Push where every pushing lane names its own queue: lanes are grouped with match.any, one atomic per distinct queue.
__device__ __forceinline__ void pipe_push_binned(const PipePool& P, PipeCtl* ctl, int q, bool pred, int slot, unsigned lane_lt)
{
    const unsigned grp = __match_any_sync(0xffffffffu, pred ? q : -1); // lanes with the same destination
    if (!__any_sync(0xffffffffu, pred)) return;
    const int leader = __ffs(grp) - 1;
    unsigned int base = 0;
    if (pred && (int)(threadIdx.x & 31u) == leader) base = atomicAdd(&ctl->tail[q], (unsigned int)__popc(grp));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (pred) {
        volatile unsigned short* e = pipe_entry(P, q, base + __popc(grp & lane_lt));
        pipe_entry_wait_empty(P, e);
        pipe_entry_publish(e, slot);
    }
}

// Warp-level pop of up to `want` entries of queue q (never more than are reserved by producers). Returns the count n and
// gives lanes with rank < n their slot id. One CAS per call; a lost race returns 0 and the caller simply tries again later.
__device__ __forceinline__ int pipe_pop(const PipePool& P, PipeCtl* ctl, int q, int want, int rank, bool take, int& slot)
{
    unsigned int h = 0;
    int n = 0, avail = 0;
    if ((threadIdx.x & 31u) == 0u) {
        h = *(volatile unsigned int*)&ctl->head[q];
        const unsigned int t = *(volatile unsigned int*)&ctl->tail[q];
        avail = (int)(t - h);
        n = min(want, avail);
        if (n > 0 && atomicCAS(&ctl->head[q], h, h + (unsigned int)n) != h) n = 0;
    }
    h = __shfl_sync(0xffffffffu, h, 0);
    n = __shfl_sync(0xffffffffu, n, 0);
    if (take && rank < n) {
        volatile unsigned short* e = pipe_entry(P, q, h + (unsigned int)rank);
        unsigned int v;
        while ((v = pipe_entry_acquire(e)) == PIPE_EMPTY) { } // the producer reserved this entry before us; its store is in flight
        *e = (unsigned short)PIPE_EMPTY;
        slot = (int)v;
    }
    return n;
}

__device__ __forceinline__ void pipe_item(const PipeArgs& A, uint32_t w, int& x, int& y, uint32_t& pixel, uint32_t& sample)
{
    // two divisions by launch constants per generated path were 1.2 % of all instructions (profiles/r02_f): multiply instead
    // (inv = floor((2^64 - 1) / d) + 1 is exact for 2 <= d < 2^32 and every 32-bit dividend; d == 1 has no such multiplier)
    const uint32_t s = A.n_pixels_cover == 1u ? w : (uint32_t)__umul64hi((unsigned long long)w, A.inv_pixels), q = w - s * A.n_pixels_cover;
    const int lr = A.R.cover_w == 1u ? (int)q : (int)__umul64hi((unsigned long long)q, A.inv_cover_w);
    x = (int)(q - (uint32_t)lr * A.R.cover_w);
    y = crt_image_row(A.R.tile_rows, A.R.tile_rank, A.R.tile_world, lr);
    pixel = (uint32_t)y * A.R.width + (uint32_t)x;
    sample = A.R.spp_begin + s * A.R.spp_stride;
}

// Claims up to `need` consecutive work items from the sharded global counters (one atomic). Warp-uniform result.
// `exhausted` becomes true (and stays true) once a scan over every shard found nothing.
__device__ __forceinline__ int pipe_claim(const PipeArgs& A, int need, unsigned int& shard_cursor, bool& exhausted, uint32_t& base)
{
    int g = 0;
    uint32_t wb = 0;
    unsigned int sc = shard_cursor;
    if ((threadIdx.x & 31u) == 0u && !exhausted) {
        for (int k = 0; k < WF_SHARDS; k++) {
            WfShard* sh = &A.shards[(sc + (unsigned)k) & (WF_SHARDS - 1)];
            const unsigned int end = sh->end;
            if (*(volatile unsigned int*)&sh->head < end) {
                const unsigned int b = atomicAdd(&sh->head, (unsigned int)need);
                if (b < end) {
                    g = (int)min((unsigned int)need, end - b);
                    wb = b;
                    sc = (sc + (unsigned)k) & (WF_SHARDS - 1);
                    break;
                }
            }
        }
    }
    g = __shfl_sync(0xffffffffu, g, 0);
    base = __shfl_sync(0xffffffffu, wb, 0);
    shard_cursor = __shfl_sync(0xffffffffu, sc, 0);
    if (g == 0) exhausted = true;
    return g;
}

template <bool SMEM_ALL> __device__ __forceinline__ uint32_t pipe_mat_bits(const SceneView& S, int i)
{
    if (SMEM_ALL || i < S.smem_mats) return __float_as_uint(S.s_mats[2 * i + 1].w);
    return __float_as_uint(__ldg(S.g_mats + 2 * (size_t)i + 1).w);
}

// One batch of queue q (n <= 32 slots, lane < n holds `slot`): generate, or shade one bounce. Returns whether the lane's slot now
// holds a ray that waits for its closest-hit query (the caller extends it in place or pushes it into EXT).
template <bool SMEM_ALL>
__device__ __forceinline__ bool pipe_batch(const PipeArgs& A, const SceneView& S, const PipePool& P, PipeCtl* ctl, int q, int n, int slot,
                                           unsigned lane, unsigned lt, unsigned int& shard_cursor, bool& exhausted)
{
    const bool mine = (int)lane < n;
    if (q == PQ_GEN) {
        int done = 0;
        bool started = false;
        while (done < n) {
            uint32_t base;
            const int g = pipe_claim(A, n - done, shard_cursor, exhausted, base);
            if (g == 0) break;
            if (mine && (int)lane >= done && (int)lane < done + g) {
                int x, y;
                uint32_t pixel, sample;
                pipe_item(A, base + (uint32_t)((int)lane - done), x, y, pixel, sample);
                float3 o, d;
                gen_primary<false>(A.R, x, y, pixel, sample, o, d);
                P.a[slot] = make_float4(o.x, o.y, o.z, d.x);
                P.b[slot] = make_float4(d.y, d.z, 0.f, 0.f);
                P.c[slot] = make_float4(1.f, 1.f, 1.f, __uint_as_float(0u));
                P.px[slot] = make_uint2(pixel, sample);
                started = true;
            }
            done += g;
        }
        if (lane == 0 && done < n) atomicSub(&ctl->live, n - done); // no work left: these slots retire
        return started;
    }
    bool alive = false, ended = false;
    if (mine) {
        const float4 a0 = P.a[slot], b0 = P.b[slot], c0 = P.c[slot];
        const uint2 ps = P.px[slot];
        float3 o = f3(a0.x, a0.y, a0.z), d = f3(a0.w, b0.x, b0.y), thr = f3(c0.x, c0.y, c0.z), add;
        uint32_t bounce = __float_as_uint(c0.w);
        Hit h;
        h.t = b0.z;
        h.leaf = __float_as_int(b0.w);
        alive = scatter_step<SMEM_ALL>(A.R, S, h, ps.x, ps.y, bounce, o, d, thr, add);
        if (alive) {
            P.a[slot] = make_float4(o.x, o.y, o.z, d.x);
            P.b[slot] = make_float4(d.y, d.z, 0.f, 0.f);
            P.c[slot] = make_float4(thr.x, thr.y, thr.z, __uint_as_float(bounce));
        }
        else { // accumulate: order-free integer atomics, fire-and-forget
            unsigned long long* px = reinterpret_cast<unsigned long long*>(A.R.accum + 4 * (size_t)ps.x);
            atomicAdd(px + 0, (unsigned long long)crt_to_fixed(add.x));
            atomicAdd(px + 1, (unsigned long long)crt_to_fixed(add.y));
            atomicAdd(px + 2, (unsigned long long)crt_to_fixed(add.z));
            ended = true;
        }
    }
    // (publishing before the atomics, so that the release fence does not wait for them, measured 0.7 % slower)
    pipe_push(P, ctl, PQ_GEN, ended, slot, lt);
    return alive;
}

// The child-interleaved copy of the packed node table that the global path of k_render_pipe reads (trace.cuh crt_pair_node).
__global__ void k_pair_nodes(const float4* __restrict__ nodes, float4* __restrict__ out, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 q0, q1, q2, q3;
    crt_pair_node(nodes[4 * (size_t)i], nodes[4 * (size_t)i + 1], nodes[4 * (size_t)i + 2], nodes[4 * (size_t)i + 3], q0, q1, q2, q3);
    out[4 * (size_t)i] = q0; out[4 * (size_t)i + 1] = q1; out[4 * (size_t)i + 2] = q2; out[4 * (size_t)i + 3] = q3;
}

// NSLOTS > 0: the pool size is a compile-time constant. The pool sits at the START of dynamic shared memory, so every pool
// and ring address is `constant + index`: with a run-time N behind the scene tables the compiler re-derived those addresses from
// kernel parameters at each use (3.7 % of the executed instructions, profiles/r01_j). NSLOTS == 0 takes N from the arguments.
// A traversal round is node, node, node, leaf: a lane that reaches a leaf sits the remaining node visits of the round out, and
// ONE leaf visit serves them all (measured against node, leaf, node, leaf: +7 % on C2, +4 % on the 1 M-sphere scene; two or
// four node visits per leaf visit are slower, profiles/r02_a_*).
// Both paths: branch-free node visit on child-interleaved nodes (FFMA2), dummy node for idle lanes, sentinel stack in PTX.
// Scene larger than shared memory: 44-byte nodes (64-byte records) from the child-interleaved global table, a prefix of it staged in 48-byte slots (node_visit_pairg).
// SMEM_ALL (the whole scene is staged): the nodes take the PAIR layout of trace.cuh (48-byte nodes: binary16 half-extents,
// 16-bit references) and node_visit_pair (FFMA2, branch-free, dummy node for idle lanes, sentinel stack pushed / popped in PTX).
// SPH: the scene holds only spheres (DevScene::spheres_only, chosen by the host): leaf visits without the type dispatch.
template <int NSLOTS, bool SMEM_ALL, bool SPH = false>
__global__ void __launch_bounds__(PIPE_THREADS, 1) k_render_pipe(const __grid_constant__ PipeArgs A)
{
    extern __shared__ float4 smem[];
    __shared__ PipeCtl ctl;
    PipePool P;
    P.n = NSLOTS > 0 ? NSLOTS : A.n_slots;
    P.mask = (unsigned int)P.n - 1u;
    P.a = smem;
    P.b = smem + P.n;
    P.c = smem + 2 * P.n;
    P.px = reinterpret_cast<uint2*>(smem + 3 * P.n);
    P.ring = reinterpret_cast<unsigned short*>(smem + 3 * P.n + P.n / 2);
    // 70 bytes per slot: N is a multiple of 64, so the scene tables that follow stay 16-byte aligned
    constexpr bool PAIR = SMEM_ALL;
    const SceneView S = stage_scene<true, PAIR, !PAIR>(A.R.scene, smem + (3 * P.n + P.n / 2) + (PQ_COUNT * P.n) / 8);
    for (int i = threadIdx.x; i < PQ_COUNT * P.n; i += PIPE_THREADS) P.ring[i] = (unsigned short)(i >= PQ_GEN * P.n ? i - PQ_GEN * P.n : PIPE_EMPTY);
    if (threadIdx.x <= PQ_COUNT) { ctl.head[threadIdx.x] = 0u; ctl.tail[threadIdx.x] = threadIdx.x == PQ_GEN ? (unsigned int)P.n : 0u; }
    if (threadIdx.x == 0) ctl.live = P.n; // every slot starts in the GEN queue
    __syncthreads();

    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, lt = (1u << lane) - 1u;
    const float t_min = 0.001f;
    int slot = -1, cur = CRT_TRAV_DONE;
    TravRay r = trav_setup<false>(f3(0.f, 0.f, 0.f), f3(0.f, 0.f, 1.f));
    Hit h;
    h.t = 3.402823466e+38f; // node_visit_pair: t_hit >= t_min in every lane, also in those that have no ray yet
    h.leaf = -1;
    int stack[CRT_STACK_SIZE];
    StackPtr top = stk_begin(stack); // sentinel below the first entry: a pop of the empty stack yields "finished"
    const StackPtr top0 = top;
    unsigned int shard_cursor = (blockIdx.x * PIPE_WARPS + warp) & (WF_SHARDS - 1);
    bool exhausted = false;

    bool refill_short = true;
    while (*(volatile int*)&ctl.live > 0) {
        const unsigned act0 = __ballot_sync(0xffffffffu, cur != CRT_TRAV_DONE);
        bool worked = false;
        // ---- shade / generate duty: lanes 1..6 each look at one queue, every iteration (looking only every 2nd..16th one, or
        //      only when the EXT queue ran low, made no difference: profiles/r02_c_sweep.txt) ----
        {
            int sz = 0;
            if (lane < PQ_COUNT) sz = (int)(*(volatile unsigned int*)&ctl.tail[lane] - *(volatile unsigned int*)&ctl.head[lane]);
            const unsigned full_q = __ballot_sync(0xffffffffu, lane >= 1u && sz >= 32);
            const unsigned some_q = __ballot_sync(0xffffffffu, lane >= 1u && sz > 0);
            const int ext_sz = __shfl_sync(0xffffffffu, sz, 0);
            int q = -1;
            if (full_q) {
                // rotate the starting queue with the warp id so that the warps do not all go for the same one
                const unsigned rot = 1u + (warp % (PQ_COUNT - 1));
                const unsigned hi = full_q >> rot << rot;
                q = __ffs(hi ? hi : full_q) - 1;
            }
            else if (some_q && act0 == 0u && ext_sz <= 0) {
                q = __ffs(some_q) - 1; // starving: nothing to traverse, so a partial batch is better than waiting
            }
            if (q >= 0) {
                int s2 = -1;
                const int n = pipe_pop(P, &ctl, q, 32, (int)lane, true, s2);
                if (n > 0) {
                    const bool cont = pipe_batch<SMEM_ALL>(A, S, P, &ctl, q, n, s2, lane, lt, shard_cursor, exhausted);
                    worked = true;
                    // A lane that is idle keeps the ray it has just shaded or generated and extends it itself: no trip through
                    // EXT (a push, a pop and the hand-over of the ring entry per ray); the other new rays go into the queue.
                    const bool adopt = cont && cur == CRT_TRAV_DONE;
                    pipe_push(P, &ctl, PQ_EXT, cont && !adopt, s2, lt);
                    if (adopt) slot = s2;
                    // Lanes in mid-traversal kept only {slot, cur, sp, t, leaf} across the batch: rebuild the per-ray constants.
                    const int si = max(slot, 0);
                    const float4 a0 = P.a[si], b0 = P.b[si];
                    r = trav_setup<false>(f3(a0.x, a0.y, a0.z), f3(a0.w, b0.x, b0.y));
                    if (adopt) {
                        h.t = 3.402823466e+38f;
                        h.leaf = -1;
                        cur = trav_begin<false, SMEM_ALL, PAIR, SPH>(S, r, t_min, h);
                        top = top0;
                    }
                }
            }
        }
        // ---- extend: refill idle lanes from EXT ----
        bool idle = cur == CRT_TRAV_DONE;
        unsigned m = __ballot_sync(0xffffffffu, idle);
        refill_short = false;
        if (m != 0u) {
            int s2 = -1;
            const int rank = __popc(m & lt);
            const int n = pipe_pop(P, &ctl, PQ_EXT, __popc(m), rank, idle, s2);
            if (idle && rank < n) {
                slot = s2;
                const float4 a0 = P.a[slot], b0 = P.b[slot];
                r = trav_setup<false>(f3(a0.x, a0.y, a0.z), f3(a0.w, b0.x, b0.y));
                h.t = 3.402823466e+38f;
                h.leaf = -1;
                cur = trav_begin<false, SMEM_ALL, PAIR, SPH>(S, r, t_min, h);
                top = top0;
            }
            refill_short = n < __popc(m);
        }
        if (__ballot_sync(0xffffffffu, cur != CRT_TRAV_DONE) == 0u) {
            if (!worked) __nanosleep(100);
            continue;
        }
        // ---- traverse: until fewer than THRESH lanes are busy (a refill is due) ----
        const int leave_below = refill_short ? 1 : A.thresh; // EXT ran dry: leaving early would only spin
        for (int rounds_left = A.max_rounds; rounds_left > 0; rounds_left--) {
            if (PAIR) {
                node_visit_pair(S, r, t_min, h.t, cur, top);
                node_visit_pair(S, r, t_min, h.t, cur, top);
                node_visit_pair(S, r, t_min, h.t, cur, top);
                if (pair_at_leaf(cur)) {
                    if (SPH) leaf_step_sphere(S, r, t_min, h, cur, top);
                    else leaf_step_pair<SMEM_ALL>(S, r, t_min, h, cur, top);
                }
            }
            else {
                node_visit_pairg(S, r, t_min, h.t, cur, top);
                node_visit_pairg(S, r, t_min, h.t, cur, top);
                node_visit_pairg(S, r, t_min, h.t, cur, top);
                if (cur < 0 && cur != CRT_TRAV_DONE) leaf_step_pairg<SMEM_ALL>(S, r, t_min, h, cur, top);
            }
            if (__popc(__ballot_sync(0xffffffffu, cur != CRT_TRAV_DONE)) < leave_below) break;
        }
            // ---- finished lanes: store the hit, push the slot into the queue of what it hit ----
        const bool fin = cur == CRT_TRAV_DONE && slot >= 0;
        {
            // the leaf code carries the material type of what was hit (crt_internal.h): no load needed to bin the ray
            // (PAIR: the traversal worked on 16-bit references; the pool gets the leaf code the shading expects)
            const int leaf = PAIR && h.leaf >= 0 ? pair_leaf_code(h.leaf) : h.leaf;
            const int cls = leaf < 0 ? PQ_MISS : PQ_LAMBERTIAN + ((leaf >> CRT_LEAF_MAT_SHIFT) & 3);
            if (fin) reinterpret_cast<float2*>(P.b + slot)[1] = make_float2(h.t, __int_as_float(leaf));
                    pipe_push_binned(P, &ctl, cls, fin, slot, lt);
            if (fin) slot = -1;
        }
    }
    // Ray / path counters of the CTA, read off the queues instead of being counted: every ray is popped once from one of the
    // five class queues when it is shaded, and every path ends exactly once, by a push into GEN (which started with N entries).
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long rays = 0;
        for (int q = PQ_MISS; q <= PQ_LIGHT; q++) rays += ctl.head[q];
        DevCounters* c = A.stat + (blockIdx.x & (CRT_COUNTER_SHARDS - 1));
        atomicAdd(&c->paths, (unsigned long long)(ctl.tail[PQ_GEN] - (unsigned int)P.n));
        atomicAdd(&c->rays, rays);
    }
}
