// lbvh.cuh — BVH construction ON THE GPU for large scenes and for edits that change the topology.
//
// Replaces the host-side `BVHNode` constructor the reference runs at start-up and after EVERY GUI edit
// (Hittable.cuh:303-385, called from CudaLayer.cpp:247-256,493-494,967-968,1375-1377; ~7 s for 1 M spheres).
// The threaded SAH builder of scene_pack.cpp needs 190 ms for that scene on 16 host cores — and every rank of a multi-GPU job
// repeats it on the same cores. Here the tree is a linear BVH (Lauterbach et al. 2009) with the fully parallel hierarchy
// construction of Karras 2012:
//   k_lbvh_keys       63-bit Morton code of every primitive's box centre (21 bits per axis over the centroid bounds)
//   cub radix sort    (key, primitive) pairs — the one library call; everything else is written here
//   k_lbvh_pack       primitive, material record, hittable index in sorted (= leaf slot) order
//   k_lbvh_hierarchy  one thread per inner node: its key range and split from common-prefix lengths -> child references,
//                     parent links, leaf references (ties between equal keys are broken by the sorted position)
//   k_bvh_refit       (refit.cuh) every box bottom-up, the same kernel that serves scene edits
//   k_lbvh_top_order + k_lbvh_permute   the top of the tree is renumbered breadth-first so that the first K nodes are the
//                     hottest ones, which is what the kernels stage in shared memory (crt_internal.h)
//   k_lbvh_stats      depth (the traversal stack is CRT_STACK_SIZE entries) and the SAH cost statistic
// The kernels only need what the closest-hit query needs: conservative boxes and one primitive per leaf. Parity is therefore
// the same bar as for the host builder: primary-hit IDs bit-exact against the oracle (tests/test_gpu_lbvh.py).
#pragma once

#include "device_common.cuh"
#include "refit.cuh"

struct LbvhRaw
{
    const crt_hittable* hittables;
    const crt_material* materials;
    const crt_texture* textures;
    const unsigned char* image_present; // per image: 1 if it has pixels
    const uint32_t* active;             // indices of the active hittables
    uint32_t n;                         // number of active hittables (>= 2)
};

__device__ __forceinline__ void lbvh_prim_box(const crt_hittable& h, float lo[3], float hi[3])
{
    if (h.type == CRT_SPHERE) { // Hittable.cuh:112-116
        const float r = fabsf(h.a);
        lo[0] = h.cx - r; lo[1] = h.cy - r; lo[2] = h.cz - r;
        hi[0] = h.cx + r; hi[1] = h.cy + r; hi[2] = h.cz + r;
        return;
    }
    const float hw = h.a / 2, hh = h.b / 2; // Hittable.cuh:142-146,171-181
    if (h.type == CRT_XYRECT) { lo[0] = h.cx - hw; hi[0] = h.cx + hw; lo[1] = h.cy - hh; hi[1] = h.cy + hh; lo[2] = h.cz - 0.0001f; hi[2] = h.cz + 0.0001f; }
    else if (h.type == CRT_XZRECT) { lo[0] = h.cx - hw; hi[0] = h.cx + hw; lo[2] = h.cz - hh; hi[2] = h.cz + hh; lo[1] = h.cy - 0.0001f; hi[1] = h.cy + 0.0001f; }
    else { lo[1] = h.cy - hh; hi[1] = h.cy + hh; lo[2] = h.cz - hw; hi[2] = h.cz + hw; lo[0] = h.cx - 0.0001f; hi[0] = h.cx + 0.0001f; }
}

// spreads the low 21 bits of v so that there are two zero bits between consecutive bits
__device__ __forceinline__ unsigned long long lbvh_spread3(unsigned int v)
{
    unsigned long long x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

// cmin / inv_extent: centroid bounds of the scene (computed on the host while it validates the hittable list).
// A primitive whose box has at least `big_area` surface (a quarter of the scene box's: a ground sphere, a room wall) gets the
// otherwise unused top key bit: such boxes overlap everything, and wherever their centre happens to fall on the curve they
// would inflate every ancestor up to the root. With the top bit they hang directly under the root, next to the tree of the
// ordinary primitives — where the SAH builder puts them too.
__global__ void k_lbvh_keys(LbvhRaw R, float3 cmin, float3 inv_extent, float big_area, unsigned long long* __restrict__ keys,
                            uint32_t* __restrict__ vals)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= R.n) return;
    const crt_hittable h = R.hittables[R.active[k]];
    float lo[3], hi[3];
    lbvh_prim_box(h, lo, hi);
    const float c[3] = {0.5f * lo[0] + 0.5f * hi[0], 0.5f * lo[1] + 0.5f * hi[1], 0.5f * lo[2] + 0.5f * hi[2]};
    const float m[3] = {cmin.x, cmin.y, cmin.z}, ie[3] = {inv_extent.x, inv_extent.y, inv_extent.z};
    unsigned int q[3];
    for (int a = 0; a < 3; a++) {
        float t = (c[a] - m[a]) * ie[a]; // [0, 1]; NaN / inf (degenerate input) land in cell 0
        t = t >= 0.f ? t : 0.f;
        t = t <= 1.f ? t : 1.f;
        q[a] = min((unsigned int)(t * 2097152.0f), 2097151u);
    }
    const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    const bool big = dx * dy + dy * dz + dz * dx >= big_area;
    keys[k] = big ? (1ull << 63 | k) : (lbvh_spread3(q[0]) << 2 | lbvh_spread3(q[1]) << 1 | lbvh_spread3(q[2]));
    vals[k] = k;
}

// Device copies of crt_pack_prim / crt_pack_material (scene_pack.cpp): same records, same bits.
__global__ void k_lbvh_pack(LbvhRaw R, const uint32_t* __restrict__ sorted, float4* __restrict__ prims, float* __restrict__ prim_k,
                            float4* __restrict__ mats, int* __restrict__ prim_id, int* __restrict__ leaf_code)
{
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= R.n) return;
    const uint32_t idx = R.active[sorted[s]];
    const crt_hittable h = R.hittables[idx];
    const float hw = h.a / 2, hh = h.b / 2;
    float4 p;
    float k = 0.f;
    if (h.type == CRT_SPHERE) p = make_float4(h.cx, h.cy, h.cz, h.a);
    else if (h.type == CRT_XYRECT) { p = make_float4(h.cx - hw, h.cx + hw, h.cy - hh, h.cy + hh); k = h.cz; }
    else if (h.type == CRT_XZRECT) { p = make_float4(h.cx - hw, h.cx + hw, h.cz - hh, h.cz + hh); k = h.cy; }
    else { p = make_float4(h.cy - hh, h.cy + hh, h.cz - hw, h.cz + hw); k = h.cx; }
    prims[s] = p;
    prim_k[s] = k;
    prim_id[s] = (int)idx;
    const crt_material m = R.materials[h.material];
    float4 m0 = make_float4(0.f, 0.f, 0.f, 0.f), m1 = make_float4(0.f, 0.f, 0.f, 0.f);
    uint32_t bits = (uint32_t)m.type;
    float param = m.param;
    if (m.type == CRT_METAL) param = m.param < 1 ? m.param : 1;
    if (m.type == CRT_DIFFUSELIGHT) param = (float)m.light_intensity;
    if (m.type != CRT_DIELECTRIC) {
        const crt_texture t = R.textures[m.texture];
        bits |= (uint32_t)t.type << 4;
        if (t.type == CRT_CONSTANT) { m0.x = t.color[0]; m0.y = t.color[1]; m0.z = t.color[2]; }
        else if (t.type == CRT_CHECKER) {
            m0.x = t.odd[0]; m0.y = t.odd[1]; m0.z = t.odd[2];
            m1.x = t.even[0]; m1.y = t.even[1]; m1.z = t.even[2];
        }
        else {
            const bool has = t.image >= 0 && R.image_present[t.image] != 0;
            bits |= (uint32_t)(has ? t.image + 1 : 0) << 8;
        }
    }
    else m0.x = m0.y = m0.z = 1.0f;
    m0.w = param;
    m1.w = __uint_as_float(bits);
    mats[2 * (size_t)s] = m0;
    mats[2 * (size_t)s + 1] = m1;
    leaf_code[s] = ~(int)(((uint32_t)h.type << CRT_LEAF_TYPE_SHIFT) | ((uint32_t)m.type << CRT_LEAF_MAT_SHIFT) | s);
}

// Length of the common prefix of the keys at sorted positions i and j (-1 outside the array); equal keys are told apart by
// their positions, so that the "keys" are distinct and the hierarchy is a proper binary tree.
__device__ __forceinline__ int lbvh_delta(const unsigned long long* __restrict__ keys, int n, int i, int j)
{
    if (j < 0 || j >= n) return -1;
    const unsigned long long a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((unsigned int)i ^ (unsigned int)j);
    return __clzll((long long)(a ^ b));
}

// Karras 2012, algorithm 3: inner node i covers the sorted range [min(i, j), max(i, j)] and splits after gamma.
__global__ void k_lbvh_hierarchy(const unsigned long long* __restrict__ keys, const int* __restrict__ leaf_code, int n, float4* nodes,
                                 int* __restrict__ parent, int* __restrict__ leaf_ref)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = lbvh_delta(keys, n, i, i - d);
    int lmax = 2;
    while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (lbvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = lbvh_delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (lbvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int first = min(i, j), last = max(i, j);
    int child[2];
    child[0] = first == gamma ? leaf_code[gamma] : gamma;
    child[1] = last == gamma + 1 ? leaf_code[gamma + 1] : gamma + 1;
    if (first == gamma) leaf_ref[gamma] = 2 * i; else parent[gamma] = 2 * i;
    if (last == gamma + 1) leaf_ref[gamma + 1] = 2 * i + 1; else parent[gamma + 1] = 2 * i + 1;
    if (i == 0) parent[0] = -1;
    nodes[4 * (size_t)i + 3] = make_float4(__int_as_float(child[0]), __int_as_float(child[1]), 0.f, 0.f);
}

// Depth of the deepest leaf (levels of inner nodes above it) and the SAH statistic of scene_pack.cpp (sum of the half areas
// of every node and leaf box over the root's). out[0] = max depth (int), out[1] = float bits of the area sum accumulator.
struct LbvhStats { int max_depth; float area_sum; float root_area; int pad; };

__device__ __forceinline__ float lbvh_half_area(const float* b) { return b[3] * b[4] + b[4] * b[5] + b[5] * b[3]; } // of half-extents: x4 cancels

__global__ void k_lbvh_stats(const float4* __restrict__ nodes, const int* __restrict__ parent, const int* __restrict__ leaf_ref, int n,
                             LbvhStats* out)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    int depth = 0;
    float area = 0.f;
    if (s < n) {
        int ref = leaf_ref[s];
        area = lbvh_half_area(reinterpret_cast<const float*>(nodes + 4 * (size_t)(ref >> 1)) + 6 * (ref & 1));
        while (true) { // climb to the root
            depth++;
            const int up = parent[ref >> 1];
            if (up < 0) break;
            ref = up;
        }
        if (s < n - 1 && s > 0) { // inner node s (not the root): its own box sits in its parent
            const int pr = parent[s];
            area += lbvh_half_area(reinterpret_cast<const float*>(nodes + 4 * (size_t)(pr >> 1)) + 6 * (pr & 1));
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        depth = max(depth, __shfl_xor_sync(0xffffffffu, depth, o));
        area += __shfl_xor_sync(0xffffffffu, area, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMax(&out->max_depth, depth);
        atomicAdd(&out->area_sum, area);
    }
    if (s == 0) { // root box = union of its two child boxes
        const float* b = reinterpret_cast<const float*>(nodes);
        float h[3];
        for (int a = 0; a < 3; a++) {
            const float lo = fminf(b[a] - b[3 + a], b[6 + a] - b[9 + a]), hi = fmaxf(b[a] + b[3 + a], b[6 + a] + b[9 + a]);
            h[a] = 0.5f * (hi - lo);
        }
        out->root_area = h[0] * h[1] + h[1] * h[2] + h[2] * h[0];
    }
}

// Breadth-first order of the top of the tree: one warp walks the levels, appending inner children in order (left before
// right, parents in frontier order) with ballots, so the numbering is deterministic. Writes the permutation patches:
// perm[] must hold the identity on entry; on exit the first `m` positions are the first m nodes of the breadth-first order
// and the nodes they displaced have taken the vacated positions.
#define LBVH_TOP_MAX 1024

__global__ void k_lbvh_top_order(const float4* __restrict__ nodes, int n_nodes, int want, int* __restrict__ perm)
{
    __shared__ int order[LBVH_TOP_MAX];
    __shared__ unsigned char in_top[LBVH_TOP_MAX];
    const int lane = threadIdx.x;
    const int m = min(min(want, LBVH_TOP_MAX), n_nodes);
    for (int i = lane; i < LBVH_TOP_MAX; i += 32) in_top[i] = 0;
    if (lane == 0) order[0] = 0;
    __syncwarp();
    int head = 0, tail = 1;
    while (head < tail && tail < m) {
        const int level_end = tail;
        for (int base = head; base < level_end && tail < m; base += 32) {
            const int q = base + lane;
            int c0 = -1, c1 = -1;
            if (q < level_end) {
                const float4 w3 = nodes[4 * (size_t)order[q] + 3];
                c0 = __float_as_int(w3.x);
                c1 = __float_as_int(w3.y);
            }
            const unsigned m0 = __ballot_sync(0xffffffffu, c0 >= 0), m1 = __ballot_sync(0xffffffffu, c1 >= 0);
            const unsigned lt = (1u << lane) - 1u;
            // position of this lane's children: all children of earlier lanes first, own left before own right
            const int before = __popc(m0 & lt) + __popc(m1 & lt);
            int pos = tail + before;
            if (c0 >= 0) { if (pos < m) order[pos] = c0; pos++; }
            if (c1 >= 0 && pos < m) order[pos] = c1;
            tail = min(m, tail + __popc(m0) + __popc(m1));
            __syncwarp();
        }
        head = level_end;
    }
    const int got = tail; // nodes placed (== m unless the tree is smaller)
    __syncwarp();
    for (int q = lane; q < got; q += 32)
        if (order[q] < got) in_top[order[q]] = 1;
    __syncwarp();
    if (lane == 0) { // pair the top nodes that come from outside [0, got) with the positions' old occupants that must leave
        int b = 0;
        for (int q = 0; q < got; q++) {
            const int old = order[q];
            perm[old] = q;
            if (old >= got) {
                while (in_top[b]) b++;
                perm[b] = old;
                b++;
            }
        }
    }
}

__global__ void k_lbvh_identity(int* perm, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) perm[i] = i;
}

// Moves node i to perm[i] and rewrites every reference to a node through perm.
__global__ void k_lbvh_permute(const float4* __restrict__ src, float4* __restrict__ dst, const int* __restrict__ perm,
                               const int* __restrict__ parent_src, int* __restrict__ parent_dst, int n_nodes)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const int ni = perm[i];
    const float4* s = src + 4 * (size_t)i;
    float4* d = dst + 4 * (size_t)ni;
    d[0] = s[0]; d[1] = s[1]; d[2] = s[2];
    const float4 w3 = s[3];
    int c0 = __float_as_int(w3.x), c1 = __float_as_int(w3.y);
    if (c0 >= 0) c0 = perm[c0];
    if (c1 >= 0) c1 = perm[c1];
    d[3] = make_float4(__int_as_float(c0), __int_as_float(c1), 0.f, 0.f);
    const int p = parent_src[i];
    parent_dst[ni] = p < 0 ? p : 2 * perm[p >> 1] + (p & 1);
}

__global__ void k_lbvh_permute_leaf_refs(int* __restrict__ leaf_ref, const int* __restrict__ perm, int n)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const int r = leaf_ref[s];
    leaf_ref[s] = 2 * perm[r >> 1] + (r & 1);
}
