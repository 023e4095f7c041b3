// trace.cuh — closest-hit query: BVH2 traversal + primitive tests.
//
// Replaces BVHNode::Hit / PerformHit (Hittable.cuh:387-439,470-485), AABB::Hit (AABB.cuh:30-50),
// Sphere::Hit (Hittable.cuh:80-110) and XY/XZ/YZRect::Hit (Hittable.cuh:139-168,195-224,251-280).
// Semantics kept from the reference: t_min = 0.001f in ray-parameter units (Kernel.cu:40), the ray
// direction is NOT normalised, spheres accept a root iff t_min < t < t_max (strict), near root first;
// rects accept iff !(t < t_min || t > t_max) and the point is inside the closed rectangle.
// Only the hit distance and the leaf code are produced here; point, normal and UV are deferred to
// shading (the reference recomputes them, with acos/atan2, for every accepted candidate).
#pragma once

#include "device_common.cuh"

struct SceneView
{
    const float4* g_nodes;
    const float4* g_nodes_pair; // child-interleaved copy of g_nodes (crt_pair_node), for node_visit_pairg
    const float4* g_prims;
    const float* g_prim_k;
    const float4* g_mats;
    const float4* s_nodes;
    const float4* s_prims;
    const float4* s_mats;
    const DevImage* images;
    int smem_nodes, smem_prims, smem_mats;
    int entry, n_top;
    int top_code[CRT_MAX_TOP];
    int top_pair[CRT_MAX_TOP]; // PAIR staging: the peeled leaves as 16-bit references
    int dummy; // PAIR staging: the dummy node that lanes which are not at an inner node visit (node_visit_pair)
};

// 32 bytes through the read-only path in one instruction (PTX 8.8 ld.global.nc.v8, 32-byte aligned address).
__device__ __forceinline__ void ldg256(const float4* p, float4& a, float4& b)
{
#ifdef CRT_HOST_SIM
    a = p[0];
    b = p[1];
#else
    asm("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
        : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
        : "l"(p));
#endif
}

// Two shared-memory layouts of the node table, both chosen so that lanes at different nodes spread over all eight 16-byte
// bank groups (the 64-byte stride of the global layout puts word 0 of every node into two of them):
//   PAD = false  word-major SoA: word k of node i at s_nodes[k * smem_nodes + i]. No wasted bytes: used by the megakernel,
//                whose several resident CTAs each hold a scene copy.
//   PAD = true   80-byte nodes: word k of node i at s_nodes[5 * i + k], the fifth word unused. 80 is an odd multiple of 16,
//                and the four loads of a visit share ONE address register with immediate offsets (the SoA form needs an
//                address computation per word): used by the one-CTA-per-SM kernels, where instruction issue is the bound.
#define CRT_SMEM_NODE_F4 5

template <bool SMEM_ALL, bool PAD>
__device__ __forceinline__ void load_node(const SceneView& S, int i, float4& w0, float4& w1, float4& w2, int2& ch)
{
    if (SMEM_ALL || i < S.smem_nodes) {
        if (PAD) {
            const float4* p = S.s_nodes + CRT_SMEM_NODE_F4 * i;
            w0 = p[0]; w1 = p[1]; w2 = p[2];
            ch = *reinterpret_cast<const int2*>(p + 3);
        }
        else {
            const float4* p = S.s_nodes + i;
            const int n = S.smem_nodes;
            w0 = p[0]; w1 = p[n]; w2 = p[2 * n];
            const float4 w3 = p[3 * n];
            ch = make_int2(__float_as_int(w3.x), __float_as_int(w3.y));
        }
    }
    else {
        // a 64-byte node is two 256-bit read-only loads (sm_100: LDG.E.256), half the L1 requests of four 128-bit ones;
        // on the scenes that do not fit in shared memory the L1 request rate is what the traversal runs into
        const float4* p = S.g_nodes + 4 * (size_t)i;
        float4 w3;
        ldg256(p, w0, w1);
        ldg256(p + 2, w2, w3);
        ch = make_int2(__float_as_int(w3.x), __float_as_int(w3.y));
    }
}

template <bool SMEM_ALL> __device__ __forceinline__ float4 load_prim(const SceneView& S, int i)
{
    if (SMEM_ALL || i < S.smem_prims)
        return S.s_prims[i];
    return __ldg(S.g_prims + i);
}

template <bool SMEM_ALL> __device__ __forceinline__ void load_mat(const SceneView& S, int i, float4& m0, float4& m1)
{
    if (SMEM_ALL || i < S.smem_mats) {
        m0 = S.s_mats[2 * i];
        m1 = S.s_mats[2 * i + 1];
    }
    else {
        m0 = __ldg(S.g_mats + 2 * (size_t)i);
        m1 = __ldg(S.g_mats + 2 * (size_t)i + 1);
    }
}

#define CRT_TRAV_DONE ((int)0x80000000)

// ---- The PAIR staging (one-CTA-per-SM kernels, whole scene in shared memory): 48-byte nodes ----
// The shared-memory pipe is what k_render_pipe runs into once its instruction count is down (profiles/r02_y: issue 73.5 %,
// l1tex 91 %; a node visit is 14 wavefronts of loads against 37 instructions = 9.3 issue cycles per SM). So the staged node is
// made SMALLER at the price of a few more instructions: half-extents as binary16 rounded UP (the box only grows: it culls, it
// never decides a hit), child references as 16-bit codes. A node is 40 bytes in a 48-byte stride (an odd multiple of 16: lanes
// at different nodes spread over all eight 16-byte bank groups), fetched by LDS.128 + LDS.128 + LDS.64 = 10 wavefronts:
//   +0  {c0.x, c1.x, c0.y, c1.y}                 centres, binary32, the two children interleaved (one FFMA2 serves both)
//   +16 {c0.z, c1.z, half2(h0.x, h1.x), half2(h0.y, h1.y)}
//   +32 {half2(h0.z, h1.z), ref0 | ref1 << 16}
// A 16-bit reference: inner node = its offset in 16-byte units (< 0x8000); leaf = 0x8000 | type << 13 | material type << 11 |
// primitive slot (< 2048: more primitives than that do not fit in shared memory next to the path pool). Inside the PAIR
// traversal `cur`, the stack entries and Hit::leaf hold these codes (CRT_TRAV_DONE stays what it is: neither test below
// matches it); pair_leaf_code() turns a leaf reference back into the leaf code of crt_internal.h when the ray is finished.
#define CRT_PAIR_NODE_F4 3
#define CRT_PAIR_LEAF 0x8000
#define CRT_PAIR_MAX_PRIMS 2048
__device__ __forceinline__ int pair_ref(int ref)
{
    if (ref == CRT_TRAV_DONE) return ref;
    if (ref >= 0) return ref * CRT_PAIR_NODE_F4;
    const int code = ~ref;
    return CRT_PAIR_LEAF | (((code >> CRT_LEAF_TYPE_SHIFT) & 3) << 13) | (((code >> CRT_LEAF_MAT_SHIFT) & 3) << 11) | (code & (CRT_PAIR_MAX_PRIMS - 1));
}
__device__ __forceinline__ int pair_leaf_code(int v)
{
    return (((v >> 13) & 3) << CRT_LEAF_TYPE_SHIFT) | (((v >> 11) & 3) << CRT_LEAF_MAT_SHIFT) | (v & (CRT_PAIR_MAX_PRIMS - 1));
}
__device__ __forceinline__ bool pair_at_node(int cur) { return (unsigned)cur < (unsigned)CRT_PAIR_LEAF; }
__device__ __forceinline__ bool pair_at_leaf(int cur) { return (cur & CRT_PAIR_LEAF) != 0; }

// binary32 >= 0 -> binary16 bits, rounded towards +inf (overflow gives +inf: the box side becomes unbounded, still correct)
__device__ __forceinline__ unsigned crt_f2h_ru(float x)
{
#ifdef CRT_HOST_SIM
    const unsigned b = __float_as_uint(fabsf(x));
    if (b >= 0x7f800000u) return 0x7c00u;
    const int e = (int)(b >> 23) - 127;
    if (e > 15) return 0x7c00u;
    if (b == 0u) return 0u;
    if (e >= -14) { // normal in binary16; a carry out of the mantissa moves into the exponent (up to 0x7c00 = +inf)
        unsigned h = ((unsigned)(e + 15) << 10) | ((b & 0x7fffffu) >> 13);
        if (b & 0x1fffu) h++;
        return h;
    }
    const unsigned m = (b & 0x7fffffu) | 0x800000u; // subnormal: units of 2^-24
    const int shift = -e - 1; // value = m * 2^(e - 23) = (m >> shift) * 2^-24
    if (shift >= 32) return 1u;
    unsigned h = m >> shift;
    if (m & ((1u << shift) - 1u)) h++;
    return h;
#else
    return (unsigned)__half_as_ushort(__float2half_ru(fabsf(x)));
#endif
}
// two binary16 in one word -> {low, high} as binary32
__device__ __forceinline__ float2 crt_h2_to_f2(unsigned w)
{
#ifdef CRT_HOST_SIM
    float2 r;
    for (int k = 0; k < 2; k++) {
        const unsigned h = (w >> (16 * k)) & 0xffffu, sgn = (h & 0x8000u) << 16, e = (h >> 10) & 31u, m = h & 0x3ffu;
        float v;
        if (e == 31u) v = __uint_as_float(sgn | 0x7f800000u | (m << 13));
        else if (e == 0u) v = (sgn ? -1.f : 1.f) * (float)m * 5.9604644775390625e-8f; // m * 2^-24
        else v = __uint_as_float(sgn | ((e + 112u) << 23) | (m << 13));
        (k ? r.y : r.x) = v;
    }
    return r;
#else
    return __half22float2(*reinterpret_cast<const __half2*>(&w));
#endif
}

// ---- The child-interleaved node of the pipelined kernel's GLOBAL path (scenes that do not fit in shared memory) ----
// The PAIR node with 32-bit child references (those of crt_internal.h), 44 bytes:
//   {c0.x, c1.x, c0.y, c1.y} {c0.z, c1.z, half2(h0.x, h1.x), half2(h0.y, h1.y)} {half2(h0.z, h1.z), child0, child1, -}
// In HBM one record per 64 bytes, fetched by ONE LDG.256 + ONE LDG.128 = 12 wavefronts of the LSU data pipe per warp visit
// instead of the 16 of two LDG.256 on a binary32 node; staged prefix: 48-byte slots, three LDS.128. That pipe is the bound on
// these scenes (profiles/r02_y3_pipe_c4scene.txt: l1tex__data_pipe_lsu_wavefronts 90 %, half of it global node loads, issue 56 %).
// The table is made on the device (k_pair_nodes) whenever the packed nodes have changed; node_visit_pairg reads it.
__device__ __forceinline__ void crt_pair_node(const float4& g0, const float4& g1, const float4& g2, const float4& g3, float4& q0, float4& q1,
                                              float4& q2, float4& q3)
{
    // packed: {c0.c.xyz, c0.h.x} {c0.h.yz, c1.c.xy} {c1.c.z, c1.h.xyz} {child0, child1}
    q0 = make_float4(g0.x, g1.z, g0.y, g1.w);
    q1 = make_float4(g0.z, g2.x, __uint_as_float(crt_f2h_ru(g0.w) | (crt_f2h_ru(g2.y) << 16)), __uint_as_float(crt_f2h_ru(g1.x) | (crt_f2h_ru(g2.z) << 16)));
    q2 = make_float4(__uint_as_float(crt_f2h_ru(g1.y) | (crt_f2h_ru(g2.w) << 16)), g3.x, g3.y, 0.f);
    q3 = make_float4(0.f, 0.f, 0.f, 0.f);
}

// Stages the first smem_* entries of each table into dynamic shared memory. Call from every thread, then __syncthreads().
// PAIR (needs PAD and the whole tree in shared memory): the 48-byte nodes above, plus a dummy node behind the last one.
// PAIRG (needs PAD): 48-byte slots holding the global path's node records; slot 0 is the dummy node, node i sits in slot i + 1.
template <bool PAD, bool PAIR = false, bool PAIRG = false> __device__ __forceinline__ SceneView stage_scene(const DevScene& D, float4* smem)
{
    SceneView S;
    S.g_nodes = D.nodes; S.g_nodes_pair = D.nodes_pair; S.g_prims = D.prims; S.g_prim_k = D.prim_k; S.g_mats = D.mats; S.images = D.images;
    S.smem_nodes = D.smem_nodes; S.smem_prims = D.smem_prims; S.smem_mats = D.smem_mats;
    S.entry = D.entry; S.n_top = D.n_top;
    for (int k = 0; k < CRT_MAX_TOP; k++) { S.top_code[k] = D.top_code[k]; S.top_pair[k] = PAIR ? pair_ref(~D.top_code[k]) : 0; }
    float4* sn = smem;
    float4* sp = sn + (PAIR ? CRT_PAIR_NODE_F4 * (size_t)(D.smem_nodes + 1) : PAIRG ? CRT_PAIR_NODE_F4 * (size_t)(D.smem_nodes + 1)
                                                                                    : (PAD ? CRT_SMEM_NODE_F4 : 4) * (size_t)D.smem_nodes);
    float4* sm = sp + D.smem_prims;
    const int tid = threadIdx.x, nt = blockDim.x;
    if (PAIR) {
        for (int i = tid; i <= D.smem_nodes; i += nt) {
            float4 w0, w1;
            uint2 w2;
            if (i < D.smem_nodes) {
                const float4 g0 = __ldg(D.nodes + 4 * i), g1 = __ldg(D.nodes + 4 * i + 1), g2 = __ldg(D.nodes + 4 * i + 2), g3 = __ldg(D.nodes + 4 * i + 3);
                // global: {c0.c.xyz, c0.h.x} {c0.h.yz, c1.c.xy} {c1.c.z, c1.h.xyz} {child0, child1}
                w0 = make_float4(g0.x, g1.z, g0.y, g1.w);
                w1 = make_float4(g0.z, g2.x, __uint_as_float(crt_f2h_ru(g0.w) | (crt_f2h_ru(g2.y) << 16)), __uint_as_float(crt_f2h_ru(g1.x) | (crt_f2h_ru(g2.z) << 16)));
                w2 = make_uint2(crt_f2h_ru(g1.y) | (crt_f2h_ru(g2.w) << 16),
                                ((unsigned)pair_ref(__float_as_int(g3.x)) & 0xffffu) | ((unsigned)pair_ref(__float_as_int(g3.y)) << 16));
            }
            else {
                // The dummy node: child 0 is the whole space (centre 0, half-extent +inf), child 1 is empty (half-extent -inf:
                // near = +inf, far = -inf). A lane that visits it hits child 0 only: no push, no pop.
                const unsigned pm = 0x7c00u | (0xfc00u << 16);
                w0 = make_float4(0.f, 0.f, 0.f, 0.f);
                w1 = make_float4(0.f, 0.f, __uint_as_float(pm), __uint_as_float(pm));
                w2 = make_uint2(pm, 0u);
            }
            sn[CRT_PAIR_NODE_F4 * i] = w0;
            sn[CRT_PAIR_NODE_F4 * i + 1] = w1;
            *reinterpret_cast<uint2*>(sn + CRT_PAIR_NODE_F4 * i + 2) = w2;
        }
        S.entry = pair_ref(D.entry);
    }
    else if (PAIRG) {
        for (int i = tid; i <= D.smem_nodes; i += nt) {
            float4 q0, q1, q2, q3;
            if (i == 0) { // the dummy node (see above) with 32-bit references
                const float pm = __uint_as_float(0x7c00u | (0xfc00u << 16));
                q0 = make_float4(0.f, 0.f, 0.f, 0.f);
                q1 = make_float4(0.f, 0.f, pm, pm);
                q2 = make_float4(pm, 0.f, 0.f, 0.f);
            }
            else {
                const float4* g = D.nodes + 4 * (size_t)(i - 1);
                crt_pair_node(__ldg(g), __ldg(g + 1), __ldg(g + 2), __ldg(g + 3), q0, q1, q2, q3);
            }
            float4* d = sn + CRT_PAIR_NODE_F4 * (size_t)i;
            d[0] = q0; d[1] = q1; d[2] = q2;
        }
    }
    else
    for (int i = tid; i < 4 * D.smem_nodes; i += nt) {
        const float4 v = __ldg(D.nodes + i);
        sn[PAD ? CRT_SMEM_NODE_F4 * (i >> 2) + (i & 3) : (i & 3) * D.smem_nodes + (i >> 2)] = v;
    }
    S.dummy = CRT_PAIR_NODE_F4 * D.smem_nodes; // PAIR: in 16-byte units, like every PAIR node reference
    for (int i = tid; i < D.smem_prims; i += nt) sp[i] = __ldg(D.prims + i);
    for (int i = tid; i < 2 * D.smem_mats; i += nt) sm[i] = __ldg(D.mats + i);
    S.s_nodes = sn; S.s_prims = sp; S.s_mats = sm;
    return S;
}

struct Hit
{
    float t;
    int leaf; // (type << 28) | primitive slot; -1 = miss
};

// Sphere::Hit, Hittable.cuh:82-107. Returns the accepted root or a negative sentinel.
template <bool EXACT>
__device__ __forceinline__ bool sphere_test(float4 sph, float3 o, float3 d, float a, float inv_a, float t_min, float t_max, float& t_out)
{
    if (EXACT) {
        float3 oc = f3(sub_rn(o.x, sph.x), sub_rn(o.y, sph.y), sub_rn(o.z, sph.z));
        float b = dot3_rn(oc, d);
        float c = sub_rn(dot3_rn(oc, oc), mul_rn(sph.w, sph.w));
        float disc = sub_rn(mul_rn(b, b), mul_rn(a, c));
        if (disc > 0.f) {
            float sq = sqrt_rn(disc);
            float t = div_rn(sub_rn(-b, sq), a);
            if (t < t_max && t > t_min) { t_out = t; return true; }
            t = div_rn(add_rn(-b, sq), a);
            if (t < t_max && t > t_min) { t_out = t; return true; }
        }
        return false;
    }
    else {
        // Fast path: the reference's formula (Hittable.cuh:83-88) with fused operations. For a sphere that is small against its
        // distance its discriminant is noisy (both terms ~ a |oc|^2): rays that pass up to |oc| * 3e-4 outside become grazing
        // "hits", and the image of such scenes is STATISTICALLY what the reference renders only with this formula (a robust
        // discriminant moved the mean of the 10 k-sphere scene by 3 % against the oracle: measured, not adopted). Such a phantom hit
        // can lie outside the leaf box, so whether it is found depends on the boxes on the way: every fast path therefore tests
        // the SAME boxes (half-extents rounded up to binary16, see slab2).
        float3 oc = o - f3(sph.x, sph.y, sph.z);
        float b = dot3(oc, d);
        float c = fmaf(-sph.w, sph.w, dot3(oc, oc));
        float disc = fmaf(-a, c, b * b);
        if (disc > 0.f) {
            float sq = fast_sqrt(disc);
            float t = (-b - sq) * inv_a;
            if (t < t_max && t > t_min) { t_out = t; return true; }
            t = (-b + sq) * inv_a;
            if (t < t_max && t > t_min) { t_out = t; return true; }
        }
        return false;
    }
}

// XYRect/XZRect/YZRect::Hit. r = {a0, a1, b0, b1}, k = plane offset. (oa, ob, ok)/(da, db, dk) are the ray
// components along the rect's two in-plane axes and its normal axis.
template <bool EXACT>
__device__ __forceinline__ bool rect_test(float4 r, float k, float oa, float ob, float ok, float da, float db, float dk, float t_min,
                                          float t_max, float& t_out)
{
    float t, pa, pb;
    if (EXACT) {
        float inv = div_rn(1.0f, dk);
        t = mul_rn(sub_rn(k, ok), inv);
        if (t < t_min || t > t_max) return false;
        pa = add_rn(oa, mul_rn(t, da));
        pb = add_rn(ob, mul_rn(t, db));
    }
    else {
        float inv = fast_rcp(dk);
        t = (k - ok) * inv;
        if (t < t_min || t > t_max) return false;
        pa = fmaf(t, da, oa);
        pb = fmaf(t, db, ob);
    }
    if (pa < r.x || pa > r.y || pb < r.z || pb > r.w) return false;
    t_out = t;
    return true;
}

template <bool EXACT, bool SMEM_ALL>
__device__ __forceinline__ bool leaf_test(const SceneView& S, int code, float3 o, float3 d, float a, float inv_a, float t_min,
                                          float t_max, float& t_out)
{
    const int slot = code & CRT_LEAF_INDEX_MASK;
    const int type = (code >> CRT_LEAF_TYPE_SHIFT) & 3;
    float4 p = load_prim<SMEM_ALL>(S, slot);
    if (type == CRT_SPHERE)
        return sphere_test<EXACT>(p, o, d, a, inv_a, t_min, t_max, t_out);
    float k = __ldg(S.g_prim_k + slot);
    if (type == CRT_XYRECT) return rect_test<EXACT>(p, k, o.x, o.y, o.z, d.x, d.y, d.z, t_min, t_max, t_out);
    if (type == CRT_XZRECT) return rect_test<EXACT>(p, k, o.x, o.z, o.y, d.x, d.z, d.y, t_min, t_max, t_out);
    return rect_test<EXACT>(p, k, o.y, o.z, o.x, d.y, d.z, d.x, t_min, t_max, t_out);
}

// Per-ray constants of a closest-hit query.
struct TravRay
{
    float3 o, d;
    float3 id;   // reciprocal direction (components clamped away from zero so that 0 * inf never appears)
    float3 ood;  // o * id: the slab midpoint of an axis is one FFMA, c * id - ood
    float3 aid;  // |id|: the slab half-width of an axis is h * aid
    float a;     // Dot(dir, dir), Hittable.cuh:83
    float inv_a; // fast path only
};

template <bool EXACT> __device__ __forceinline__ TravRay trav_setup(float3 o, float3 d)
{
    TravRay r;
    r.o = o;
    r.d = d;
    const float tiny = 1e-18f;
    float3 dc = f3(fabsf(d.x) < tiny ? copysignf(tiny, d.x) : d.x, fabsf(d.y) < tiny ? copysignf(tiny, d.y) : d.y,
                   fabsf(d.z) < tiny ? copysignf(tiny, d.z) : d.z);
    if (!EXACT) dc = f3(fabsf(d.x) < tiny ? tiny : d.x, fabsf(d.y) < tiny ? tiny : d.y, fabsf(d.z) < tiny ? tiny : d.z);
    if (EXACT) {
        r.a = dot3_rn(d, d);
        r.inv_a = 0.f;
    }
    else {
        r.a = dot3(d, d);
        r.inv_a = fast_rcp(r.a);
    }
    // The box test only culls (the boxes are padded on the host), so it uses the same fast arithmetic in both paths;
    // EXACT governs the primitive tests, which decide the result.
    r.id = f3(fast_rcp(dc.x), fast_rcp(dc.y), fast_rcp(dc.z));
    r.ood = f3(o.x * r.id.x, o.y * r.id.y, o.z * r.id.z);
    r.aid = f3(fabsf(r.id.x), fabsf(r.id.y), fabsf(r.id.z));
    return r;
}

// Slab intervals of the two child boxes of a node (centre / half-extent form): per axis t_mid = c * id - ood and the
// half-width h * |id|, so near = t_mid - hw, far = t_mid + hw for either sign of the direction: 9 FFMA + 2 three-input
// min/max per box, instead of 6 FFMA + 10 min/max with lo/hi planes (the ALU pipe that runs min/max was the busiest one).
// H16: the half-extents are first rounded UP to binary16, which makes the boxes of this visit exactly those of the PAIR nodes
// (node_visit_pair / node_visit_pairg store them that way). Every fast path sets it, so that all variants cull with the same
// boxes: the reference's sphere formula admits grazing hits outside the leaf box (sphere_test), and a frame can only be
// independent of the variant if such a hit is found by all of them or by none. The EXACT paths keep the binary32 boxes.
__device__ __forceinline__ float crt_h16_up(float x)
{
#ifdef CRT_HOST_SIM
    return crt_h2_to_f2(crt_f2h_ru(x)).x;
#else
    return __half2float(__float2half_ru(x));
#endif
}
template <bool H16>
__device__ __forceinline__ void slab2(const TravRay& r, float4 w0, float4 w1, float4 w2, float t_min, float t_hit, float& tn0, float& tf0,
                                      float& tn1, float& tf1)
{
    if (H16) {
        w0.w = crt_h16_up(w0.w), w1.x = crt_h16_up(w1.x), w1.y = crt_h16_up(w1.y);
        w2.y = crt_h16_up(w2.y), w2.z = crt_h16_up(w2.z), w2.w = crt_h16_up(w2.w);
    }
    const float mx0 = fmaf(w0.x, r.id.x, -r.ood.x), my0 = fmaf(w0.y, r.id.y, -r.ood.y), mz0 = fmaf(w0.z, r.id.z, -r.ood.z);
    const float mx1 = fmaf(w1.z, r.id.x, -r.ood.x), my1 = fmaf(w1.w, r.id.y, -r.ood.y), mz1 = fmaf(w2.x, r.id.z, -r.ood.z);
    tn0 = fmaxf(fmaxf(fmaf(-w0.w, r.aid.x, mx0), fmaf(-w1.x, r.aid.y, my0)), fmaxf(fmaf(-w1.y, r.aid.z, mz0), t_min));
    tf0 = fminf(fminf(fmaf(w0.w, r.aid.x, mx0), fmaf(w1.x, r.aid.y, my0)), fminf(fmaf(w1.y, r.aid.z, mz0), t_hit));
    tn1 = fmaxf(fmaxf(fmaf(-w2.y, r.aid.x, mx1), fmaf(-w2.z, r.aid.y, my1)), fmaxf(fmaf(-w2.w, r.aid.z, mz1), t_min));
    tf1 = fminf(fminf(fmaf(w2.y, r.aid.x, mx1), fmaf(w2.z, r.aid.y, my1)), fminf(fmaf(w2.w, r.aid.z, mz1), t_hit));
}

// d = a * {s, s} + c on both halves of a register pair: one FFMA2 (fma.rn.f32x2). Each half is the IEEE fused multiply-add
// fmaf() computes, so the paired visit produces the bits of slab2().
__device__ __forceinline__ float2 fma2s(float2 a, float s, float2 c)
{
#ifdef CRT_HOST_SIM
    return make_float2(fmaf(a.x, s, c.x), fmaf(a.y, s, c.y));
#else
    float2 d;
    asm("{.reg .b64 ra, rb, rc, rd; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%4}; mov.b64 rc, {%5,%6}; fma.rn.f32x2 rd, ra, rb, rc; mov.b64 {%0,%1}, rd;}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(s), "f"(c.x), "f"(c.y));
    return d;
#endif
}
// d = a * {s, s} + {t, t}
__device__ __forceinline__ float2 fma2ss(float2 a, float s, float t)
{
#ifdef CRT_HOST_SIM
    return make_float2(fmaf(a.x, s, t), fmaf(a.y, s, t));
#else
    float2 d;
    asm("{.reg .b64 ra, rb, rc, rd; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%4}; mov.b64 rc, {%5,%5}; fma.rn.f32x2 rd, ra, rb, rc; mov.b64 {%0,%1}, rd;}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(s), "f"(t));
    return d;
#endif
}

// The traversal stack of the PAIR kernels: a pointer to the next free entry of a per-thread array whose entry 0 is a
// CRT_TRAV_DONE sentinel, so that a pop needs no emptiness test. On the device the pointer is the 32-bit local-window address
// and the push / pop of a node visit are four PREDICATED instructions (STL, LDL, two adds) written in PTX: from `if (push) sp++`
// the compiler makes an add and a predicated move per case, plus an address computation per access (10 instructions of the 42
// of a visit, profiles/r02_v_pipe_c2_by_line.txt). The host build (tests/sim) runs the same logic on a plain pointer.
#ifdef CRT_HOST_SIM
typedef int* StackPtr;
__device__ __forceinline__ StackPtr stk_begin(int* stack) { stack[0] = CRT_TRAV_DONE; return stack + 1; }
__device__ __forceinline__ void stk_visit(StackPtr& top, bool push, bool pop, int other, int& next)
{
    if (push) *top = other;
    if (pop) next = top[-1];
    if (push) top++;
    if (pop) top--;
}
__device__ __forceinline__ int stk_pop(StackPtr& top) { return *--top; }
#else
typedef unsigned StackPtr;
__device__ __forceinline__ StackPtr stk_begin(int* stack)
{
    stack[0] = CRT_TRAV_DONE;
    return (unsigned)__cvta_generic_to_local(stack) + 4u;
}
// push and pop exclude each other
__device__ __forceinline__ void stk_visit(StackPtr& top, bool push, bool pop, int other, int& next)
{
    asm volatile("{ .reg .pred p, q; setp.ne.s32 p, %2, 0; setp.ne.s32 q, %3, 0;\n"
                 " @p st.local.b32 [%0], %4;\n @q ld.local.b32 %1, [%0+-4];\n @p add.u32 %0, %0, 4;\n @q add.u32 %0, %0, -4; }"
                 : "+r"(top), "+r"(next) : "r"((int)push), "r"((int)pop), "r"(other) : "memory");
}
__device__ __forceinline__ int stk_pop(StackPtr& top)
{
    int v;
    asm volatile("add.u32 %0, %0, -4;\n ld.local.b32 %1, [%0];" : "+r"(top), "=r"(v) : : "memory");
    return v;
}
#endif

// The inner-node visit of the kernels that hold the WHOLE tree in shared memory in the PAIR layout (stage_scene<true, true>),
// executed by every lane of the warp like node_step_all. `cur` is a 16-bit PAIR reference (or CRT_TRAV_DONE).
// A lane that is NOT at an inner node (it waits for the leaf visit of the round, or it has no ray) visits the dummy node staged
// behind the tree: child 0 always hits, child 1 never does, so it neither pushes nor pops and the hit tests need no `at` term -
// PROVIDED t_hit >= t_min in every lane at all times (a lane without a ray keeps the t of its last one, or FLT_MAX).
// Per visit: 3 LDS (10 wavefronts) + 6 binary16 conversions + 9 FFMA2 + 8 min/max + 3 FSETP + the reference select, the four
// stack instructions, no branch.
__device__ __forceinline__ void node_visit_pair(const SceneView& S, const TravRay& r, float t_min, float t_hit, int& cur, StackPtr& top)
{
    const bool at = pair_at_node(cur);
    const char* p = reinterpret_cast<const char*>(S.s_nodes) + 16 * (at ? cur : S.dummy);
    const float4 w0 = *reinterpret_cast<const float4*>(p), w1 = *reinterpret_cast<const float4*>(p + 16);
    const uint2 w2 = *reinterpret_cast<const uint2*>(p + 32);
    const float2 hx = crt_h2_to_f2(__float_as_uint(w1.z)), hy = crt_h2_to_f2(__float_as_uint(w1.w)), hz = crt_h2_to_f2(w2.x);
    const float2 mx = fma2ss(make_float2(w0.x, w0.y), r.id.x, -r.ood.x);
    const float2 my = fma2ss(make_float2(w0.z, w0.w), r.id.y, -r.ood.y);
    const float2 mz = fma2ss(make_float2(w1.x, w1.y), r.id.z, -r.ood.z);
    const float2 nx = fma2s(make_float2(-hx.x, -hx.y), r.aid.x, mx), fx = fma2s(hx, r.aid.x, mx);
    const float2 ny = fma2s(make_float2(-hy.x, -hy.y), r.aid.y, my), fy = fma2s(hy, r.aid.y, my);
    const float2 nz = fma2s(make_float2(-hz.x, -hz.y), r.aid.z, mz), fz = fma2s(hz, r.aid.z, mz);
    const float tn0 = fmaxf(fmaxf(nx.x, ny.x), fmaxf(nz.x, t_min)), tf0 = fminf(fminf(fx.x, fy.x), fminf(fz.x, t_hit));
    const float tn1 = fmaxf(fmaxf(nx.y, ny.y), fmaxf(nz.y, t_min)), tf1 = fminf(fminf(fx.y, fy.y), fminf(fz.y, t_hit));
    const bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
    const bool second_first = h1 && (tn1 < tn0 || !h0); // child 1 is the nearer hit child, or the only one
    const int r0 = (int)(w2.y & 0xffffu), r1 = (int)(w2.y >> 16);
    const int first = second_first ? r1 : r0, other = second_first ? r0 : r1;
    int next = first;
    // `at` on the push: a lane that is not at a node reads the dummy node, whose child 1 can only "hit" if the lane's ray constants
    // are not finite (NaN drops out of min / max); such a lane must still not move its stack
    stk_visit(top, at && h0 && h1, !(h0 || h1), other, next);
    if (at) cur = next;
}

// The same visit for scenes that do NOT fit in shared memory (stage_scene<true, false, true>): `cur` >= 0 is a node index, the
// child references are those of crt_internal.h. Nodes below smem_nodes come from the staged prefix (slot i + 1; slot 0 is the
// dummy node, which lanes that are not at an inner node visit as index -1), the others from the global table of crt_pair_node
// with one LDG.256 and one LDG.128. Everything after the loads is the arithmetic of node_visit_pair.
__device__ __forceinline__ void node_visit_pairg(const SceneView& S, const TravRay& r, float t_min, float t_hit, int& cur, StackPtr& top)
{
    const bool at = cur >= 0;
    const int i = at ? cur : -1;
    float4 q0, q1, q2;
    if (i < S.smem_nodes) {
        const float4* p = S.s_nodes + CRT_PAIR_NODE_F4 * (i + 1);
        q0 = p[0]; q1 = p[1]; q2 = p[2];
    }
    else {
        const float4* p = S.g_nodes_pair + 4 * (size_t)i;
        ldg256(p, q0, q1);
        q2 = __ldg(p + 2);
    }
    const float2 hx = crt_h2_to_f2(__float_as_uint(q1.z)), hy = crt_h2_to_f2(__float_as_uint(q1.w)), hz = crt_h2_to_f2(__float_as_uint(q2.x));
    const float2 mx = fma2ss(make_float2(q0.x, q0.y), r.id.x, -r.ood.x);
    const float2 my = fma2ss(make_float2(q0.z, q0.w), r.id.y, -r.ood.y);
    const float2 mz = fma2ss(make_float2(q1.x, q1.y), r.id.z, -r.ood.z);
    const float2 nx = fma2s(make_float2(-hx.x, -hx.y), r.aid.x, mx), fx = fma2s(hx, r.aid.x, mx);
    const float2 ny = fma2s(make_float2(-hy.x, -hy.y), r.aid.y, my), fy = fma2s(hy, r.aid.y, my);
    const float2 nz = fma2s(make_float2(-hz.x, -hz.y), r.aid.z, mz), fz = fma2s(hz, r.aid.z, mz);
    const float tn0 = fmaxf(fmaxf(nx.x, ny.x), fmaxf(nz.x, t_min)), tf0 = fminf(fminf(fx.x, fy.x), fminf(fz.x, t_hit));
    const float tn1 = fmaxf(fmaxf(nx.y, ny.y), fmaxf(nz.y, t_min)), tf1 = fminf(fminf(fx.y, fy.y), fminf(fz.y, t_hit));
    const bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
    const bool second_first = h1 && (tn1 < tn0 || !h0); // child 1 is the nearer hit child, or the only one
    const int r0 = __float_as_int(q2.y), r1 = __float_as_int(q2.z);
    const int first = second_first ? r1 : r0, other = second_first ? r0 : r1;
    int next = first;
    stk_visit(top, at && h0 && h1, !(h0 || h1), other, next);
    if (at) cur = next;
}
// ... and its leaf visit: `cur` = ~leaf code, sentinel stack.
template <bool SMEM_ALL>
__device__ __forceinline__ void leaf_step_pairg(const SceneView& S, const TravRay& r, float t_min, Hit& h, int& cur, StackPtr& top)
{
    const int code = ~cur;
    float t;
    if (leaf_test<false, SMEM_ALL>(S, code, r.o, r.d, r.a, r.inv_a, t_min, h.t, t)) {
        h.t = t;
        h.leaf = code;
    }
    cur = stk_pop(top);
}

// One inner-node visit: slab-test both child boxes against (t_min, t_hit), descend into the nearer hit child and push the
// other. `cur` >= 0 on entry; on exit it is the next node (>= 0), a leaf (~code, < 0) or CRT_TRAV_DONE.
template <bool EXACT, bool SMEM_ALL, bool PAD = false>
__device__ __forceinline__ void node_step(const SceneView& S, const TravRay& r, float t_min, float t_hit, int& cur, int& sp, int* stack)
{
    float4 w0, w1, w2;
    int2 ch;
    load_node<SMEM_ALL, PAD>(S, cur, w0, w1, w2, ch);
    float tn0, tf0, tn1, tf1;
    slab2<!EXACT>(r, w0, w1, w2, t_min, t_hit, tn0, tf0, tn1, tf1);
    // The boxes are padded on the host (scene_pack.cpp); <= keeps degenerate (flat) boxes alive.
    const bool h0 = tn0 <= tf0;
    const bool h1 = tn1 <= tf1;
    if (h0 && h1) {
        const bool swap = tn1 < tn0;
        cur = swap ? ch.y : ch.x;
        stack[sp++] = swap ? ch.x : ch.y;
    }
    else if (h0) cur = ch.x;
    else if (h1) cur = ch.y;
    else cur = sp ? stack[--sp] : CRT_TRAV_DONE;
}

// The same visit executed by EVERY lane of the warp (fast path): lanes that are not at an inner node run it on node 0 and
// discard the result. Inactive lanes cost a SIMT machine the same issue slots either way, and without the branch around the
// body the loop loses its convergence-barrier bookkeeping (profiles/r01_c: 20 uniform instructions per round before this).
template <bool SMEM_ALL>
__device__ __forceinline__ void node_step_all(const SceneView& S, const TravRay& r, float t_min, float t_hit, int& cur, int& sp, int* stack)
{
    const bool at = cur >= 0;
    float4 w0, w1, w2;
    int2 ch;
    load_node<SMEM_ALL, true>(S, at ? cur : 0, w0, w1, w2, ch);
    float tn0, tf0, tn1, tf1;
    slab2<true>(r, w0, w1, w2, t_min, t_hit, tn0, tf0, tn1, tf1);
    const bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
    const bool swap = tn1 < tn0;
    const int first = (h0 && !(h1 && swap)) ? ch.x : ch.y; // the nearer hit child (or the only one)
    const int second = swap ? ch.x : ch.y;
    if (at) {
        if (h0 && h1) stack[sp++] = second;
        if (h0 || h1) cur = first;
        else cur = sp ? stack[--sp] : CRT_TRAV_DONE;
    }
}

// One leaf visit: test the primitive, shrink the hit distance, pop. `cur` < 0 and != CRT_TRAV_DONE on entry.
template <bool EXACT, bool SMEM_ALL>
__device__ __forceinline__ void leaf_step(const SceneView& S, const TravRay& r, float t_min, Hit& h, int& cur, int& sp, const int* stack)
{
    const int code = ~cur;
    float t;
    if (leaf_test<EXACT, SMEM_ALL>(S, code, r.o, r.d, r.a, r.inv_a, t_min, h.t, t)) {
        h.t = t;
        h.leaf = code;
    }
    cur = sp ? stack[--sp] : CRT_TRAV_DONE;
}

// The leaf visit of the PAIR kernels (sentinel stack: the pop is unconditional), any primitive type. `cur` is a PAIR leaf
// reference, and so is what goes into Hit::leaf.
template <bool SMEM_ALL>
__device__ __forceinline__ void leaf_step_pair(const SceneView& S, const TravRay& r, float t_min, Hit& h, int& cur, StackPtr& top)
{
    float t;
    if (leaf_test<false, SMEM_ALL>(S, pair_leaf_code(cur), r.o, r.d, r.a, r.inv_a, t_min, h.t, t)) {
        h.t = t;
        h.leaf = cur;
    }
    cur = stk_pop(top);
}

// The sphere test of a scene that holds only spheres, in shared memory: no type dispatch, no branch (the two roots are tried by
// selects). The arithmetic is sphere_test<false>'s, operation for operation. `ref` is a PAIR leaf reference.
__device__ __forceinline__ void sphere_visit(const SceneView& S, int ref, const TravRay& r, float t_min, Hit& h)
{
    const float4 sph = S.s_prims[ref & (CRT_PAIR_MAX_PRIMS - 1)];
    const float3 oc = r.o - f3(sph.x, sph.y, sph.z);
    const float b = dot3(oc, r.d);
    const float c = fmaf(-sph.w, sph.w, dot3(oc, oc));
    const float disc = fmaf(-r.a, c, b * b);
    const float sq = fast_sqrt(fmaxf(disc, 0.f)); // only used where disc > 0, where it is fast_sqrt(disc)
    const float t1 = (-b - sq) * r.inv_a, t2 = (-b + sq) * r.inv_a;
    const bool ok1 = t1 < h.t && t1 > t_min, ok2 = t2 < h.t && t2 > t_min;
    const bool hit = disc > 0.f && (ok1 || ok2);
    if (hit) {
        h.t = ok1 ? t1 : t2;
        h.leaf = ref;
    }
}
__device__ __forceinline__ void leaf_step_sphere(const SceneView& S, const TravRay& r, float t_min, Hit& h, int& cur, StackPtr& top)
{
    sphere_visit(S, cur, r, t_min, h);
    cur = stk_pop(top);
}

// Start of a closest-hit query: the primitives peeled off the top of the tree (DevScene::top_code) are tested here, without
// their boxes, and the descent starts below them. Returns the node to start at.
// PAIR: Hit::leaf and the returned entry are PAIR references (pipelined kernel with the whole scene staged); SPH: that scene
// holds only spheres.
template <bool EXACT, bool SMEM_ALL, bool PAIR = false, bool SPH = false>
__device__ __forceinline__ int trav_begin(const SceneView& S, const TravRay& r, float t_min, Hit& h)
{
#pragma unroll
    for (int k = 0; k < CRT_MAX_TOP; k++) {
        if (k < S.n_top) {
            const int code = S.top_code[k];
            if (SPH) sphere_visit(S, S.top_pair[k], r, t_min, h); // spheres only, all staged: no type dispatch, no branch
            else {
                float t;
                if (leaf_test<EXACT, SMEM_ALL>(S, code, r.o, r.d, r.a, r.inv_a, t_min, h.t, t)) {
                    h.t = t;
                    h.leaf = PAIR ? S.top_pair[k] : code;
                }
            }
        }
    }
    return S.entry;
}

// Closest hit in (t_min, t_max). Near-first ordered descent with a per-thread stack.
template <bool EXACT, bool COUNT, bool SMEM_ALL>
__device__ __forceinline__ Hit closest_hit(const SceneView& S, float3 o, float3 d, float t_min, float t_max, unsigned& n_box,
                                           unsigned& n_prim)
{
    Hit h;
    h.t = t_max;
    h.leaf = -1;
    const TravRay r = trav_setup<EXACT>(o, d);
    int stack[CRT_STACK_SIZE];
    int sp = 0;
    int cur = trav_begin<EXACT, SMEM_ALL>(S, r, t_min, h); // >= 0: inner node, < 0: leaf (~code), CRT_TRAV_DONE: finished
    if (COUNT) n_prim += (unsigned)S.n_top;
    while (true) {
        node_step<EXACT, SMEM_ALL>(S, r, t_min, h.t, cur, sp, stack);
        if (COUNT) n_box += 2;
        while (cur < 0 && cur != CRT_TRAV_DONE) {
            if (COUNT) n_prim++;
            leaf_step<EXACT, SMEM_ALL>(S, r, t_min, h, cur, sp, stack);
        }
        if (cur == CRT_TRAV_DONE) break;
    }
    return h;
}

// The same query on the PAIR staging (stage_scene<true, true>: 48-byte nodes, 16-bit references) for the one-thread-per-pixel
// kernel when the whole scene is staged: node_visit_pair + the sentinel stack, 44 instead of 53 instructions per node visit.
// Returns the leaf code of crt_internal.h like closest_hit.
template <bool SPH> __device__ __forceinline__ Hit closest_hit_pair(const SceneView& S, float3 o, float3 d, float t_min, float t_max)
{
    Hit h;
    h.t = t_max;
    h.leaf = -1;
    const TravRay r = trav_setup<false>(o, d);
    int stack[CRT_STACK_SIZE];
    StackPtr top = stk_begin(stack);
    int cur = trav_begin<false, true, true, SPH>(S, r, t_min, h);
    while (cur != CRT_TRAV_DONE) {
        node_visit_pair(S, r, t_min, h.t, cur, top); // a no-op on the dummy node if the query starts at a leaf
        while (pair_at_leaf(cur)) {
            if (SPH) leaf_step_sphere(S, r, t_min, h, cur, top);
            else leaf_step_pair<true>(S, r, t_min, h, cur, top);
        }
    }
    if (h.leaf >= 0) h.leaf = pair_leaf_code(h.leaf);
    return h;
}
