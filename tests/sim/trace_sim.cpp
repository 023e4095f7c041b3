// trace_sim.cpp — TEST INFRASTRUCTURE: cudaraytracer_b200/csrc/trace.cuh (the BVH traversal of every kernel) compiled for the
// HOST with g++, so that its logic can be checked without a GPU: the closest hit of the traversal — in each of its forms: the
// megakernel's node_step loop, the pipelined kernel's node, node, node, leaf rounds with node_step_all (80-byte staged or global
// nodes) and with node_visit_pair (child-interleaved nodes, FFMA2 emulated by two fmaf, sentinel stack; generic and
// sphere-only leaf visits) — must be the closest hit of a brute-force loop over all primitives, bit for bit,
// on the trees crt_pack_scene builds. Only the approximate reciprocal / square root differ from the device (IEEE here), which
// no comparison in this file depends on: every form runs the same arithmetic.
//
//   trace_sim <scene kind 1..5> <size hint> <rays> <seed>      exit code 0 = all forms agree
#define CRT_HOST_SIM 1
#include <cuda_runtime.h> // vector types only

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#undef __device__
#undef __host__
#undef __forceinline__
#undef __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __global__

// the few device intrinsics the headers use, with their host meaning
static inline int __float_as_int(float f) { int i; memcpy(&i, &f, 4); return i; }
static inline float __int_as_float(int i) { float f; memcpy(&f, &i, 4); return f; }
static inline unsigned __float_as_uint(float f) { unsigned i; memcpy(&i, &f, 4); return i; }
static inline float __uint_as_float(unsigned i) { float f; memcpy(&f, &i, 4); return f; }
template <class T> static inline T __ldg(const T* p) { return *p; }
static inline float __fmul_rn(float a, float b) { volatile float r = a * b; return r; }
static inline float __fadd_rn(float a, float b) { volatile float r = a + b; return r; }
static inline float __fsub_rn(float a, float b) { volatile float r = a - b; return r; }
static inline float __fdiv_rn(float a, float b) { volatile float r = a / b; return r; }
static inline float __fsqrt_rn(float a) { return sqrtf(a); }
static inline float __fdividef(float a, float b) { return a / b; }
static inline long long __float2ll_rn(float x) { return llrintf(x); }
static inline unsigned __umulhi(unsigned a, unsigned b) { return (unsigned)(((unsigned long long)a * b) >> 32); }
using std::max;
using std::min;
static struct { unsigned x, y, z; } threadIdx = {0, 0, 0}, blockDim = {1, 1, 1};

#include "crt_internal.h"
#include "trace.cuh"

namespace {

uint64_t g_state;
double rnd() // splitmix64 -> [0, 1)
{
    uint64_t z = (g_state += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (double)(z >> 11) * (1.0 / 9007199254740992.0);
}

Hit brute_force(const SceneView& S, int n_prims, const std::vector<int>& codes, const TravRay& r, float t_min)
{
    Hit h;
    h.t = 3.402823466e+38f;
    h.leaf = -1;
    for (int i = 0; i < n_prims; i++) {
        float t;
        if (leaf_test<false, true>(S, codes[i], r.o, r.d, r.a, r.inv_a, t_min, h.t, t)) { h.t = t; h.leaf = codes[i]; }
    }
    return h;
}

// one lane of k_render_pipe's traversal: rounds of three node visits and one leaf visit (pipeline.cuh)
long g_node_visits = 0, g_leaf_visits = 0; // statistics of the tree, printed at the end
long g_lane_nodes = 0; // inner nodes the last lane call really visited: every fast form must walk the SAME nodes (one set of boxes)

template <bool SMEM_ALL> Hit pipe_lane(const SceneView& S, const TravRay& r, float t_min, long& visits)
{
    Hit h;
    h.t = 3.402823466e+38f;
    h.leaf = -1;
    int stack[CRT_STACK_SIZE], sp = 0;
    int cur = trav_begin<false, SMEM_ALL>(S, r, t_min, h);
    while (cur != CRT_TRAV_DONE) {
        for (int k = 0; k < 3; k++) {
            if (cur >= 0) { g_node_visits++; g_lane_nodes++; }
            node_step_all<SMEM_ALL>(S, r, t_min, h.t, cur, sp, stack);
        }
        if (cur < 0 && cur != CRT_TRAV_DONE) { g_leaf_visits++; leaf_step<false, SMEM_ALL>(S, r, t_min, h, cur, sp, stack); }
        if (++visits > 100000000L) { fprintf(stderr, "traversal does not terminate\n"); exit(3); }
        if (sp < 0 || sp > CRT_STACK_SIZE) { fprintf(stderr, "stack pointer out of range: %d\n", sp); exit(3); }
    }
    return h;
}

// the same lane on the PAIR staging: node_visit_pair, sentinel stack, generic or sphere-only leaf visit
template <bool SPH> Hit pair_lane(const SceneView& S, const TravRay& r, float t_min, long& visits)
{
    Hit h;
    h.t = 3.402823466e+38f;
    h.leaf = -1;
    int stack[CRT_STACK_SIZE];
    StackPtr top = stk_begin(stack);
    int cur = trav_begin<false, true, true, SPH>(S, r, t_min, h);
    while (cur != CRT_TRAV_DONE) {
        // a lane that reached a leaf sits the remaining node visits of the round out ON THE DUMMY NODE: it must neither push nor pop
        for (int k = 0; k < 3; k++) {
            StackPtr before = top;
            const bool at = pair_at_node(cur);
            g_lane_nodes += at;
            node_visit_pair(S, r, t_min, h.t, cur, top);
            if (!at && top != before) { fprintf(stderr, "the dummy node moved the stack\n"); exit(3); }
        }
        if (pair_at_leaf(cur)) {
            if (SPH) leaf_step_sphere(S, r, t_min, h, cur, top);
            else leaf_step_pair<true>(S, r, t_min, h, cur, top);
        }
        if (++visits > 100000000L) { fprintf(stderr, "traversal does not terminate\n"); exit(3); }
        if (top - stack < 0 || top - stack > CRT_STACK_SIZE) { fprintf(stderr, "stack pointer out of range: %d\n", (int)(top - stack)); exit(3); }
    }
    // a lane without a ray (cur == CRT_TRAV_DONE) keeps visiting the dummy node: nothing may move
    for (int k = 0; k < 3; k++) {
        StackPtr before = top;
        node_visit_pair(S, r, t_min, h.t, cur, top);
        if (cur != CRT_TRAV_DONE || top != before) { fprintf(stderr, "an idle lane moved on the dummy node\n"); exit(3); }
    }
    if (h.leaf >= 0) h.leaf = pair_leaf_code(h.leaf); // the traversal worked on 16-bit PAIR references
    return h;
}

// the lane of the scenes that do not fit in shared memory: node_visit_pairg on the child-interleaved global table, part of it staged
Hit pairg_lane(const SceneView& S, const TravRay& r, float t_min, long& visits)
{
    Hit h;
    h.t = 3.402823466e+38f;
    h.leaf = -1;
    int stack[CRT_STACK_SIZE];
    StackPtr top = stk_begin(stack);
    int cur = trav_begin<false, false>(S, r, t_min, h);
    while (cur != CRT_TRAV_DONE) {
        for (int k = 0; k < 3; k++) {
            StackPtr before = top;
            const bool at = cur >= 0;
            g_lane_nodes += at;
            node_visit_pairg(S, r, t_min, h.t, cur, top);
            if (!at && top != before) { fprintf(stderr, "the dummy node moved the stack (global path)\n"); exit(3); }
        }
        if (cur < 0 && cur != CRT_TRAV_DONE) leaf_step_pairg<false>(S, r, t_min, h, cur, top);
        if (++visits > 100000000L) { fprintf(stderr, "traversal does not terminate\n"); exit(3); }
        if (top - stack < 0 || top - stack > CRT_STACK_SIZE) { fprintf(stderr, "stack pointer out of range: %d\n", (int)(top - stack)); exit(3); }
    }
    for (int k = 0; k < 3; k++) {
        StackPtr before = top;
        node_visit_pairg(S, r, t_min, h.t, cur, top);
        if (cur != CRT_TRAV_DONE || top != before) { fprintf(stderr, "an idle lane moved on the dummy node (global path)\n"); exit(3); }
    }
    return h;
}

// Statistic only: what a 4-wide tree would cost. The binary tree is collapsed on the fly - a 4-wide node is a
// binary node with each inner child replaced by that child's two children - and traversed near-first with the same slab
// arithmetic; counted are 4-wide node visits, box tests and primitive tests. The closest hit must be the binary tree's.
long g_wide_visits = 0, g_wide_boxes = 0, g_wide_prims = 0;
struct WideCand { float tn; int ref; };
void wide_visit(const SceneView& S, const CrtPackedNode* nodes, int node, const TravRay& r, float t_min, Hit& h)
{
    g_wide_visits++;
    WideCand cand[4];
    int nc = 0;
    auto test = [&](int owner, int c) { // child c of node `owner`: its box is stored in the owner
        float4 w0, w1, w2; int2 ch;
        load_node<false, false>(S, owner, w0, w1, w2, ch);
        float tn0, tf0, tn1, tf1;
        slab2<true>(r, w0, w1, w2, t_min, h.t, tn0, tf0, tn1, tf1);
        g_wide_boxes++;
        if (c == 0 ? tn0 <= tf0 : tn1 <= tf1) { cand[nc].tn = c == 0 ? tn0 : tn1; cand[nc].ref = nodes[owner].child[c]; nc++; }
    };
    for (int c = 0; c < 2; c++) {
        const int ref = nodes[node].child[c];
        if (ref >= 0) { test(ref, 0); test(ref, 1); } // inner child: its two children join the wide node
        else if (ref != CRT_TRAV_DONE) test(node, c);
    }
    std::sort(cand, cand + nc, [](const WideCand& a, const WideCand& b) { return a.tn < b.tn; });
    for (int k = 0; k < nc; k++) {
        if (cand[k].tn > h.t) continue; // culled by a hit found in a nearer child
        if (cand[k].ref >= 0) wide_visit(S, nodes, cand[k].ref, r, t_min, h);
        else {
            g_wide_prims++;
            float t;
            const int code = ~cand[k].ref;
            if (leaf_test<false, false>(S, code, r.o, r.d, r.a, r.inv_a, t_min, h.t, t)) { h.t = t; h.leaf = code; }
        }
    }
}
Hit wide_lane(const SceneView& S, const CrtPackedNode* nodes, const TravRay& r, float t_min)
{
    Hit h;
    h.t = 3.402823466e+38f;
    h.leaf = -1;
    const int cur = trav_begin<false, false>(S, r, t_min, h);
    if (cur >= 0) wide_visit(S, nodes, cur, r, t_min, h);
    else if (cur != CRT_TRAV_DONE) {
        float t;
        if (leaf_test<false, false>(S, ~cur, r.o, r.d, r.a, r.inv_a, t_min, h.t, t)) { h.t = t; h.leaf = ~cur; }
    }
    return h;
}

} // namespace

// "half" mode: the host forms of the binary16 helpers (crt_f2h_ru, crt_h2_to_f2, crt_h16_up) on special values and random bit
// patterns, one line of hex words per value: input, binary16 bits, the decoded low / high halves of (h | h << 16), crt_h16_up.
static int half_mode(int n)
{
    std::vector<uint32_t> in = {0u, 0x00000001u, 0x33000000u, 0x33800000u, 0x33800001u, 0x387fc000u, 0x387fffffu, 0x38800000u, 0x38800001u,
                                0x3f800000u, 0x3f800001u, 0x3f801fffu, 0x3f802000u, 0x3f802001u, 0x477fe000u, 0x477fe001u, 0x477fffffu, 0x47800000u,
                                0x7f7fffffu, 0x7f800000u, 0x7fc00000u, 0xbf800001u};
    uint64_t st = 12345;
    while ((int)in.size() < n) {
        st = st * 6364136223846793005ull + 1442695040888963407ull;
        uint32_t b = (uint32_t)(st >> 32);
        if (in.size() % 3 == 0) b = (b & 0x807fffffu) | ((uint32_t)(100 + (st >> 20) % 45) << 23); // exponents around binary16's range
        in.push_back(b);
    }
    for (uint32_t b : in) {
        const float x = __uint_as_float(b);
        const unsigned h = crt_f2h_ru(x);
        const float2 f = crt_h2_to_f2(h | (h << 16));
        printf("%08x %04x %08x %08x %08x\n", b, h, __float_as_uint(f.x), __float_as_uint(f.y), __float_as_uint(crt_h16_up(x)));
    }
    return 0;
}

int main(int argc, char** argv)
{
    if (argc > 1 && !strcmp(argv[1], "half")) return half_mode(argc > 2 ? atoi(argv[2]) : 1000);
    const int kind = argc > 1 ? atoi(argv[1]) : CRT_SCENE_RTIOW;
    const uint64_t hint = argc > 2 ? (uint64_t)atoll(argv[2]) : 0;
    const int n_rays = argc > 3 ? atoi(argv[3]) : 20000;
    g_state = argc > 4 ? (uint64_t)atoll(argv[4]) : 1;

    crt_scene_builder* sb = nullptr;
    if (crt_scene_generate(kind, 1984, hint, &sb) != CRT_OK) { fprintf(stderr, "scene: %s\n", crt_last_error_string(nullptr)); return 2; }
    CrtPackedScene P;
    std::string err;
    if (!crt_pack_scene(crt_scene_builder_desc(sb), &P, &err)) { fprintf(stderr, "pack: %s\n", err.c_str()); return 2; }
    const int nn = (int)P.nodes.size(), np = (int)P.stats.n_prims;

    DevScene D;
    memset(&D, 0, sizeof(D));
    D.nodes = reinterpret_cast<const float4*>(P.nodes.data());
    D.prims = reinterpret_cast<const float4*>(P.prims.data());
    D.prim_k = P.prim_k.data();
    D.mats = reinterpret_cast<const float4*>(P.mats.data());
    D.prim_id = P.prim_id.data();
    D.n_nodes = nn; D.n_prims = np;
    int32_t entry = 0, codes[CRT_MAX_TOP] = {0, 0, 0, 0};
    D.n_top = crt_peel_top(P.nodes.data(), P.nodes.size(), (uint64_t)np, CRT_MAX_TOP, &entry, codes);
    D.entry = entry;
    for (int k = 0; k < CRT_MAX_TOP; k++) D.top_code[k] = k < D.n_top ? codes[k] : 0;

    // leaf codes of every primitive (for the brute-force loop), read off the tree
    std::vector<int> leaf_code((size_t)std::max(np, 1), 0);
    int found = 0;
    for (const CrtPackedNode& n : P.nodes)
        for (int c = 0; c < 2; c++)
            if (n.child[c] < 0 && np > 0) {
                const int code = ~n.child[c];
                const int slot = code & CRT_LEAF_INDEX_MASK;
                if (slot < np && (np >= 2 || c == 0)) { leaf_code[(size_t)slot] = code; found++; }
            }
    if (found != np) { fprintf(stderr, "tree names %d leaves, scene has %d primitives\n", found, np); return 2; }

    // the stagings: everything behind 'global' pointers, the megakernel's word-major shared-memory table, the padded (80-byte)
    // table of the one-CTA-per-SM kernels, and that table child-interleaved with byte-offset child references (PAIR)
    D.smem_nodes = D.smem_prims = D.smem_mats = 0;
    std::vector<float4> smem0(16), smem1((size_t)nn * CRT_SMEM_NODE_F4 + (size_t)np * 3 + 16), smem2(smem1.size());
    const SceneView Sg = stage_scene<false>(D, smem0.data());
    D.smem_nodes = nn; D.smem_prims = np; D.smem_mats = np;
    std::vector<float4> smem3(smem1.size());
    const SceneView Sw = stage_scene<false>(D, smem3.data()); // word-major table: what the megakernel stages
    const SceneView Sp = stage_scene<true, false>(D, smem1.data());
    // the PAIR staging exists for scenes that fit in shared memory: 16-bit references (trace.cuh)
    const bool pair_ok = np <= CRT_PAIR_MAX_PRIMS && CRT_PAIR_NODE_F4 * (nn + 1) < CRT_PAIR_LEAF;
    const SceneView So = pair_ok ? stage_scene<true, true>(D, smem2.data()) : Sp;
    // the global path of the pipelined kernel: child-interleaved table behind the 'global' pointer, the first half of it staged
    std::vector<float4> pair_tab((size_t)std::max(nn, 1) * 4);
    for (int i = 0; i < nn; i++) {
        const float4* g = D.nodes + 4 * (size_t)i;
        crt_pair_node(g[0], g[1], g[2], g[3], pair_tab[4 * (size_t)i], pair_tab[4 * (size_t)i + 1], pair_tab[4 * (size_t)i + 2], pair_tab[4 * (size_t)i + 3]);
    }
    D.nodes_pair = pair_tab.data();
    D.smem_nodes = nn / 2; D.smem_prims = 0; D.smem_mats = 0;
    std::vector<float4> smem4((size_t)(nn / 2 + 1) * CRT_PAIR_NODE_F4 + 16);
    const SceneView Sq = stage_scene<true, false, true>(D, smem4.data());
    D.smem_nodes = nn; D.smem_prims = np; D.smem_mats = np;
    bool spheres_only = true;
    for (int c : leaf_code) if (((c >> CRT_LEAF_TYPE_SHIFT) & 3) != CRT_SPHERE) spheres_only = false;

    // rays: from a camera-like point towards the scene, from random points inside the scene box in random directions, and
    // axis-parallel ones (zero direction components take the clamped-reciprocal path of trav_setup)
    float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
    for (int i = 0; i < np; i++)
        for (int a = 0; a < 3; a++) {
            const float c = P.prims[4 * (size_t)i + a];
            if (fabsf(c) < 200.f) { lo[a] = std::min(lo[a], c); hi[a] = std::max(hi[a], c); }
        }
    for (int a = 0; a < 3; a++) if (!(lo[a] <= hi[a])) { lo[a] = -1.f; hi[a] = 1.f; }
    long mism = 0, hits = 0, visits = 0;
    const float t_min = 0.001f;
    for (int i = 0; i < n_rays; i++) {
        float3 o, d;
        const int mode = i % 3;
        if (mode == 0) {
            o = make_float3(13.f + (float)rnd(), 2.f + (float)rnd(), 3.f + (float)rnd());
            d = make_float3((float)(lo[0] + (hi[0] - lo[0]) * rnd()) - o.x, (float)(lo[1] + (hi[1] - lo[1]) * rnd()) - o.y,
                            (float)(lo[2] + (hi[2] - lo[2]) * rnd()) - o.z);
        }
        else {
            o = make_float3((float)(lo[0] + (hi[0] - lo[0]) * rnd()), (float)(lo[1] + (hi[1] - lo[1]) * rnd()) + 0.5f,
                            (float)(lo[2] + (hi[2] - lo[2]) * rnd()));
            d = make_float3((float)(2 * rnd() - 1), (float)(2 * rnd() - 1), (float)(2 * rnd() - 1));
            if (mode == 2) { const int z = (int)(rnd() * 3); (z == 0 ? d.x : z == 1 ? d.y : d.z) = 0.f; }
        }
        const TravRay r = trav_setup<false>(o, d);
        const Hit want = brute_force(Sp, np, leaf_code, r, t_min);
        unsigned nb = 0, npr = 0;
        Hit got[7];
        got[0] = closest_hit<false, false, false>(Sg, o, d, t_min, 3.402823466e+38f, nb, npr); // megakernel form, global tables
        got[1] = closest_hit<false, false, true>(Sw, o, d, t_min, 3.402823466e+38f, nb, npr);  // ... tables in 'shared memory'
        long walked[5];
        g_lane_nodes = 0; got[2] = pipe_lane<true>(Sp, r, t_min, visits); walked[0] = g_lane_nodes;
        g_lane_nodes = 0; got[3] = pipe_lane<false>(Sg, r, t_min, visits); walked[1] = g_lane_nodes; // nodes behind the 'global' pointers (scenes that do not fit)
        g_lane_nodes = 0; got[4] = pair_ok ? pair_lane<false>(So, r, t_min, visits) : got[2]; walked[2] = pair_ok ? g_lane_nodes : walked[0];
        g_lane_nodes = 0; got[5] = pair_ok && spheres_only ? pair_lane<true>(So, r, t_min, visits) : got[4]; walked[3] = pair_ok && spheres_only ? g_lane_nodes : walked[2];
        g_lane_nodes = 0; got[6] = pairg_lane(Sq, r, t_min, visits); walked[4] = g_lane_nodes;
        // Every fast form culls with the binary16-rounded boxes (slab2<H16> / the PAIR tables) and descends near child first:
        // they must visit the same inner nodes, not merely find the same hit - that is what keeps a grazing phantom hit of the
        // reference's sphere formula (trace.cuh sphere_test) either in all variants' frames or in none.
        for (int k = 1; k < 5; k++)
            if (walked[k] != walked[0]) {
                if (mism++ < 5) fprintf(stderr, "ray %d: form %d visited %ld inner nodes, the word-major form %ld\n", i, k + 2, walked[k], walked[0]);
            }
        hits += want.leaf >= 0;
        {
            const Hit w = wide_lane(Sg, P.nodes.data(), r, t_min);
            if (w.leaf != want.leaf || memcmp(&w.t, &want.t, 4) != 0) { if (mism++ < 5) fprintf(stderr, "ray %d 4-wide walk: leaf %d t %.9g, brute force: leaf %d t %.9g\n", i, w.leaf, w.t, want.leaf, want.t); }
        }
        for (int k = 0; k < 7; k++)
            if (got[k].leaf != want.leaf || memcmp(&got[k].t, &want.t, 4) != 0) {
                if (mism++ < 5)
                    fprintf(stderr, "ray %d form %d: leaf %d t %.9g, brute force: leaf %d t %.9g\n", i, k, got[k].leaf, got[k].t, want.leaf, want.t);
            }
    }
    printf("scene %d (%d primitives, %d nodes, %d peeled): %d rays, %ld hits, %ld mismatches\n", kind, np, nn, D.n_top, n_rays, hits, mism);
    printf("tree: SAH cost %.3f, depth %d; per ray %.3f node visits, %.3f leaf visits (+ %d peeled tests)\n", P.stats.sah_cost, (int)P.stats.max_depth,
           (double)g_node_visits / (2.0 * n_rays), (double)g_leaf_visits / (2.0 * n_rays), D.n_top);
    printf("the same tree collapsed 4-wide: per ray %.3f node visits, %.3f box tests, %.3f primitive tests\n", (double)g_wide_visits / n_rays,
           (double)g_wide_boxes / n_rays, (double)g_wide_prims / n_rays);
    crt_scene_builder_free(sb);
    return mism ? 1 : 0;
}
