// scene_pack.cpp — host BVH build + packing of the scene into the device layout of crt_internal.h.
//
// Replaces the reference's BVHNode constructor (Hittable.cuh:303-385), which splits at the median of
// the LIST ORDER and allocates one managed block per split. The closest hit does not depend on the
// tree shape, so this builder is free to use a spatial binned-SAH split; only the
// bounding boxes (Sphere::BoundingBox Hittable.cuh:112-116, *Rect::BoundingBox :171-181) and the
// inactive-object filter (:311-320) follow the reference.

#include "crt_internal.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <thread>

namespace {

struct Box
{
    float lo[3], hi[3];
    void reset()
    {
        for (int i = 0; i < 3; i++) {
            lo[i] = std::numeric_limits<float>::infinity();
            hi[i] = -std::numeric_limits<float>::infinity();
        }
    }
    void grow(const Box& b)
    {
        for (int i = 0; i < 3; i++) {
            lo[i] = std::min(lo[i], b.lo[i]);
            hi[i] = std::max(hi[i], b.hi[i]);
        }
    }
    void grow(const float* p)
    {
        for (int i = 0; i < 3; i++) {
            lo[i] = std::min(lo[i], p[i]);
            hi[i] = std::max(hi[i], p[i]);
        }
    }
    float half_area() const
    {
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        if (!(dx >= 0.f) || !(dy >= 0.f) || !(dz >= 0.f))
            return 0.f;
        return dx * dy + dy * dz + dz * dx;
    }
};

struct BuildNode
{
    Box box;
    int32_t child[2]; // >= 0 build-node index, < 0: ~prim (index into the active list)
};

struct Builder
{
    std::vector<Box> pbox;
    std::vector<float> cent; // 3 per prim
    std::vector<uint32_t> idx;
    std::vector<BuildNode> nodes;
    std::atomic<int> max_depth{0};
    int par_depth = 10; // levels whose left subtree gets its own thread (CRT_BVH_PAR_DEPTH; 0 = single-threaded build)

    static constexpr int kBins = 64;

    Box range_box(uint32_t b, uint32_t e) const
    {
        Box r;
        r.reset();
        for (uint32_t i = b; i < e; i++)
            r.grow(pbox[idx[i]]);
        return r;
    }

    // Returns the child reference for the range [b, e). A subtree over k primitives has exactly k - 1 inner nodes, so node
    // indices are assigned by range (root of [b, e) at `base`, left subtree from base + 1, right subtree from base + (mid - b)):
    // no shared counter, and big subtrees are built by their own threads.
    int32_t build(uint32_t b, uint32_t e, const Box& bounds, int depth, uint32_t base)
    {
        for (int cur = max_depth.load(std::memory_order_relaxed); depth > cur && !max_depth.compare_exchange_weak(cur, depth);) { }
        if (e - b == 1)
            return ~(int32_t)idx[b];

        Box cb;
        cb.reset();
        for (uint32_t i = b; i < e; i++)
            cb.grow(&cent[3 * (size_t)idx[i]]);

        int best_axis = -1, best_split = -1;
        float best_cost = std::numeric_limits<float>::infinity();
        const uint32_t n = e - b;
        Box lb, rb; // child boxes; for a SAH split they fall out of the bins, no extra pass over the range
        lb.reset();
        rb.reset();
        if (n > 2) {
            // one pass over the range bins all three axes
            Box bins[3][kBins];
            uint32_t cnt[3][kBins];
            float scale[3];
            for (int ax = 0; ax < 3; ax++) {
                float ext = cb.hi[ax] - cb.lo[ax];
                scale[ax] = ext > 0.f ? (float)kBins / ext : 0.f;
                for (int k = 0; k < kBins; k++) {
                    bins[ax][k].reset();
                    cnt[ax][k] = 0;
                }
            }
            for (uint32_t i = b; i < e; i++) {
                const uint32_t p = idx[i];
                const Box& pb = pbox[p];
                for (int ax = 0; ax < 3; ax++) {
                    int k = (int)((cent[3 * (size_t)p + ax] - cb.lo[ax]) * scale[ax]);
                    k = std::min(std::max(k, 0), kBins - 1);
                    bins[ax][k].grow(pb);
                    cnt[ax][k]++;
                }
            }
            for (int ax = 0; ax < 3; ax++) {
                if (!(scale[ax] > 0.f))
                    continue;
                float right_area[kBins];
                uint32_t right_cnt[kBins];
                Box acc;
                acc.reset();
                uint32_t c = 0;
                for (int k = kBins - 1; k > 0; k--) {
                    acc.grow(bins[ax][k]);
                    c += cnt[ax][k];
                    right_area[k] = acc.half_area();
                    right_cnt[k] = c;
                }
                acc.reset();
                c = 0;
                for (int k = 0; k < kBins - 1; k++) {
                    acc.grow(bins[ax][k]);
                    c += cnt[ax][k];
                    if (c == 0 || right_cnt[k + 1] == 0)
                        continue;
                    float cost = acc.half_area() * (float)c + right_area[k + 1] * (float)right_cnt[k + 1];
                    if (cost < best_cost) {
                        best_cost = cost;
                        best_axis = ax;
                        best_split = k;
                    }
                }
            }
            if (best_axis >= 0) {
                for (int k = 0; k < kBins; k++) (k <= best_split ? lb : rb).grow(bins[best_axis][k]);
            }
        }

        uint32_t mid;
        if (best_axis >= 0) {
            float ext = cb.hi[best_axis] - cb.lo[best_axis];
            float scale = (float)kBins / ext;
            float lo = cb.lo[best_axis];
            int ax = best_axis, sp = best_split;
            auto it = std::partition(idx.begin() + b, idx.begin() + e, [&](uint32_t p) {
                int k = (int)((cent[3 * (size_t)p + ax] - lo) * scale);
                k = std::min(std::max(k, 0), kBins - 1);
                return k <= sp;
            });
            mid = (uint32_t)(it - idx.begin());
        }
        else {
            // n == 2, or all centroids coincide: split the range in half along the widest centroid axis.
            int ax = 0;
            float e0 = cb.hi[0] - cb.lo[0], e1 = cb.hi[1] - cb.lo[1], e2 = cb.hi[2] - cb.lo[2];
            if (e1 > e0 && e1 >= e2) ax = 1;
            else if (e2 > e0 && e2 > e1) ax = 2;
            mid = b + n / 2;
            std::nth_element(idx.begin() + b, idx.begin() + mid, idx.begin() + e, [&](uint32_t p, uint32_t q) {
                float cp = cent[3 * (size_t)p + ax], cq = cent[3 * (size_t)q + ax];
                return cp < cq || (cp == cq && p < q);
            });
        }
        bool boxes_from_bins = best_axis >= 0;
        if (mid == b || mid == e) {
            mid = b + n / 2;
            boxes_from_bins = false;
        }
        if (!boxes_from_bins) {
            lb = range_box(b, mid);
            rb = range_box(mid, e);
        }

        int32_t me = (int32_t)base;
        int32_t l, r;
        if (n >= 8192 && depth <= par_depth) { // top of a big tree: left subtree on another thread
            std::thread t([&] { l = build(b, mid, lb, depth + 1, base + 1); });
            r = build(mid, e, rb, depth + 1, base + (mid - b));
            t.join();
        }
        else {
            l = build(b, mid, lb, depth + 1, base + 1);
            r = build(mid, e, rb, depth + 1, base + (mid - b));
        }
        nodes[me].box = bounds;
        nodes[me].child[0] = l;
        nodes[me].child[1] = r;
        return me;
    }
};

// Conservative outward padding so that the fused/approximate slab test on the GPU never culls a
// primitive whose exact test would hit: a few ulps relative to the coordinate, plus an absolute term
// proportional to the scene extent (the error of lo*idir - o*idir grows with |o|, and ray origins
// live inside the scene bounds after the first bounce).
inline float pad_lo(float v, float abs_pad) { return v - (fabsf(v) * 4e-6f + abs_pad); }
inline float pad_hi(float v, float abs_pad) { return v + (fabsf(v) * 4e-6f + abs_pad); }

// Boxes are stored as CENTRE + HALF-EXTENT {c.x, c.y, c.z, h.x, h.y, h.z} (crt_internal.h): the slab test then needs no
// per-plane min/max (trace.cuh). The conversion rounds h UP, so [c - h, c + h] contains the padded [lo, hi].
void crt_box_to_center_half(const float lo[3], const float hi[3], float out[6])
{
    for (int i = 0; i < 3; i++) {
        const float c = 0.5f * lo[i] + 0.5f * hi[i];
        const double hd = std::max((double)hi[i] - (double)c, (double)c - (double)lo[i]);
        float h = (float)hd;
        if ((double)h < hd) h = nextafterf(h, std::numeric_limits<float>::infinity());
        out[i] = c;
        out[3 + i] = h;
    }
}

void store_box(float* w, int slot, const Box& b, bool empty, float abs_pad)
{
    float v[6];
    if (empty) { // always hit: infinite half-extent
        for (int i = 0; i < 3; i++) {
            v[i] = 0.f;
            v[3 + i] = std::numeric_limits<float>::infinity();
        }
    }
    else {
        float lo[3], hi[3];
        for (int i = 0; i < 3; i++) {
            lo[i] = pad_lo(b.lo[i], abs_pad);
            hi[i] = pad_hi(b.hi[i], abs_pad);
        }
        crt_box_to_center_half(lo, hi, v);
    }
    for (int i = 0; i < 6; i++)
        w[slot * 6 + i] = v[i];
}

} // namespace

void crt_pack_prim(const crt_hittable& h, float pr[4], float* k)
{
    const float hw = h.a / 2, hh = h.b / 2; // the reference computes center -/+ (width / 2) (Hittable.cuh:142-146)
    *k = 0.f;
    switch (h.type) {
    case CRT_SPHERE: pr[0] = h.cx; pr[1] = h.cy; pr[2] = h.cz; pr[3] = h.a; break;
    case CRT_XYRECT: pr[0] = h.cx - hw; pr[1] = h.cx + hw; pr[2] = h.cy - hh; pr[3] = h.cy + hh; *k = h.cz; break;
    case CRT_XZRECT: pr[0] = h.cx - hw; pr[1] = h.cx + hw; pr[2] = h.cz - hh; pr[3] = h.cz + hh; *k = h.cy; break;
    default:         pr[0] = h.cy - hh; pr[1] = h.cy + hh; pr[2] = h.cz - hw; pr[3] = h.cz + hw; *k = h.cx; break;
    }
}

void crt_pack_material(const crt_scene_desc* sd, const crt_material& m, float mr[8])
{
    for (int i = 0; i < 8; i++) mr[i] = 0.f;
    uint32_t bits = (uint32_t)m.type;
    float param = m.param;
    if (m.type == CRT_METAL)
        param = m.param < 1 ? m.param : 1; // Material.cuh:71
    if (m.type == CRT_DIFFUSELIGHT)
        param = (float)m.light_intensity;  // int * Vec3 in the reference (Material.cuh:168-172)
    if (m.type != CRT_DIELECTRIC) {
        const crt_texture& t = sd->textures[m.texture];
        bits |= (uint32_t)t.type << 4;
        if (t.type == CRT_CONSTANT) {
            mr[0] = t.color[0]; mr[1] = t.color[1]; mr[2] = t.color[2];
        }
        else if (t.type == CRT_CHECKER) {
            mr[0] = t.odd[0]; mr[1] = t.odd[1]; mr[2] = t.odd[2];
            mr[4] = t.even[0]; mr[5] = t.even[1]; mr[6] = t.even[2];
        }
        else {
            bool has = t.image >= 0 && sd->images[t.image].pixels != nullptr;
            bits |= (uint32_t)(has ? t.image + 1 : 0) << 8;
        }
    }
    else {
        mr[0] = mr[1] = mr[2] = 1.0f; // attenuation (1,1,1), Material.cuh:112
    }
    mr[3] = param;
    memcpy(&mr[7], &bits, 4);
}

int crt_peel_top(const CrtPackedNode* nodes, size_t n_nodes, uint64_t n_prims, int max_top, int32_t* entry, int32_t* codes)
{
    *entry = 0;
    if (n_prims < 3 || !n_nodes) return 0; // one / two primitives: nothing below a peeled leaf
    int n_top = 0;
    size_t node = 0;
    while (n_top < max_top && node < n_nodes) {
        const CrtPackedNode& nd = nodes[node];
        const bool leaf0 = nd.child[0] < 0, leaf1 = nd.child[1] < 0;
        if (leaf0 == leaf1) break; // two inner children: the tree proper starts here; two leaves: nothing below
        const int lc = leaf0 ? 0 : 1;
        // boxes are {c.xyz, h.xyz}: the node's own box is the union of its two child boxes
        const float* a = nd.w + 6 * lc;
        const float* b = nd.w + 6 * (1 - lc);
        float ext[3];
        for (int i = 0; i < 3; i++)
            ext[i] = 0.5f * (std::max(a[i] + a[3 + i], b[i] + b[3 + i]) - std::min(a[i] - a[3 + i], b[i] - b[3 + i]));
        const float area_leaf = a[3] * a[4] + a[4] * a[5] + a[5] * a[3], area_node = ext[0] * ext[1] + ext[1] * ext[2] + ext[2] * ext[0];
        if (!(area_leaf >= 0.25f * area_node)) break; // a small leaf next to the root: its box test culls it, keep it in the tree
        codes[n_top++] = ~nd.child[lc];
        node = (size_t)nd.child[1 - lc];
        *entry = (int32_t)node;
    }
    if (node >= n_nodes) { *entry = 0; return 0; } // chain longer than the prefix we were given: do not peel
    return n_top;
}

// Validation of one material and of the texture / image it references. Shared by the build paths (through check_hittable)
// and by the refit path of crt_scene_commit, which packs the material records of edited hittables without a rebuild.
bool crt_check_material(const crt_scene_desc* sd, int64_t index, std::string* err)
{
    if (index < 0 || (uint64_t)index >= sd->n_materials) {
        *err = "scene: material index " + std::to_string(index) + " out of range";
        return false;
    }
    const crt_material& m = sd->materials[index];
    if (m.type < CRT_LAMBERTIAN || m.type > CRT_DIFFUSELIGHT) {
        *err = "scene: material " + std::to_string(index) + " has unknown type";
        return false;
    }
    if (m.type != CRT_DIELECTRIC) {
        if (m.texture < 0 || (uint64_t)m.texture >= sd->n_textures) {
            *err = "scene: material " + std::to_string(index) + " has texture index out of range";
            return false;
        }
        const crt_texture& t = sd->textures[m.texture];
        if (t.type < CRT_CONSTANT || t.type > CRT_IMAGE) {
            *err = "scene: texture " + std::to_string(m.texture) + " has unknown type";
            return false;
        }
        if (t.type == CRT_IMAGE && t.image >= 0 && (uint64_t)t.image >= sd->n_images) {
            *err = "scene: texture " + std::to_string(m.texture) + " has image index out of range";
            return false;
        }
    }
    return true;
}

// Validation of one ACTIVE hittable and of everything it references (shared by the host and the GPU build paths).
static bool check_hittable(const crt_scene_desc* sd, uint64_t i, std::string* err)
{
    const crt_hittable& h = sd->hittables[i];
    if (h.type < CRT_SPHERE || h.type > CRT_YZRECT) {
        *err = "scene: hittable " + std::to_string(i) + " has unknown type " + std::to_string(h.type);
        return false;
    }
    if (h.material < 0 || (uint64_t)h.material >= sd->n_materials) {
        *err = "scene: hittable " + std::to_string(i) + " has material index out of range";
        return false;
    }
    // A NaN or infinite coordinate is never hit by the reference (every comparison of Sphere::Hit fails), but inside a BVH it
    // poisons the boxes of all its ancestors: refuse it instead of rendering a silently broken tree.
    if (!std::isfinite(h.cx) || !std::isfinite(h.cy) || !std::isfinite(h.cz) || !std::isfinite(h.a) || (h.type != CRT_SPHERE && !std::isfinite(h.b))) {
        *err = "scene: hittable " + std::to_string(i) + " has a non-finite coordinate or size (deactivate it instead)";
        return false;
    }
    if (!crt_check_material(sd, h.material, err)) return false;
    return true;
}

// Host-side preparation of a GPU build (lbvh.cuh): validates, lists the active hittables and measures the scene (bounds of
// the box centres for the Morton grid, largest coordinate magnitude for the absolute box padding). Threaded for large scenes.
bool crt_prepare_gpu_build(const crt_scene_desc* sd, std::vector<uint32_t>* active, float cmin[3], float cmax[3], float blo[3],
                           float bhi[3], float* abs_pad, std::string* err)
{
    if (!sd || (!sd->hittables && sd->n_hittables)) {
        *err = "scene: NULL hittable list";
        return false;
    }
    if (sd->n_hittables > (uint64_t)CRT_LEAF_INDEX_MASK) {
        *err = "scene: too many hittables (limit 2^26 - 1)";
        return false;
    }
    const uint64_t n = sd->n_hittables;
    unsigned nt = n >= 65536 ? std::min(16u, std::max(1u, std::thread::hardware_concurrency())) : 1u;
    struct Part { std::vector<uint32_t> act; float cmin[3], cmax[3], blo[3], bhi[3], extent; bool ok; std::string err; };
    std::vector<Part> parts(nt);
    auto work = [&](unsigned t) {
        Part& P = parts[t];
        P.ok = true;
        P.extent = 0.f;
        for (int a = 0; a < 3; a++) {
            P.cmin[a] = P.blo[a] = std::numeric_limits<float>::infinity();
            P.cmax[a] = P.bhi[a] = -std::numeric_limits<float>::infinity();
        }
        const uint64_t i0 = n * t / nt, i1 = n * (t + 1) / nt;
        P.act.reserve((size_t)(i1 - i0));
        for (uint64_t i = i0; i < i1; i++) {
            const crt_hittable& h = sd->hittables[i];
            if (!h.active) continue;
            if (!check_hittable(sd, i, &P.err)) { P.ok = false; return; }
            P.act.push_back((uint32_t)i);
            float lo[3], hi[3];
            const float hw = h.a / 2, hh = h.b / 2, r = fabsf(h.a);
            switch (h.type) {
            case CRT_SPHERE: lo[0] = h.cx - r; hi[0] = h.cx + r; lo[1] = h.cy - r; hi[1] = h.cy + r; lo[2] = h.cz - r; hi[2] = h.cz + r; break;
            case CRT_XYRECT: lo[0] = h.cx - hw; hi[0] = h.cx + hw; lo[1] = h.cy - hh; hi[1] = h.cy + hh; lo[2] = h.cz - 0.0001f; hi[2] = h.cz + 0.0001f; break;
            case CRT_XZRECT: lo[0] = h.cx - hw; hi[0] = h.cx + hw; lo[2] = h.cz - hh; hi[2] = h.cz + hh; lo[1] = h.cy - 0.0001f; hi[1] = h.cy + 0.0001f; break;
            default:         lo[1] = h.cy - hh; hi[1] = h.cy + hh; lo[2] = h.cz - hw; hi[2] = h.cz + hw; lo[0] = h.cx - 0.0001f; hi[0] = h.cx + 0.0001f; break;
            }
            for (int a = 0; a < 3; a++) {
                const float c = 0.5f * lo[a] + 0.5f * hi[a];
                P.cmin[a] = std::min(P.cmin[a], c);
                P.cmax[a] = std::max(P.cmax[a], c);
                P.blo[a] = std::min(P.blo[a], lo[a]);
                P.bhi[a] = std::max(P.bhi[a], hi[a]);
                P.extent = std::max(P.extent, std::max(fabsf(lo[a]), fabsf(hi[a])));
            }
        }
    };
    if (nt == 1) work(0);
    else {
        std::vector<std::thread> pool;
        for (unsigned t = 1; t < nt; t++) pool.emplace_back(work, t);
        work(0);
        for (std::thread& t : pool) t.join();
    }
    float extent = 0.f;
    for (int a = 0; a < 3; a++) {
        cmin[a] = blo[a] = std::numeric_limits<float>::infinity();
        cmax[a] = bhi[a] = -std::numeric_limits<float>::infinity();
    }
    size_t total = 0;
    for (const Part& P : parts) {
        if (!P.ok) { *err = P.err; return false; }
        total += P.act.size();
    }
    active->clear();
    active->reserve(total);
    for (const Part& P : parts) {
        active->insert(active->end(), P.act.begin(), P.act.end());
        extent = std::max(extent, P.extent);
        for (int a = 0; a < 3; a++) {
            cmin[a] = std::min(cmin[a], P.cmin[a]); cmax[a] = std::max(cmax[a], P.cmax[a]);
            blo[a] = std::min(blo[a], P.blo[a]); bhi[a] = std::max(bhi[a], P.bhi[a]);
        }
    }
    *abs_pad = 1e-7f + 4e-7f * extent; // same rule as the host build below
    return true;
}

bool crt_pack_scene(const crt_scene_desc* sd, CrtPackedScene* out, std::string* err)
{
    auto t0 = std::chrono::steady_clock::now();
    if (!sd || (!sd->hittables && sd->n_hittables)) {
        *err = "scene: NULL hittable list";
        return false;
    }
    if (sd->n_hittables > (uint64_t)CRT_LEAF_INDEX_MASK) {
        *err = "scene: too many hittables (limit 2^26 - 1)";
        return false;
    }
    // 1. active filter + validation
    std::vector<uint32_t> active;
    active.reserve(sd->n_hittables);
    for (uint64_t i = 0; i < sd->n_hittables; i++) {
        if (!sd->hittables[i].active)
            continue;
        if (!check_hittable(sd, i, err))
            return false;
        active.push_back((uint32_t)i);
    }
    const uint32_t n = (uint32_t)active.size();

    // 2. primitive boxes
    Builder B;
    B.pbox.resize(n);
    B.cent.resize(3 * (size_t)n);
    B.idx.resize(n);
    Box all;
    all.reset();
    for (uint32_t k = 0; k < n; k++) {
        const crt_hittable& h = sd->hittables[active[k]];
        Box& b = B.pbox[k];
        float hw = h.a / 2, hh = h.b / 2; // the reference computes center -/+ (width / 2) (Hittable.cuh:142-146)
        switch (h.type) {
        case CRT_SPHERE:
            b.lo[0] = h.cx - h.a; b.lo[1] = h.cy - h.a; b.lo[2] = h.cz - h.a;
            b.hi[0] = h.cx + h.a; b.hi[1] = h.cy + h.a; b.hi[2] = h.cz + h.a;
            if (h.a < 0.f) // negative radius (hollow-glass trick): box must still be ordered
                for (int i = 0; i < 3; i++)
                    std::swap(b.lo[i], b.hi[i]);
            break;
        case CRT_XYRECT:
            b.lo[0] = h.cx - hw; b.hi[0] = h.cx + hw; b.lo[1] = h.cy - hh; b.hi[1] = h.cy + hh;
            b.lo[2] = h.cz - 0.0001f; b.hi[2] = h.cz + 0.0001f;
            break;
        case CRT_XZRECT:
            b.lo[0] = h.cx - hw; b.hi[0] = h.cx + hw; b.lo[2] = h.cz - hh; b.hi[2] = h.cz + hh;
            b.lo[1] = h.cy - 0.0001f; b.hi[1] = h.cy + 0.0001f;
            break;
        default: // CRT_YZRECT: height spans y and width spans z (Hittable.cuh:253-257)
            b.lo[1] = h.cy - hh; b.hi[1] = h.cy + hh; b.lo[2] = h.cz - hw; b.hi[2] = h.cz + hw;
            b.lo[0] = h.cx - 0.0001f; b.hi[0] = h.cx + 0.0001f;
            break;
        }
        for (int i = 0; i < 3; i++)
            B.cent[3 * (size_t)k + i] = 0.5f * (b.lo[i] + b.hi[i]);
        B.idx[k] = k;
        all.grow(b);
    }

    auto t_boxes = std::chrono::steady_clock::now();
    // 3. build
    if (const char* e = getenv("CRT_BVH_PAR_DEPTH")) B.par_depth = atoi(e);
    if (n >= 2) {
        B.nodes.resize(n - 1);
        B.build(0, n, all, 1, 0); // the root is always build-node 0
    }
    else {
        B.nodes.emplace_back();
        B.nodes[0].box = all;
        B.nodes[0].child[0] = ~0;
        B.nodes[0].child[1] = ~0; // single object: left == right, like the reference (Hittable.cuh:326-328)
    }

    // EXPERIMENT (CRT_BVH_ROTATE=<passes>, off by default): tree rotations after the top-down build - a node's child is swapped
    // with a grandchild on the other side when that shrinks the grandchild's parent box. Counted on the real ray distribution
    // with tests/sim/path_sim (box / primitive tests per ray): RTIOW 18.35 -> 18.18 / 1.98, 10 k glass-metal and the sphere
    // grid unchanged (< 0.1 %), textured (67 primitives) 5.66 -> 4.04 / 1.72 -> 2.21 (the earth sphere moves to the top and is
    // peeled). Not adopted until the B200 has timed it: fewer node visits against more primitive tests.
    if (n >= 3 && getenv("CRT_BVH_ROTATE") && atoi(getenv("CRT_BVH_ROTATE")) > 0) {
        auto child_box = [&](int32_t c) -> const Box& { return c >= 0 ? B.nodes[c].box : B.pbox[(uint32_t)~c]; };
        // post-order: children before their parent
        std::vector<int32_t> post;
        {
            std::vector<int32_t> st{0};
            while (!st.empty()) {
                const int32_t v = st.back();
                st.pop_back();
                post.push_back(v);
                for (int c = 0; c < 2; c++)
                    if (B.nodes[v].child[c] >= 0) st.push_back(B.nodes[v].child[c]);
            }
            std::reverse(post.begin(), post.end());
        }
        const int passes = std::max(1, atoi(getenv("CRT_BVH_ROTATE")));
        long applied = 0;
        for (int pass = 0; pass < passes; pass++) {
            long changed = 0;
            for (int32_t v : post) {
                BuildNode& N = B.nodes[v];
                float best = -1e-7f * N.box.half_area();
                int best_s = -1, best_g = -1;
                for (int sd_ = 0; sd_ < 2; sd_++) {
                    const int32_t xc = N.child[sd_];
                    if (xc < 0) continue;
                    const BuildNode& X = B.nodes[xc];
                    const Box& Y = child_box(N.child[1 - sd_]);
                    for (int g = 0; g < 2; g++) { // swap Y with X.child[g]: X keeps child[1 - g] and gains Y
                        Box nb = child_box(X.child[1 - g]);
                        nb.grow(Y);
                        const float delta = nb.half_area() - X.box.half_area();
                        if (delta < best) { best = delta; best_s = sd_; best_g = g; }
                    }
                }
                if (best_s >= 0) {
                    BuildNode& X = B.nodes[N.child[best_s]];
                    std::swap(N.child[1 - best_s], X.child[best_g]);
                    Box nb = child_box(X.child[0]);
                    nb.grow(child_box(X.child[1]));
                    X.box = nb;
                    changed++;
                }
            }
            applied += changed;
            if (!changed) break;
        }
        // depth after the rotations
        int depth = 1;
        std::vector<std::pair<int32_t, int>> st{{0, 1}};
        while (!st.empty()) {
            auto [v, d] = st.back();
            st.pop_back();
            depth = std::max(depth, d + 1);
            for (int c = 0; c < 2; c++)
                if (B.nodes[v].child[c] >= 0) st.push_back({B.nodes[v].child[c], d + 1});
        }
        B.max_depth.store(depth);
        if (getenv("CRT_BVH_TIMING")) fprintf(stderr, "[crt] bvh: %ld rotations, depth %d\n", applied, depth);
    }

    auto t_built = std::chrono::steady_clock::now();
    // 4. breadth-first emission; leaves get their final primitive slot in BFS discovery order so that
    //    primitives referenced by neighbouring nodes are neighbours in memory too.
    const uint32_t nn = (uint32_t)B.nodes.size();
    out->nodes.assign(nn, CrtPackedNode());
    const size_t np_alloc = n ? n : 1;
    out->prims.assign(4 * np_alloc, 0.f);
    out->prim_k.assign(np_alloc, 0.f);
    out->mats.assign(8 * np_alloc, 0.f);
    out->prim_id.assign(np_alloc, -1);
    out->parent.assign(nn, -1);
    out->leaf_ref.assign(np_alloc, 0);
    std::vector<int32_t> order;   // BFS order -> build node
    std::vector<int32_t> newidx(nn, -1);
    order.reserve(nn);
    order.push_back(0);
    newidx[0] = 0;
    for (size_t q = 0; q < order.size(); q++) {
        const BuildNode& bn = B.nodes[order[q]];
        for (int c = 0; c < 2; c++) {
            if (n >= 2 && bn.child[c] >= 0) {
                newidx[bn.child[c]] = (int32_t)order.size();
                order.push_back(bn.child[c]);
            }
        }
    }
    // first leaf slot of every packed node (leaves get their slots in BFS order), so that the packing loop below has no
    // loop-carried state and can be split over threads
    std::vector<uint32_t> slot0(nn + 1, 0);
    for (uint32_t q = 0; q < nn; q++) {
        const BuildNode& bn = B.nodes[order[q]];
        uint32_t leaves = 0;
        if (n >= 2) leaves = (bn.child[0] < 0) + (bn.child[1] < 0);
        else if (n == 1) leaves = 1;
        slot0[q + 1] = slot0[q] + leaves;
    }
    const float root_area = all.half_area();
    float extent = 0.f;
    if (n)
        for (int i = 0; i < 3; i++) extent = std::max(extent, std::max(fabsf(all.lo[i]), fabsf(all.hi[i])));
    const float abs_pad = 1e-7f + 4e-7f * extent;
    out->abs_pad = abs_pad;
    const uint32_t n_chunks = nn >= 65536 ? 16u : 1u;
    std::vector<double> sah_part(n_chunks, 0.0);
    auto pack_range = [&](uint32_t chunk) {
    double sah = 0.0;
    const uint32_t q_lo = (uint32_t)((uint64_t)nn * chunk / n_chunks), q_hi = (uint32_t)((uint64_t)nn * (chunk + 1) / n_chunks);
    for (uint32_t q = q_lo; q < q_hi; q++) {
        const BuildNode& bn = B.nodes[order[q]];
        CrtPackedNode& pn = out->nodes[q];
        uint32_t next_prim = slot0[q];
        memset(&pn, 0, sizeof(pn));
        if (root_area > 0.f)
            sah += 1.0 * bn.box.half_area() / root_area;
        for (int c = 0; c < 2; c++) {
            if (n == 0 && c == 0) {
                // Empty scene: one dummy sphere with a NaN centre (its discriminant is NaN, so it can never be hit)
                // behind an always-true box, so that the kernels need no special case.
                store_box(pn.w, c, all, true, abs_pad);
                pn.child[c] = ~(int32_t)(((uint32_t)CRT_SPHERE << CRT_LEAF_TYPE_SHIFT) | 0u);
                out->prims[0] = out->prims[1] = out->prims[2] = std::numeric_limits<float>::quiet_NaN();
                out->prims[3] = 0.f;
                out->prim_id[0] = -1;
                continue;
            }
            if (n < 2 && c == 1) { // second slot repeats the first (its test fails the strict t < t_max)
                for (int i = 0; i < 6; i++) pn.w[6 + i] = pn.w[i];
                pn.child[1] = pn.child[0];
                continue;
            }
            int32_t ch = bn.child[c];
            if (ch >= 0) {
                store_box(pn.w, c, B.nodes[ch].box, false, abs_pad);
                pn.child[c] = newidx[ch];
                out->parent[newidx[ch]] = (int32_t)(q * 2 + c);
            }
            else {
                uint32_t k = (uint32_t)~ch;          // index into the active list
                uint32_t slot = next_prim++;
                const crt_hittable& h = sd->hittables[active[k]];
                store_box(pn.w, c, B.pbox[k], false, abs_pad);
                if (root_area > 0.f)
                    sah += 1.0 * B.pbox[k].half_area() / root_area;
                pn.child[c] = ~(int32_t)(((uint32_t)h.type << CRT_LEAF_TYPE_SHIFT) |
                                         ((uint32_t)sd->materials[h.material].type << CRT_LEAF_MAT_SHIFT) | slot);
                crt_pack_prim(h, &out->prims[4 * (size_t)slot], &out->prim_k[slot]);
                out->prim_id[slot] = (int32_t)active[k];
                out->leaf_ref[slot] = (int32_t)(q * 2 + c);
                crt_pack_material(sd, sd->materials[h.material], &out->mats[8 * (size_t)slot]);
            }
        }
    }
    sah_part[chunk] = sah;
    };
    if (n_chunks == 1) {
        pack_range(0);
    }
    else {
        std::vector<std::thread> pool;
        for (uint32_t c = 1; c < n_chunks; c++) pool.emplace_back(pack_range, c);
        pack_range(0);
        for (std::thread& t : pool) t.join();
    }
    double sah = 0.0;
    for (double v : sah_part) sah += v; // fixed chunking, fixed order: the statistic is deterministic

    auto t1 = std::chrono::steady_clock::now();
    if (getenv("CRT_BVH_TIMING"))
        fprintf(stderr, "[crt] bvh: boxes %.1f ms, build %.1f ms, emit %.1f ms\n", std::chrono::duration<double, std::milli>(t_boxes - t0).count(),
                std::chrono::duration<double, std::milli>(t_built - t_boxes).count(), std::chrono::duration<double, std::milli>(t1 - t_built).count());
    memset(&out->stats, 0, sizeof(out->stats));
    out->stats.n_prims = n;
    out->stats.n_nodes = nn;
    out->stats.max_depth = (uint64_t)B.max_depth.load();
    out->stats.sah_cost = sah;
    out->stats.build_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    out->stats.node_bytes = (uint64_t)nn * sizeof(CrtPackedNode);
    out->stats.prim_bytes = (uint64_t)n * (16 + 4 + 32 + 4);
    return true;
}
