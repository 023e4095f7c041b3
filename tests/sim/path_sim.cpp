// path_sim.cpp — TEST INFRASTRUCTURE: the device code of a whole camera path (cudaraytracer_b200/csrc/path.cuh: gen_primary,
// closest_hit, scatter_step -> shade.cuh, trace.cuh, Philox) compiled for the HOST with g++ and run as the megakernel runs it
// (one pixel after the other, all its samples, flattened bounce loop). No GPU. Used by tests/test_path_sim.py to
//   * compare the image of the DEVICE shading code with the oracle's on the CPU (statistically: the RNG streams differ), and
//   * report what the tree costs a real ray distribution: rays per path, box tests and primitive tests per ray - the same
//     counters crt_get_counters returns for a CRT_FLAG_COUNT_TESTS render on the GPU (bench.py: roofline.box_tests_per_ray).
// The approximate reciprocal / square root / sine of the device are IEEE / libm here, so single paths differ in the last bits;
// nothing below depends on that.
//
//   path_sim <scene kind> <size hint> <width> <height> <spp> <max depth> [out.f32: width*height*3 mean radiance] [threads]
//            [events file: the node / leaf visit sequence of every ray, for tools/sched_model.py | ""]
//            [ids file: primary-hit hittable index per pixel with the EXACT arithmetic (what crt_primary_ids computes)]
#define CRT_HOST_SIM 1
#include <cuda_runtime.h> // vector types only

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#undef __device__
#undef __host__
#undef __forceinline__
#undef __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __global__

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
// glibc's <math.h> declares __sinf / __log2f / __sincosf itself (internal aliases): map the device names with macros
static inline void crt_sim_sincosf(float x, float* s, float* c) { *s = sinf(x); *c = cosf(x); }
#define __sinf(x) sinf(x)
#define __log2f(x) log2f(x)
#define __sincosf(x, s, c) crt_sim_sincosf((x), (s), (c))
static inline float rsqrtf(float x) { return 1.0f / sqrtf(x); }
using std::max;
using std::min;
static struct { unsigned x, y, z; } threadIdx = {0, 0, 0}, blockDim = {1, 1, 1};

#include "crt_internal.h"
#include "path.cuh"

int main(int argc, char** argv)
{
    const int kind = argc > 1 ? atoi(argv[1]) : CRT_SCENE_RTIOW;
    const uint64_t hint = argc > 2 ? (uint64_t)atoll(argv[2]) : 0;
    const uint32_t W = argc > 3 ? (uint32_t)atoi(argv[3]) : 128, H = argc > 4 ? (uint32_t)atoi(argv[4]) : 72;
    const uint32_t spp = argc > 5 ? (uint32_t)atoi(argv[5]) : 8, depth = argc > 6 ? (uint32_t)atoi(argv[6]) : 50;
    const char* out = argc > 7 ? argv[7] : "";
    unsigned nt = argc > 8 ? (unsigned)atoi(argv[8]) : std::max(1u, std::thread::hardware_concurrency());
    // optional: the visit sequence of every ray ('N' / 'L' bytes, 0-terminated) for tools/sched_model.py; forces one thread
    FILE* ev = argc > 9 && argv[9][0] ? fopen(argv[9], "wb") : nullptr;
    if (ev) nt = 1;

    crt_scene_builder* sb = nullptr;
    if (crt_scene_generate(kind, 1984, hint, &sb) != CRT_OK) { fprintf(stderr, "scene: %s\n", crt_last_error_string(nullptr)); return 2; }
    const crt_scene_desc* sd = crt_scene_builder_desc(sb);
    CrtPackedScene P;
    std::string err;
    if (!crt_pack_scene(sd, &P, &err)) { fprintf(stderr, "pack: %s\n", err.c_str()); return 2; }
    const int nn = (int)P.nodes.size(), np = (int)P.stats.n_prims;

    // image textures: RGB8 -> RGBA8, what k_rgb_to_rgba does on the device
    std::vector<std::vector<uchar4>> texels(sd->n_images);
    std::vector<DevImage> images(std::max<size_t>(sd->n_images, 1));
    for (uint64_t i = 0; i < sd->n_images; i++) {
        const crt_image& im = sd->images[i];
        images[i].texels = nullptr; images[i].width = im.width; images[i].height = im.height;
        if (!im.pixels || im.width <= 0 || im.height <= 0) continue;
        const size_t n = (size_t)im.width * im.height;
        texels[i].resize(n);
        for (size_t k = 0; k < n; k++) texels[i][k] = make_uchar4(im.pixels[3 * k], im.pixels[3 * k + 1], im.pixels[3 * k + 2], 255);
        images[i].texels = texels[i].data();
    }

    RenderArgs A;
    memset(&A, 0, sizeof(A));
    DevScene& D = A.scene;
    D.nodes = reinterpret_cast<const float4*>(P.nodes.data());
    D.prims = reinterpret_cast<const float4*>(P.prims.data());
    D.prim_k = P.prim_k.data();
    D.mats = reinterpret_cast<const float4*>(P.mats.data());
    D.prim_id = P.prim_id.data();
    D.images = images.data();
    D.n_nodes = nn; D.n_prims = np;
    D.smem_nodes = nn; D.smem_prims = np; D.smem_mats = np;
    int32_t entry = 0, codes[CRT_MAX_TOP] = {0, 0, 0, 0};
    D.n_top = crt_peel_top(P.nodes.data(), P.nodes.size(), (uint64_t)np, CRT_MAX_TOP, &entry, codes);
    D.entry = entry;
    for (int k = 0; k < CRT_MAX_TOP; k++) D.top_code[k] = k < D.n_top ? codes[k] : 0;
    std::vector<float4> smem((size_t)nn * 4 + (size_t)np * 3 + 16);
    const SceneView S = stage_scene<false>(D, smem.data()); // the megakernel's word-major tables

    // camera: crt_set_camera's derivation (crt_api.cu derive_camera; Kernel.cu:130-146, CudaLayer.cpp:45-62)
    crt_input_struct in;
    if (crt_scene_builder_camera(sb, W, H, &in) != CRT_OK) { fprintf(stderr, "camera: %s\n", crt_last_error_string(nullptr)); return 2; }
    DevCamera& C = A.cam;
    C.ox = in.origin_x; C.oy = in.origin_y; C.oz = in.origin_z;
    C.fx = in.orientation_x; C.fy = in.orientation_y; C.fz = in.orientation_z;
    C.ux = in.up_x; C.uy = in.up_y; C.uz = in.up_z;
    {
        volatile float cx = C.uy * C.fz - C.uz * C.fy, cy = -(C.ux * C.fz - C.uz * C.fx), cz = C.ux * C.fy - C.uy * C.fx;
        volatile float dd = cx * cx + cy * cy + cz * cz;
        volatile float inv = 1 / sqrtf(dd);
        C.rx = inv * cx; C.ry = inv * cy; C.rz = inv * cz;
    }
    C.far_plane = in.far_plane; C.near_plane = in.near_plane; C.fov = in.fov;
    { volatile float k = 1.0f / in.fov; C.far_k = k * 10.0f; }
    for (int i = 0; i < 3; i++) { C.bg0[i] = in.background_start[i]; C.bg1[i] = in.background_end[i]; }
    C.lens_radius = 0.f; C.focus_dist = 1.f;
    A.width = W; A.height = H; A.cover_w = W; A.cover_h = H;
    A.spp_begin = 0; A.spp_count = spp; A.spp_stride = 1; A.max_depth = depth;
    A.seed_lo = 1984u; A.seed_hi = 0u;
    crt_philox_round_keys(A.seed_lo, A.seed_hi, A.philox_keys);
    A.jitter = CRT_JITTER_RANDOM;

    std::vector<float> mean((size_t)W * H * 3, 0.f);
    std::atomic<unsigned long long> t_paths{0}, t_rays{0}, t_box{0}, t_prim{0};
    std::atomic<uint32_t> next_row{0};
    auto work = [&] {
        unsigned long long paths = 0, rays = 0, box = 0, prim = 0;
        for (uint32_t y; (y = next_row.fetch_add(1)) < H;) {
            for (uint32_t x = 0; x < W; x++) {
                const uint32_t pixel = y * W + x;
                double acc[3] = {0, 0, 0};
                for (uint32_t s = 0; s < spp; s++) {
                    float3 o, d, thr = f3(1.f, 1.f, 1.f), add;
                    uint32_t bounce = 0;
                    gen_primary<false>(A, (int)x, (int)y, pixel, s, o, d);
                    paths++;
                    while (true) {
                        unsigned nb = 0, npr = 0;
                        rays++;
                        Hit h;
                        if (!ev) h = closest_hit<false, true, true>(S, o, d, 0.001f, 3.402823466e+38f, nb, npr);
                        else { // the same walk, written out: one byte per visit ('N' inner node, 'L' leaf), 0 ends the ray
                            h.t = 3.402823466e+38f;
                            h.leaf = -1;
                            const TravRay r = trav_setup<false>(o, d);
                            int stack[CRT_STACK_SIZE], sp = 0;
                            int cur = trav_begin<false, true>(S, r, 0.001f, h);
                            npr += (unsigned)S.n_top;
                            while (cur != CRT_TRAV_DONE) {
                                if (cur >= 0) { node_step<false, true>(S, r, 0.001f, h.t, cur, sp, stack); nb += 2; fputc('N', ev); }
                                else { leaf_step<false, true>(S, r, 0.001f, h, cur, sp, stack); npr++; fputc('L', ev); }
                            }
                            fputc(0, ev);
                        }
                        box += nb; prim += npr;
                        if (!scatter_step<true>(A, S, h, pixel, s, bounce, o, d, thr, add)) break;
                    }
                    // what crt_to_fixed does to a sample before it is added: NaN -> 0, clamp
                    const float v[3] = {add.x, add.y, add.z};
                    for (int c = 0; c < 3; c++) acc[c] += (v[c] == v[c]) ? std::min(std::max(v[c], -CRT_FIXED_MAX), CRT_FIXED_MAX) : 0.f;
                }
                for (int c = 0; c < 3; c++) mean[3 * (size_t)pixel + c] = (float)(acc[c] / spp);
            }
        }
        t_paths += paths; t_rays += rays; t_box += box; t_prim += prim;
    };
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < nt; t++) pool.emplace_back(work);
    work();
    for (std::thread& t : pool) t.join();

    const double rays = (double)t_rays.load();
    printf("{\"scene\": %d, \"prims\": %d, \"nodes\": %d, \"peeled\": %d, \"width\": %u, \"height\": %u, \"spp\": %u, \"paths\": %llu, \"rays\": %llu, "
           "\"rays_per_path\": %.6f, \"box_tests_per_ray\": %.6f, \"prim_tests_per_ray\": %.6f}\n",
           kind, np, nn, D.n_top, W, H, spp, t_paths.load(), t_rays.load(), rays / (double)t_paths.load(), (double)t_box.load() / rays,
           (double)t_prim.load() / rays);
    if (ev) fclose(ev);
    if (argc > 10) { // k_primary_ids on the host: exact arithmetic, jitter (0.5, 0.5) -> hittable index per pixel (int32)
        std::vector<int32_t> ids((size_t)W * H);
        for (uint32_t y = 0; y < H; y++)
            for (uint32_t x = 0; x < W; x++) {
                float3 o, d;
                camera_ray<true>(A.cam, W, H, (int)x, (int)y, 0.5f, 0.5f, o, d);
                unsigned nb = 0, npr = 0;
                const Hit h = closest_hit<true, false, true>(S, o, d, 0.001f, 3.402823466e+38f, nb, npr);
                ids[(size_t)y * W + x] = h.leaf < 0 ? -1 : D.prim_id[h.leaf & CRT_LEAF_INDEX_MASK];
            }
        FILE* f = fopen(argv[10], "wb");
        if (!f || fwrite(ids.data(), 4, ids.size(), f) != ids.size()) { fprintf(stderr, "cannot write %s\n", argv[10]); return 2; }
        fclose(f);
    }
    if (out[0]) {
        FILE* f = fopen(out, "wb");
        if (!f || fwrite(mean.data(), sizeof(float), mean.size(), f) != mean.size()) { fprintf(stderr, "cannot write %s\n", out); return 2; }
        fclose(f);
    }
    crt_scene_builder_free(sb);
    return 0;
}
