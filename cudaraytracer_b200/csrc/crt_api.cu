// crt_api.cu — the C ABI of include/crt.h: context, scene upload, camera, framebuffer, render, resolve.
// Everything here runs on the GPU or fails with an error code; there is no CPU rendering path.

#include "crt.h"
#include "crt_internal.h"

#include "kernels.cuh"
#include "pipeline.cuh"
#include "refit.cuh"
#include "lbvh.cuh"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>

// NVTX v3 is header-only (the injection library is looked up lazily; without a profiler attached a range costs a load and a
// branch). Ranges around the phases of a frame: upload / commit / render / close-out / resolve / read-back (SURVEY.md §5 tracing).
#include <nvtx3/nvToolsExt.h>
namespace {
struct NvtxRange
{
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};
} // namespace
#include <new>
#include <string>
#include <vector>

struct crt_context
{
    int device = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool timed = false;
    int sm_count = 0;
    int smem_optin = 0, smem_per_sm = 0;
    int smem_budget_override = -1;
    int bvh_builder = CRT_BUILDER_AUTO;      // crt_set_bvh_builder / CRT_BVH_BUILDER
    uint64_t gpu_build_min_prims = 65536;   // AUTO: scenes from this size up are built on the GPU (lbvh.cuh)
    long pipe_prefix_bytes = 24 * 1024; // scene prefix staged by the pipelined kernel when the whole scene does not fit

    // scene
    float4* d_nodes = nullptr;
    // the node table in the child-interleaved layout of the pipelined kernel's global path (trace.cuh node_visit_pairg): made on
    // the device from d_nodes when a render needs it and d_nodes has changed since (nodes_gen != pair_gen)
    float4* d_nodes_pair = nullptr;
    size_t cap_pair = 0;
    uint64_t nodes_gen = 1, pair_gen = 0;
    // CRT_VARIANT_AUTO on scenes both kernels can stage completely: the choice is MEASURED on the first samples of a frame and
    // kept while the tree, the frame size and the depth stay the same (crt_render)
    uint64_t auto_key = 0;
    int auto_variant = 0, last_variant = 0;
    int auto_thresh = 0, auto_rounds = 0;    // refill parameters of the pipelined kernel that won the measurement (0 = static)
    int cal_thresh = 0, cal_rounds = 0;      // ... the pair the NEXT render_pipeline call runs with (set and reset by crt_render)
    bool auto_calibrate = true; // CRT_AUTO_CALIBRATE=0: the static rule (pipeline from 256 primitives up)
    cudaEvent_t ev_cal[4] = {nullptr, nullptr, nullptr, nullptr};
    float4* d_prims = nullptr;
    float* d_prim_k = nullptr;
    float4* d_mats = nullptr;
    int* d_prim_id = nullptr;
    DevImage* d_images = nullptr;
    std::vector<uchar4*> d_texels;
    DevScene scene{};
    bool have_scene = false;
    // host copy of the uploaded description (edits are staged here) + what the GPU refit needs
    CrtFlatScene host;               // images[].pixels is only ever tested against NULL after the upload
    std::vector<int32_t> slot_of;    // hittable index -> primitive slot, -1 if inactive
    std::vector<uint8_t> dirty;      // per hittable
    std::vector<uint64_t> dirty_list;
    bool need_rebuild = false;
    int* d_parent = nullptr;
    int* d_leaf_ref = nullptr;
    unsigned int* d_visits = nullptr;
    int32_t top_hittable[CRT_MAX_TOP] = {-1, -1, -1, -1}; // hittable index behind each peeled leaf (their codes carry the material type)
    size_t cap_nodes = 0, cap_prims = 0; // element capacity of the geometry arrays (re-uploads of a same-sized scene reuse them)
    unsigned char* build_arena = nullptr; // scratch of the GPU builder, kept between builds (grow-only)
    size_t build_arena_bytes = 0;
    void* d_edit_recs = nullptr;     // staging for crt_scene_commit's refit path (grow-only: no allocation per commit)
    size_t edit_recs_cap = 0;
    float abs_pad = 0.f, scene_extent = 0.f;
    crt_bvh_stats stats{};
    uint64_t scene_h2d_bytes = 0;
    // Unchanged-scene cache: crt_scene_upload hashes the description it is given; when it is the scene already on the device,
    // nothing is validated, built or copied again (the reference builds its world once and only sends the camera per frame,
    // CudaLayer.cpp:372-375). crt_set_scene_cache / CRT_SCENE_CACHE=0 turn it off.
    int scene_cache = 1;
    bool scene_hash_valid = false;
    uint64_t scene_hash = 0;
    unsigned char* pinned_stage = nullptr; // page-locked staging of the raw description for the GPU builder's upload (grow-only)
    size_t pinned_stage_bytes = 0;

    // camera
    crt_input_struct inputs{};
    DevCamera cam{};
    bool have_cam = false;
    float aperture = 0.f, focus_dist = 1.f;

    // framebuffer
    uint32_t width = 0, height = 0;
    long long* d_accum_own = nullptr; // 4 x int64 per pixel {R, G, B, samples}, Q31.32 (device_common.cuh)
    long long* d_accum = nullptr;
    float4* d_accum_f = nullptr;      // staging for crt_read_accum
    uint32_t* d_rgba = nullptr;
    int* d_ids = nullptr;

    DevCounters* d_counters = nullptr; // CRT_COUNTER_SHARDS entries
    WfShard* d_shards = nullptr;       // work counters of the pipelined megakernel
    int pipe_thresh = 0, pipe_rounds = 0, // 0 = chosen per scene (render_pipeline)
         pipe_slots = 0; // CRT_PIPE_THRESH / _ROUNDS / _SLOTS
    uint64_t launches = 0;

    WavefrontState wf;
    size_t wf_slots = 1u << 20; // path slots of the wavefront pool (CRT_WF_SLOTS)
    int wf_thresh = 24;         // extend refills a warp when fewer lanes than this are traversing (CRT_WF_THRESH)
    int wf_batch_iters = 32;    // extend+shade iterations enqueued between two reads of the active count

    std::vector<void*> ipc_open; // peer mappings opened by crt_ipc_open

    std::string err;
};

static std::mutex g_err_mu;
static std::string g_err;

crt_status crt_set_error(crt_context* ctx, crt_status st, const char* msg)
{
    if (ctx) ctx->err = msg ? msg : "";
    std::lock_guard<std::mutex> lk(g_err_mu);
    g_err = msg ? msg : "";
    return st;
}

static crt_status cuda_fail(crt_context* ctx, cudaError_t e, const char* what)
{
    char buf[512];
    snprintf(buf, sizeof(buf), "%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
    return crt_set_error(ctx, CRT_ERR_CUDA, buf);
}

#define CU(call)                                                   \
    do {                                                           \
        cudaError_t e__ = (call);                                  \
        if (e__ != cudaSuccess) return cuda_fail(ctx, e__, #call); \
    } while (0)

static void free_geometry(crt_context* ctx)
{
    cudaFree(ctx->d_nodes); ctx->d_nodes = nullptr;
    cudaFree(ctx->d_nodes_pair); ctx->d_nodes_pair = nullptr; ctx->cap_pair = 0;
    cudaFree(ctx->d_prims); ctx->d_prims = nullptr;
    cudaFree(ctx->d_prim_k); ctx->d_prim_k = nullptr;
    cudaFree(ctx->d_mats); ctx->d_mats = nullptr;
    cudaFree(ctx->d_prim_id); ctx->d_prim_id = nullptr;
    cudaFree(ctx->d_parent); ctx->d_parent = nullptr;
    cudaFree(ctx->d_leaf_ref); ctx->d_leaf_ref = nullptr;
    cudaFree(ctx->d_visits); ctx->d_visits = nullptr;
    ctx->cap_nodes = ctx->cap_prims = 0;
    ctx->have_scene = false;
}

// Device arrays for a tree of nn nodes over np primitives. A scene of the same size (the usual re-upload: the same scene
// again, or after an edit) keeps the allocations: cudaMalloc / cudaFree of hundreds of MB synchronise the device and have
// been seen to stall for hundreds of ms when several processes share the box.
static crt_status ensure_geometry(crt_context* ctx, size_t nn, size_t np)
{
    nn = std::max<size_t>(nn, 1);
    np = std::max<size_t>(np, 1);
    ctx->have_scene = false;
    ctx->nodes_gen++; // whoever asks for geometry is about to write d_nodes
    if (ctx->d_nodes && nn <= ctx->cap_nodes && np <= ctx->cap_prims && nn * 2 > ctx->cap_nodes && np * 2 > ctx->cap_prims) return CRT_OK;
    free_geometry(ctx);
    CU(cudaMalloc(&ctx->d_nodes, nn * 64));
    CU(cudaMalloc(&ctx->d_parent, nn * 4));
    CU(cudaMalloc(&ctx->d_visits, nn * 4));
    CU(cudaMalloc(&ctx->d_leaf_ref, np * 4));
    CU(cudaMalloc(&ctx->d_prims, np * 16));
    CU(cudaMalloc(&ctx->d_prim_k, np * 4));
    CU(cudaMalloc(&ctx->d_mats, np * 32));
    CU(cudaMalloc(&ctx->d_prim_id, np * 4));
    ctx->cap_nodes = nn;
    ctx->cap_prims = np;
    return CRT_OK;
}

static void free_images(crt_context* ctx)
{
    cudaFree(ctx->d_images); ctx->d_images = nullptr;
    for (uchar4* p : ctx->d_texels) cudaFree(p);
    ctx->d_texels.clear();
}

static void free_scene(crt_context* ctx)
{
    free_geometry(ctx);
    free_images(ctx);
}

extern "C" {

int crt_abi_version(void)
{
    return CRT_ABI_VERSION;
}

const char* crt_last_error_string(const crt_context* ctx)
{
    if (ctx) return ctx->err.c_str();
    static thread_local std::string copy;
    std::lock_guard<std::mutex> lk(g_err_mu);
    copy = g_err;
    return copy.c_str();
}

crt_status crt_context_create(int device, crt_context** out)
{
    if (!out) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_context_create: out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        // Loud failure: this library has no CPU path.
        return crt_set_error(nullptr, CRT_ERR_CUDA, "crt_context_create: no CUDA device available (this library has no CPU fallback)");
    }
    if (device < 0 || device >= n) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_context_create: device index out of range");
    crt_context* ctx = new (std::nothrow) crt_context();
    if (!ctx) return crt_set_error(nullptr, CRT_ERR_NOMEM, "crt_context_create: out of memory");
    ctx->device = device;
    if ((e = cudaSetDevice(device)) != cudaSuccess) { crt_status s = cuda_fail(nullptr, e, "cudaSetDevice"); delete ctx; return s; }
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) { crt_status s = cuda_fail(nullptr, e, "cudaGetDeviceProperties"); delete ctx; return s; }
    if (prop.major != 10) {
        delete ctx;
        return crt_set_error(nullptr, CRT_ERR_UNSUPPORTED, "crt_context_create: kernels are built for sm_100a (B200) only");
    }
    ctx->sm_count = prop.multiProcessorCount;
    ctx->smem_optin = (int)prop.sharedMemPerBlockOptin;
    ctx->smem_per_sm = (int)prop.sharedMemPerMultiprocessor;
    const char* env = getenv("CRT_SMEM_KB");
    if (env) ctx->smem_budget_override = atoi(env) * 1024;
    if ((env = getenv("CRT_WF_SLOTS")) && atol(env) > 0) ctx->wf_slots = (size_t)atol(env);
    if ((env = getenv("CRT_WF_THRESH")) && atoi(env) > 0) ctx->wf_thresh = atoi(env);
    if ((env = getenv("CRT_WF_BATCH_ITERS")) && atoi(env) > 0) ctx->wf_batch_iters = atoi(env);
    if ((env = getenv("CRT_PIPE_THRESH")) && atoi(env) > 0) ctx->pipe_thresh = atoi(env);
    if ((env = getenv("CRT_PIPE_ROUNDS")) && atoi(env) > 0) ctx->pipe_rounds = atoi(env);
    if ((env = getenv("CRT_PIPE_SLOTS")) && atoi(env) > 0) ctx->pipe_slots = atoi(env);
    if ((env = getenv("CRT_AUTO_CALIBRATE"))) ctx->auto_calibrate = atoi(env) != 0;
    if ((env = getenv("CRT_PIPE_PREFIX_KB")) && atoi(env) >= 0) ctx->pipe_prefix_bytes = 1024L * atoi(env);
    if ((env = getenv("CRT_SCENE_CACHE"))) ctx->scene_cache = atoi(env) != 0;
    if ((env = getenv("CRT_BVH_BUILDER"))) {
        if (!strcmp(env, "host")) ctx->bvh_builder = CRT_BUILDER_HOST_SAH;
        else if (!strcmp(env, "gpu") || !strcmp(env, "lbvh")) ctx->bvh_builder = CRT_BUILDER_GPU_LBVH;
    }
    if ((env = getenv("CRT_GPU_BUILD_MIN_PRIMS")) && atol(env) >= 2) ctx->gpu_build_min_prims = (uint64_t)atol(env);
    if ((e = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaEventCreate(&ctx->ev0)) != cudaSuccess || (e = cudaEventCreate(&ctx->ev1)) != cudaSuccess ||
        (e = cudaMalloc(&ctx->d_counters, CRT_COUNTER_SHARDS * sizeof(DevCounters))) != cudaSuccess ||
        (e = cudaMemset(ctx->d_counters, 0, CRT_COUNTER_SHARDS * sizeof(DevCounters))) != cudaSuccess ||
        (e = cudaMalloc(&ctx->d_shards, WF_SHARDS * sizeof(WfShard))) != cudaSuccess) {
        crt_status s = cuda_fail(nullptr, e, "crt_context_create");
        crt_context_destroy(ctx);
        return s;
    }
    ctx->stream = ctx->own_stream;
    *out = ctx;
    return CRT_OK;
}

void crt_context_destroy(crt_context* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    free_scene(ctx);
    wavefront_free(ctx->wf);
    for (void* p : ctx->ipc_open) cudaIpcCloseMemHandle(p);
    cudaFree(ctx->d_accum_own);
    cudaFree(ctx->d_accum_f);
    cudaFree(ctx->d_rgba);
    cudaFree(ctx->d_ids);
    cudaFree(ctx->d_counters);
    cudaFree(ctx->d_shards);
    cudaFree(ctx->d_edit_recs);
    cudaFree(ctx->build_arena);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    for (int k = 0; k < 4; k++) if (ctx->ev_cal[k]) cudaEventDestroy(ctx->ev_cal[k]);
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    delete ctx;
}

crt_status crt_set_stream(crt_context* ctx, void* cuda_stream)
{
    if (!ctx) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_set_stream: ctx is NULL");
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
    return CRT_OK;
}

crt_status crt_synchronize(crt_context* ctx)
{
    if (!ctx) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_synchronize: ctx is NULL");
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    return CRT_OK;
}

} // extern "C"

// ------------------------------------------------------------------------------------------
// scene
// ------------------------------------------------------------------------------------------
// images: uploaded as the RGB8 the caller holds and expanded on the device to RGBA8, so a texel is one aligned 32-bit load.
// The staging buffers are returned in raw_rgb; the caller frees them after it has synchronised the stream.
// ---- host-side helpers of the upload path: threaded copy and a threaded 64-bit content hash (a 96 MB description of a
// million spheres takes one core 10 ms to copy; the upload is inside every end-to-end step) ----
static unsigned host_threads(size_t bytes)
{
    if (bytes < (1u << 20)) return 1;
    return std::max(1u, std::min(8u, std::thread::hardware_concurrency()));
}

static void parallel_copy(void* dst, const void* src, size_t bytes)
{
    const unsigned nt = host_threads(bytes);
    if (nt == 1) { if (bytes) memcpy(dst, src, bytes); return; }
    std::vector<std::thread> pool;
    const size_t chunk = ((bytes + nt - 1) / nt + 63) & ~(size_t)63;
    for (unsigned t = 0; t < nt; t++) {
        const size_t b = std::min(bytes, (size_t)t * chunk), e = std::min(bytes, b + chunk);
        if (e > b) pool.emplace_back([=]() { memcpy((char*)dst + b, (const char*)src + b, e - b); });
    }
    for (std::thread& t : pool) t.join();
}

static inline uint64_t mix64(uint64_t h, uint64_t w)
{
    h ^= w;
    h *= 0x9E3779B97F4A7C15ull;
    return (h << 29) | (h >> 35);
}

static uint64_t hash_range(const unsigned char* p, size_t bytes, uint64_t seed)
{
    uint64_t a = seed ^ 0xD6E8FEB86659FD93ull, b = seed + 0xA0761D6478BD642Full, c = ~seed, d = seed * 0xE7037ED1A0B428DBull + 1;
    size_t i = 0;
    for (; i + 32 <= bytes; i += 32) {
        uint64_t w[4];
        memcpy(w, p + i, 32);
        a = mix64(a, w[0]); b = mix64(b, w[1]); c = mix64(c, w[2]); d = mix64(d, w[3]);
    }
    for (; i < bytes; i++) a = mix64(a, p[i]);
    return mix64(mix64(mix64(mix64(bytes, a), b), c), d);
}

static uint64_t parallel_hash(const void* data, size_t bytes, uint64_t seed)
{
    const unsigned nt = host_threads(bytes);
    if (!data || !bytes) return mix64(seed, bytes);
    if (nt == 1) return hash_range((const unsigned char*)data, bytes, seed);
    std::vector<uint64_t> part(nt, 0);
    std::vector<std::thread> pool;
    const size_t chunk = ((bytes + nt - 1) / nt + 63) & ~(size_t)63;
    for (unsigned t = 0; t < nt; t++) {
        const size_t b = std::min(bytes, (size_t)t * chunk), e = std::min(bytes, b + chunk);
        pool.emplace_back([=, &part]() { part[t] = hash_range((const unsigned char*)data + b, e - b, seed + t); });
    }
    for (std::thread& t : pool) t.join();
    uint64_t h = seed;
    for (uint64_t v : part) h = mix64(h, v);
    return h;
}

// Everything crt_scene_upload looks at, texels included: two descriptions with the same hash are taken to be the same scene.
static uint64_t scene_content_hash(const crt_scene_desc* sd)
{
    uint64_t h = 0x243F6A8885A308D3ull;
    h = mix64(h, parallel_hash(sd->hittables, sd->n_hittables * sizeof(crt_hittable), 1));
    h = mix64(h, parallel_hash(sd->materials, sd->n_materials * sizeof(crt_material), 2));
    h = mix64(h, parallel_hash(sd->textures, sd->n_textures * sizeof(crt_texture), 3));
    for (uint64_t i = 0; i < sd->n_images; i++) {
        const crt_image& im = sd->images[i];
        h = mix64(h, ((uint64_t)(uint32_t)im.width << 32) | (uint32_t)im.height);
        const bool present = im.pixels && im.width > 0 && im.height > 0;
        h = mix64(h, present ? parallel_hash(im.pixels, (size_t)im.width * im.height * 3, 4 + i) : 0x5555u);
    }
    return h;
}

static crt_status upload_images(crt_context* ctx, const crt_scene_desc* sd, std::vector<unsigned char*>& raw_rgb, uint64_t& h2d)
{
    std::vector<DevImage> himg(sd->n_images);
    for (uint64_t i = 0; i < himg.size(); i++) {
        const crt_image& im = sd->images[i];
        himg[i].texels = nullptr;
        himg[i].width = im.width;
        himg[i].height = im.height;
        if (!im.pixels || im.width <= 0 || im.height <= 0) continue;
        size_t npx = (size_t)im.width * im.height;
        uchar4* d = nullptr;
        unsigned char* raw = nullptr; // RGB8 as given; expanded on the device
        CU(cudaMalloc(&d, npx * 4));
        ctx->d_texels.push_back(d);
        CU(cudaMalloc(&raw, (npx * 3 + 15) & ~(size_t)15));
        raw_rgb.push_back(raw);
        CU(cudaMemcpyAsync(raw, im.pixels, npx * 3, cudaMemcpyHostToDevice, ctx->stream));
        k_rgb_to_rgba<<<(unsigned)((npx / 4 + 1 + 255) / 256), 256, 0, ctx->stream>>>(reinterpret_cast<const uint32_t*>(raw), d, npx);
        CU(cudaGetLastError());
        h2d += npx * 3;
        himg[i].texels = d;
    }
    CU(cudaMalloc(&ctx->d_images, std::max<size_t>(himg.size(), 1) * sizeof(DevImage)));
    if (!himg.empty()) CU(cudaMemcpyAsync(ctx->d_images, himg.data(), himg.size() * sizeof(DevImage), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream)); // himg is a local
    return CRT_OK;
}

// Records the peeled top of the tree in the device scene. prim_id: leaf slot -> hittable index (host copy).
static void install_peel(crt_context* ctx, const CrtPackedNode* top_nodes, size_t n_top_nodes, uint64_t n_prims, const int32_t* prim_id)
{
    DevScene& S = ctx->scene;
    int32_t entry = 0, codes[CRT_MAX_TOP] = {0, 0, 0, 0};
    const int n = getenv("CRT_NO_PEEL") ? 0 : crt_peel_top(top_nodes, n_top_nodes, n_prims, CRT_MAX_TOP, &entry, codes);
    S.entry = entry;
    S.n_top = n;
    for (int k = 0; k < CRT_MAX_TOP; k++) {
        S.top_code[k] = k < n ? codes[k] : 0;
        ctx->top_hittable[k] = k < n ? prim_id[codes[k] & CRT_LEAF_INDEX_MASK] : -1;
    }
}

static crt_status upload_packed(crt_context* ctx, const crt_scene_desc* sd, const CrtPackedScene& P, bool keep_images)
{
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    if (!keep_images) free_images(ctx);
    const size_t nn = P.nodes.size(), np = P.prim_id.size();
    uint64_t h2d = 0;
    {
        crt_status gst = ensure_geometry(ctx, nn, np);
        if (gst != CRT_OK) return gst;
    }
    CU(cudaMemcpyAsync(ctx->d_parent, P.parent.data(), nn * 4, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_leaf_ref, P.leaf_ref.data(), np * 4, cudaMemcpyHostToDevice, ctx->stream));
    h2d += (nn + np) * 4;
    ctx->abs_pad = P.abs_pad;
    CU(cudaMemcpyAsync(ctx->d_nodes, P.nodes.data(), nn * 64, cudaMemcpyHostToDevice, ctx->stream));
    h2d += nn * 64;
    if (np) {
        CU(cudaMemcpyAsync(ctx->d_prims, P.prims.data(), np * 16, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(ctx->d_prim_k, P.prim_k.data(), np * 4, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(ctx->d_mats, P.mats.data(), np * 32, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(ctx->d_prim_id, P.prim_id.data(), np * 4, cudaMemcpyHostToDevice, ctx->stream));
        h2d += np * (16 + 4 + 32 + 4);
    }
    std::vector<unsigned char*> raw_rgb;
    if (!keep_images) {
        crt_status ist = upload_images(ctx, sd, raw_rgb, h2d);
        if (ist != CRT_OK) return ist;
    }
    CU(cudaStreamSynchronize(ctx->stream));
    for (unsigned char* p : raw_rgb) cudaFree(p);
    // bookkeeping for edits
    ctx->slot_of.assign(sd->n_hittables, -1);
    for (size_t sl = 0; sl < np; sl++)
        if (P.prim_id[sl] >= 0) ctx->slot_of[P.prim_id[sl]] = (int32_t)sl;
    ctx->dirty.assign(sd->n_hittables, 0);
    ctx->dirty_list.clear();
    ctx->need_rebuild = false;

    DevScene& S = ctx->scene;
    S.nodes = ctx->d_nodes; S.prims = ctx->d_prims; S.prim_k = ctx->d_prim_k; S.mats = ctx->d_mats;
    S.prim_id = ctx->d_prim_id; S.images = ctx->d_images;
    S.n_nodes = (int)nn; S.n_prims = (int)np;
    S.smem_nodes = S.smem_prims = S.smem_mats = 0;
    install_peel(ctx, P.nodes.data(), nn, np, P.prim_id.data());
    ctx->stats = P.stats;
    ctx->stats.h2d_bytes = h2d;
    ctx->scene_h2d_bytes = h2d;
    ctx->have_scene = true;
    return CRT_OK;
}

// ---- BVH built on the device (lbvh.cuh) ----
// *outcome: 0 = nothing touched (scene too small for this path), 1 = gave up after the upload (tree too deep; the image
// textures of `sd` are on the device), 2 = the scene is in use.
static crt_status upload_gpu_built(crt_context* ctx, const crt_scene_desc* sd, bool keep_images, int* outcome)
{
    NvtxRange nvtx_range("bvh build: GPU LBVH");
    *outcome = 0;
    auto t0 = std::chrono::steady_clock::now();
    std::vector<uint32_t> active;
    float cmin[3], cmax[3], blo[3], bhi[3], abs_pad = 0.f;
    std::string err;
    if (!crt_prepare_gpu_build(sd, &active, cmin, cmax, blo, bhi, &abs_pad, &err)) return crt_set_error(ctx, CRT_ERR_INVALID, err.c_str());
    const uint32_t n = (uint32_t)active.size();
    auto t_prep = std::chrono::steady_clock::now();
    if (n < 2) return CRT_OK; // the degenerate trees (empty scene, one primitive) are the host builder's business
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    if (!keep_images) free_images(ctx);
    cudaStream_t st = ctx->stream;
    uint64_t h2d = 0;
    // everything that only lives during the build comes out of ONE device allocation (20 cudaMalloc/cudaFree pairs cost more
    // than the kernels): the raw description, sort keys and values, leaf codes, the unpermuted tree, cub's scratch
    std::vector<unsigned char> present(std::max<uint64_t>(sd->n_images, 1), 0);
    for (uint64_t i = 0; i < sd->n_images; i++) present[i] = sd->images[i].pixels && sd->images[i].width > 0 && sd->images[i].height > 0;
    std::vector<unsigned char*> raw_rgb;
    const size_t nn = (size_t)n - 1;
    size_t tmp_bytes = 0;
    {
        cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                                        (uint32_t*)nullptr, (uint32_t*)nullptr, (int)n, 0, 64, st);
        if (e != cudaSuccess) return cuda_fail(ctx, e, "cub::DeviceRadixSort::SortPairs (size query)");
    }
    size_t arena_bytes = 0;
    auto reserve = [&](size_t bytes) { const size_t at = arena_bytes; arena_bytes += (bytes + 255) & ~(size_t)255; return at; };
    const size_t o_h = reserve(sd->n_hittables * sizeof(crt_hittable)), o_m = reserve(std::max<uint64_t>(sd->n_materials, 1) * sizeof(crt_material)),
                 o_t = reserve(std::max<uint64_t>(sd->n_textures, 1) * sizeof(crt_texture)), o_ip = reserve(present.size()),
                 o_act = reserve((size_t)n * 4), o_k0 = reserve((size_t)n * 8), o_k1 = reserve((size_t)n * 8), o_v0 = reserve((size_t)n * 4),
                 o_v1 = reserve((size_t)n * 4), o_code = reserve((size_t)n * 4), o_perm = reserve(nn * 4), o_nodes2 = reserve(nn * 64),
                 o_parent2 = reserve(nn * 4), o_stats = reserve(sizeof(LbvhStats)), o_tmp = reserve(std::max<size_t>(tmp_bytes, 16)),
                 o_end = reserve(16);
    auto cleanup = [&]() {
        for (unsigned char* p : raw_rgb) cudaFree(p);
    };
#define CUB_(call)                                                                     \
    do {                                                                               \
        cudaError_t e__ = (call);                                                      \
        if (e__ != cudaSuccess) { cleanup(); return cuda_fail(ctx, e__, #call); }      \
    } while (0)
    if (arena_bytes > ctx->build_arena_bytes) { // kept between builds: see ensure_geometry
        cudaFree(ctx->build_arena);
        ctx->build_arena = nullptr;
        ctx->build_arena_bytes = 0;
        CUB_(cudaMalloc(&ctx->build_arena, arena_bytes));
        ctx->build_arena_bytes = arena_bytes;
    }
    unsigned char* arena = ctx->build_arena;
    crt_hittable* d_h = reinterpret_cast<crt_hittable*>(arena + o_h);
    crt_material* d_m = reinterpret_cast<crt_material*>(arena + o_m);
    crt_texture* d_t = reinterpret_cast<crt_texture*>(arena + o_t);
    unsigned char* d_ip = arena + o_ip;
    uint32_t* d_act = reinterpret_cast<uint32_t*>(arena + o_act);
    unsigned long long *d_k0 = reinterpret_cast<unsigned long long*>(arena + o_k0), *d_k1 = reinterpret_cast<unsigned long long*>(arena + o_k1);
    uint32_t *d_v0 = reinterpret_cast<uint32_t*>(arena + o_v0), *d_v1 = reinterpret_cast<uint32_t*>(arena + o_v1);
    int* d_code = reinterpret_cast<int*>(arena + o_code);
    int* d_perm = reinterpret_cast<int*>(arena + o_perm);
    float4* d_nodes2 = reinterpret_cast<float4*>(arena + o_nodes2);
    int* d_parent2 = reinterpret_cast<int*>(arena + o_parent2);
    LbvhStats* d_stats = reinterpret_cast<LbvhStats*>(arena + o_stats);
    void* d_tmp = arena + o_tmp;
    {
        crt_status gst = ensure_geometry(ctx, nn, n);
        if (gst != CRT_OK) { cleanup(); return gst; }
    }
    {
        // The caller's arrays are pageable: cudaMemcpyAsync from them is a synchronous, single-threaded staging copy at ~10 GB/s
        // (22 ms of a 34 ms upload for a million spheres). Copy them with several threads into a page-locked buffer the context
        // keeps, and let the DMA engine take it from there.
        const size_t b_h = sd->n_hittables * sizeof(crt_hittable), b_m = sd->n_materials * sizeof(crt_material),
                     b_t = sd->n_textures * sizeof(crt_texture), b_a = (size_t)n * 4;
        const size_t s_h = 0, s_m = (s_h + b_h + 255) & ~(size_t)255, s_t = (s_m + b_m + 255) & ~(size_t)255,
                     s_a = (s_t + b_t + 255) & ~(size_t)255, s_end = s_a + b_a;
        if (s_end > ctx->pinned_stage_bytes) {
            if (ctx->pinned_stage) cudaFreeHost(ctx->pinned_stage);
            ctx->pinned_stage = nullptr;
            ctx->pinned_stage_bytes = 0;
            CUB_(cudaHostAlloc((void**)&ctx->pinned_stage, s_end, cudaHostAllocDefault));
            ctx->pinned_stage_bytes = s_end;
        }
        unsigned char* ps = ctx->pinned_stage;
        parallel_copy(ps + s_h, sd->hittables, b_h);
        CUB_(cudaMemcpyAsync(d_h, ps + s_h, b_h, cudaMemcpyHostToDevice, st));
        parallel_copy(ps + s_m, sd->materials, b_m);
        CUB_(cudaMemcpyAsync(d_m, ps + s_m, b_m, cudaMemcpyHostToDevice, st));
        parallel_copy(ps + s_t, sd->textures, b_t);
        CUB_(cudaMemcpyAsync(d_t, ps + s_t, b_t, cudaMemcpyHostToDevice, st));
        parallel_copy(ps + s_a, active.data(), b_a);
        CUB_(cudaMemcpyAsync(d_act, ps + s_a, b_a, cudaMemcpyHostToDevice, st));
    }
    CUB_(cudaMemcpyAsync(d_ip, present.data(), present.size(), cudaMemcpyHostToDevice, st));
    h2d += sd->n_hittables * sizeof(crt_hittable) + sd->n_materials * sizeof(crt_material) + sd->n_textures * sizeof(crt_texture) + (uint64_t)n * 4;

    auto t_alloc = std::chrono::steady_clock::now();
    LbvhRaw R;
    R.hittables = d_h; R.materials = d_m; R.textures = d_t; R.image_present = d_ip; R.active = d_act; R.n = n;
    float3 c0 = make_float3(cmin[0], cmin[1], cmin[2]), ie;
    ie.x = cmax[0] > cmin[0] ? 1.0f / (cmax[0] - cmin[0]) : 0.f;
    ie.y = cmax[1] > cmin[1] ? 1.0f / (cmax[1] - cmin[1]) : 0.f;
    ie.z = cmax[2] > cmin[2] ? 1.0f / (cmax[2] - cmin[2]) : 0.f;
    const unsigned gb = (n + 255) / 256, gn = (unsigned)((nn + 255) / 256);
    const float sx = bhi[0] - blo[0], sy = bhi[1] - blo[1], sz = bhi[2] - blo[2];
    const float big_area = 0.25f * (sx * sy + sy * sz + sz * sx); // of the half-area; NaN / inf: nothing qualifies, or everything
    k_lbvh_keys<<<gb, 256, 0, st>>>(R, c0, ie, big_area, d_k0, d_v0);
    CUB_(cub::DeviceRadixSort::SortPairs(d_tmp, tmp_bytes, d_k0, d_k1, d_v0, d_v1, (int)n, 0, 64, st));
    k_lbvh_pack<<<gb, 256, 0, st>>>(R, d_v1, ctx->d_prims, ctx->d_prim_k, ctx->d_mats, ctx->d_prim_id, d_code);
    k_lbvh_hierarchy<<<gn, 256, 0, st>>>(d_k1, d_code, (int)n, d_nodes2, d_parent2, ctx->d_leaf_ref);
    CUB_(cudaMemsetAsync(ctx->d_visits, 0, nn * 4, st));
    k_bvh_refit<<<gb, 256, 0, st>>>(d_nodes2, ctx->d_prims, ctx->d_prim_k, d_parent2, ctx->d_leaf_ref, ctx->d_visits, (int)n, abs_pad);
    CUB_(cudaMemsetAsync(d_stats, 0, sizeof(LbvhStats), st));
    k_lbvh_stats<<<gb, 256, 0, st>>>(d_nodes2, d_parent2, ctx->d_leaf_ref, (int)n, d_stats);
    // top of the tree breadth-first (what the kernels stage in shared memory), the rest keeps the sorted order
    k_lbvh_identity<<<gn, 256, 0, st>>>(d_perm, (int)nn);
    k_lbvh_top_order<<<1, 32, 0, st>>>(d_nodes2, (int)nn, LBVH_TOP_MAX, d_perm);
    k_lbvh_permute<<<gn, 256, 0, st>>>(d_nodes2, ctx->d_nodes, d_perm, d_parent2, ctx->d_parent, (int)nn);
    k_lbvh_permute_leaf_refs<<<gb, 256, 0, st>>>(ctx->d_leaf_ref, d_perm, (int)n);
    CUB_(cudaGetLastError());
    if (!keep_images) {
        crt_status ist = upload_images(ctx, sd, raw_rgb, h2d);
        if (ist != CRT_OK) { cleanup(); return ist; }
    }
    LbvhStats hs;
    std::vector<int32_t> prim_id(n);
    CrtPackedNode top_nodes[16]; // the chain of huge leaves under the root lives in the first few breadth-first nodes
    const size_t n_top_nodes = std::min<size_t>(16, nn);
    CUB_(cudaMemcpyAsync(top_nodes, ctx->d_nodes, n_top_nodes * sizeof(CrtPackedNode), cudaMemcpyDeviceToHost, st));
    CUB_(cudaMemcpyAsync(&hs, d_stats, sizeof(hs), cudaMemcpyDeviceToHost, st));
    CUB_(cudaMemcpyAsync(prim_id.data(), ctx->d_prim_id, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    CUB_(cudaStreamSynchronize(st));
    auto t_gpu = std::chrono::steady_clock::now();
    cleanup();
#undef CUB_
    if (hs.max_depth >= CRT_STACK_SIZE) { // pathological key distribution: let the SAH builder (depth-bounded by its splits) do it
        free_geometry(ctx);
        *outcome = 1;
        return CRT_OK;
    }
    ctx->abs_pad = abs_pad;
    ctx->slot_of.assign(sd->n_hittables, -1);
    for (size_t sl = 0; sl < n; sl++) ctx->slot_of[prim_id[sl]] = (int32_t)sl;
    ctx->dirty.assign(sd->n_hittables, 0);
    ctx->dirty_list.clear();
    ctx->need_rebuild = false;
    DevScene& S = ctx->scene;
    S.nodes = ctx->d_nodes; S.prims = ctx->d_prims; S.prim_k = ctx->d_prim_k; S.mats = ctx->d_mats;
    S.prim_id = ctx->d_prim_id; S.images = ctx->d_images;
    S.n_nodes = (int)nn; S.n_prims = (int)n;
    S.smem_nodes = S.smem_prims = S.smem_mats = 0;
    install_peel(ctx, top_nodes, n_top_nodes, n, prim_id.data());
    memset(&ctx->stats, 0, sizeof(ctx->stats));
    ctx->stats.n_prims = n;
    ctx->stats.n_nodes = nn;
    ctx->stats.max_depth = (uint64_t)hs.max_depth;
    ctx->stats.sah_cost = hs.root_area > 0.f ? 1.0 + (double)hs.area_sum / hs.root_area : 0.0;
    ctx->stats.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    ctx->stats.node_bytes = (uint64_t)nn * 64;
    ctx->stats.prim_bytes = (uint64_t)n * (16 + 4 + 32 + 4);
    ctx->stats.h2d_bytes = h2d;
    ctx->stats.builder = CRT_BUILDER_GPU_LBVH;
    ctx->scene_h2d_bytes = h2d;
    ctx->have_scene = true;
    *outcome = 2;
    if (getenv("CRT_BVH_TIMING")) {
        auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
        fprintf(stderr, "[crt] gpu bvh (LBVH): host prepare %.1f ms, free + alloc + enqueue copies %.1f ms, kernels + copies + sync %.1f ms, free + bookkeeping %.1f ms\n",
                ms(t0, t_prep), ms(t_prep, t_alloc), ms(t_alloc, t_gpu), ms(t_gpu, std::chrono::steady_clock::now()));
    }
    return CRT_OK;
}

// Builds the BVH of `sd` with the builder the context is set to and makes it the scene in use.
static crt_status build_and_upload_tree(crt_context* ctx, const crt_scene_desc* sd, bool keep_images, const char* who);
static crt_status build_and_upload(crt_context* ctx, const crt_scene_desc* sd, bool keep_images, const char* who)
{
    crt_status st = build_and_upload_tree(ctx, sd, keep_images, who);
    if (st != CRT_OK) return st;
    // scenes without rectangles run leaf visits without the type dispatch (trace.cuh leaf_step SPH)
    int only = 1;
    for (uint64_t i = 0; i < sd->n_hittables && only; i++)
        if (sd->hittables[i].active && sd->hittables[i].type != CRT_SPHERE) only = 0;
    ctx->scene.spheres_only = only;
    return CRT_OK;
}

static crt_status build_and_upload_tree(crt_context* ctx, const crt_scene_desc* sd, bool keep_images, const char* who)
{
    uint64_t n_active_bound = sd ? sd->n_hittables : 0;
    const bool gpu = ctx->bvh_builder == CRT_BUILDER_GPU_LBVH ||
                     (ctx->bvh_builder == CRT_BUILDER_AUTO && n_active_bound >= ctx->gpu_build_min_prims);
    if (gpu && sd && sd->hittables) {
        int outcome = 0;
        crt_status st = upload_gpu_built(ctx, sd, keep_images, &outcome);
        if (st != CRT_OK || outcome == 2) return st;
        if (outcome == 1) keep_images = true; // the textures are already on the device
    }
    CrtPackedScene P;
    std::string err;
    if (!crt_pack_scene(sd, &P, &err)) return crt_set_error(ctx, CRT_ERR_INVALID, err.c_str());
    if (P.stats.max_depth >= CRT_STACK_SIZE)
        return crt_set_error(ctx, CRT_ERR_UNSUPPORTED, (std::string(who) + ": BVH deeper than the traversal stack").c_str());
    crt_status st = upload_packed(ctx, sd, P, keep_images);
    if (st == CRT_OK) ctx->stats.builder = CRT_BUILDER_HOST_SAH;
    return st;
}

extern "C" {

crt_status crt_set_bvh_builder(crt_context* ctx, int32_t builder)
{
    if (!ctx) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_set_bvh_builder: ctx is NULL");
    if (builder < CRT_BUILDER_AUTO || builder > CRT_BUILDER_GPU_LBVH) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_set_bvh_builder: unknown builder");
    ctx->bvh_builder = builder;
    return CRT_OK;
}

crt_status crt_scene_upload(crt_context* ctx, const crt_scene_desc* scene)
{
    NvtxRange nvtx_range("crt_scene_upload");
    if (!ctx || !scene) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_scene_upload: NULL argument");
    if ((scene->n_hittables && !scene->hittables) || (scene->n_materials && !scene->materials) || (scene->n_textures && !scene->textures) ||
        (scene->n_images && !scene->images))
        return crt_set_error(ctx, CRT_ERR_INVALID, "crt_scene_upload: NULL array with a non-zero count");
    auto t0 = std::chrono::steady_clock::now();
    uint64_t hash = 0;
    if (ctx->scene_cache) {
        hash = scene_content_hash(scene);
        if (ctx->have_scene && ctx->scene_hash_valid && hash == ctx->scene_hash && ctx->dirty_list.empty() && !ctx->need_rebuild &&
            scene->n_hittables == ctx->host.hittables.size() && scene->n_materials == ctx->host.materials.size() &&
            scene->n_textures == ctx->host.textures.size() && scene->n_images == ctx->host.images.size()) {
            // the scene already on the device: nothing to validate, build or copy
            ctx->stats.h2d_bytes = 0;
            ctx->stats.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            ctx->stats.cache_hits++;
            return CRT_OK;
        }
    }
    ctx->scene_hash_valid = false;
    crt_status st = build_and_upload(ctx, scene, false, "crt_scene_upload");
    if (st != CRT_OK) return st;
    ctx->stats.cache_hits = 0;
    ctx->host.hittables.resize(scene->n_hittables);
    parallel_copy(ctx->host.hittables.data(), scene->hittables, scene->n_hittables * sizeof(crt_hittable));
    ctx->host.materials.resize(scene->n_materials);
    parallel_copy(ctx->host.materials.data(), scene->materials, scene->n_materials * sizeof(crt_material));
    ctx->host.textures.resize(scene->n_textures);
    parallel_copy(ctx->host.textures.data(), scene->textures, scene->n_textures * sizeof(crt_texture));
    ctx->host.images.assign(scene->images, scene->images + scene->n_images);
    ctx->scene_hash = hash;
    ctx->scene_hash_valid = ctx->scene_cache != 0;
    return CRT_OK;
}

crt_status crt_set_scene_cache(crt_context* ctx, int32_t enabled)
{
    if (!ctx) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_set_scene_cache: ctx is NULL");
    ctx->scene_cache = enabled != 0;
    if (!enabled) ctx->scene_hash_valid = false;
    return CRT_OK;
}

crt_status crt_scene_from_reference(crt_context* ctx, const void* world)
{
    if (!ctx || !world) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_scene_from_reference: NULL argument");
    CrtFlatScene flat;
    std::string err;
    if (!crt_flatten_reference_world(world, &flat, &err)) return crt_set_error(ctx, CRT_ERR_INVALID, err.c_str());
    crt_scene_desc d = flat.desc();
    return crt_scene_upload(ctx, &d);
}

// ---- scene edits (the caller side of the boundary: CudaLayer.cpp:493-494,918-1387 rebuild the BVH after every edit) ----
static void mark_dirty(crt_context* ctx, uint64_t index)
{
    if (!ctx->dirty[index]) {
        ctx->dirty[index] = 1;
        ctx->dirty_list.push_back(index);
    }
}

crt_status crt_scene_edit_hittable(crt_context* ctx, uint64_t index, const crt_hittable* h)
{
    if (!ctx || !h) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_scene_edit_hittable: NULL argument");
    if (!ctx->have_scene) return crt_set_error(ctx, CRT_ERR_STATE, "crt_scene_edit_hittable: no scene uploaded");
    if (index >= ctx->host.hittables.size()) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_scene_edit_hittable: index out of range");
    if (h->type < CRT_SPHERE || h->type > CRT_YZRECT) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_scene_edit_hittable: unknown type");
    if (h->material < 0 || (uint64_t)h->material >= ctx->host.materials.size())
        return crt_set_error(ctx, CRT_ERR_INVALID, "crt_scene_edit_hittable: material index out of range");
    if (h->active && (!std::isfinite(h->cx) || !std::isfinite(h->cy) || !std::isfinite(h->cz) || !std::isfinite(h->a) ||
                      (h->type != CRT_SPHERE && !std::isfinite(h->b))))
        return crt_set_error(ctx, CRT_ERR_INVALID, "crt_scene_edit_hittable: non-finite coordinate or size (deactivate the hittable instead)");
    const crt_hittable old = ctx->host.hittables[index];
    // switching a primitive on/off or changing its kind changes the tree; everything else is a refit
    if ((old.active != 0) != (h->active != 0) || (h->active && old.type != h->type)) ctx->need_rebuild = true;
    ctx->host.hittables[index] = *h;
    mark_dirty(ctx, index);
    ctx->scene_hash_valid = false; // the scene on the device is about to differ from the description that was hashed
    return CRT_OK;
}

crt_status crt_scene_edit_material(crt_context* ctx, uint64_t index, const crt_material* m)
{
    if (!ctx || !m) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_scene_edit_material: NULL argument");
    if (!ctx->have_scene) return crt_set_error(ctx, CRT_ERR_STATE, "crt_scene_edit_material: no scene uploaded");
    if (index >= ctx->host.materials.size()) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_scene_edit_material: index out of range");
    if (m->type < CRT_LAMBERTIAN || m->type > CRT_DIFFUSELIGHT) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_scene_edit_material: unknown type");
    if (m->type != CRT_DIELECTRIC && (m->texture < 0 || (uint64_t)m->texture >= ctx->host.textures.size()))
        return crt_set_error(ctx, CRT_ERR_INVALID, "crt_scene_edit_material: texture index out of range");
    ctx->host.materials[index] = *m;
    ctx->scene_hash_valid = false;
    for (uint64_t i = 0; i < ctx->host.hittables.size(); i++) // every primitive carries its own packed material record
        if (ctx->host.hittables[i].material == (int32_t)index) mark_dirty(ctx, i);
    return CRT_OK;
}

crt_status crt_scene_commit(crt_context* ctx, int force_rebuild, int* out_rebuilt)
{
    NvtxRange nvtx_range("crt_scene_commit");
    if (!ctx) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_scene_commit: ctx is NULL");
    if (!ctx->have_scene) return crt_set_error(ctx, CRT_ERR_STATE, "crt_scene_commit: no scene uploaded");
    if (out_rebuilt) *out_rebuilt = 0;
    if (ctx->dirty_list.empty() && !force_rebuild) return CRT_OK;
    CU(cudaSetDevice(ctx->device));
    const crt_scene_desc d = ctx->host.desc();
    if (force_rebuild || ctx->need_rebuild || ctx->scene.n_prims < 2) {
        if (out_rebuilt) *out_rebuilt = 1;
        return build_and_upload(ctx, &d, true, "crt_scene_commit"); // the image textures stay where they are
    }
    // refit path: scatter the edited records, then recompute every box bottom-up on the GPU
    auto t0 = std::chrono::steady_clock::now();
    // The rebuild path validates through crt_pack_scene. Here the records are packed directly, and a hittable may have been
    // re-pointed at a material that no active hittable used at upload, so nothing has looked at its texture chain yet: check
    // before anything is touched (on failure the edits stay pending; fix the material and commit again).
    for (uint64_t idx : ctx->dirty_list) {
        const crt_hittable& h = ctx->host.hittables[idx];
        std::string merr;
        if (h.active && !crt_check_material(&d, h.material, &merr)) return crt_set_error(ctx, CRT_ERR_INVALID, ("crt_scene_commit: " + merr).c_str());
    }
    std::vector<EditRec> recs;
    recs.reserve(ctx->dirty_list.size());
    float extent = 0.f;
    for (uint64_t idx : ctx->dirty_list) {
        ctx->dirty[idx] = 0;
        const crt_hittable& h = ctx->host.hittables[idx];
        const int32_t slot = ctx->slot_of[idx];
        if (!h.active || slot < 0) continue;
        EditRec r;
        float pr[4], mr[8];
        crt_pack_prim(h, pr, &r.k);
        crt_pack_material(&d, ctx->host.materials[h.material], mr);
        r.prim = make_float4(pr[0], pr[1], pr[2], pr[3]);
        r.m0 = make_float4(mr[0], mr[1], mr[2], mr[3]);
        r.m1 = make_float4(mr[4], mr[5], mr[6], mr[7]);
        r.slot = slot;
        r.leaf_ref = 0; // looked up on the device (d_leaf_ref[slot])
        r.child_code = ~(int32_t)(((uint32_t)h.type << CRT_LEAF_TYPE_SHIFT) |
                                  ((uint32_t)ctx->host.materials[h.material].type << CRT_LEAF_MAT_SHIFT) | (uint32_t)slot);
        recs.push_back(r);
        for (int k = 0; k < ctx->scene.n_top; k++) // a peeled leaf is named by its code in the kernel arguments, not by the node
            if (ctx->top_hittable[k] == (int32_t)idx) ctx->scene.top_code[k] = ~r.child_code;
        const float reach = fabsf(h.a) + fabsf(h.b);
        extent = std::max(extent, std::max(fabsf(h.cx), std::max(fabsf(h.cy), fabsf(h.cz))) + reach);
    }
    ctx->dirty_list.clear();
    // the absolute box padding scales with the scene extent (scene_pack.cpp); edits may only grow it
    ctx->abs_pad = std::max(ctx->abs_pad, 1e-7f + 4e-7f * extent);
    if (!recs.empty()) {
        if (recs.size() > ctx->edit_recs_cap) {
            cudaFree(ctx->d_edit_recs);
            ctx->d_edit_recs = nullptr;
            ctx->edit_recs_cap = 0;
            const size_t cap = std::max<size_t>(4096, recs.size() + recs.size() / 2);
            CU(cudaMalloc(&ctx->d_edit_recs, cap * sizeof(EditRec)));
            ctx->edit_recs_cap = cap;
        }
        EditRec* d_recs = static_cast<EditRec*>(ctx->d_edit_recs);
        CU(cudaMemcpyAsync(d_recs, recs.data(), recs.size() * sizeof(EditRec), cudaMemcpyHostToDevice, ctx->stream));
        const int n = (int)recs.size();
        k_apply_edits<<<(n + 255) / 256, 256, 0, ctx->stream>>>(d_recs, n, ctx->d_prims, ctx->d_prim_k, ctx->d_mats, ctx->d_nodes, ctx->d_leaf_ref);
        CU(cudaGetLastError());
        CU(cudaMemsetAsync(ctx->d_visits, 0, (size_t)ctx->scene.n_nodes * 4, ctx->stream));
        const int np = ctx->scene.n_prims;
        k_bvh_refit<<<(np + 255) / 256, 256, 0, ctx->stream>>>(ctx->d_nodes, ctx->d_prims, ctx->d_prim_k, ctx->d_parent, ctx->d_leaf_ref,
                                                              ctx->d_visits, np, ctx->abs_pad);
        ctx->nodes_gen++;
        CU(cudaGetLastError());
        ctx->launches += 2;
        CU(cudaStreamSynchronize(ctx->stream)); // recs is a local
        ctx->stats.h2d_bytes = recs.size() * sizeof(EditRec);
    }
    ctx->stats.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return CRT_OK;
}

crt_status crt_bvh_build_check(const crt_scene_desc* scene, crt_bvh_stats* out)
{
    if (!scene || !out) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_bvh_build_check: NULL argument");
    CrtPackedScene P;
    std::string err;
    if (!crt_pack_scene(scene, &P, &err)) return crt_set_error(nullptr, CRT_ERR_INVALID, err.c_str());
    *out = P.stats;
    const size_t nn = P.nodes.size(), np = P.stats.n_prims;
    std::vector<int> seen(np ? np : 1, 0);
    struct Item { int node; double lo[3], hi[3]; };
    std::vector<Item> st;
    Item root;
    root.node = 0;
    for (int i = 0; i < 3; i++) { root.lo[i] = -INFINITY; root.hi[i] = INFINITY; }
    st.push_back(root);
    size_t visited = 0;
    while (!st.empty()) {
        Item it = st.back();
        st.pop_back();
        if (it.node < 0 || (size_t)it.node >= nn) return crt_set_error(nullptr, CRT_ERR_INVALID, "bvh check: child index out of range");
        if (++visited > nn) return crt_set_error(nullptr, CRT_ERR_INVALID, "bvh check: cycle or shared node");
        const CrtPackedNode& n = P.nodes[it.node];
        for (int c = 0; c < 2; c++) {
            // stored as centre / half-extent; a child may stick out of its parent by the rounding of the two conversions
            double b[6];
            for (int i = 0; i < 3; i++) {
                b[i] = (double)n.w[6 * c + i] - (double)n.w[6 * c + 3 + i];
                b[3 + i] = (double)n.w[6 * c + i] + (double)n.w[6 * c + 3 + i];
            }
            if (np >= 2)
                for (int i = 0; i < 3; i++) {
                    const double tol = 1e-6 * (fabs(b[i]) + fabs(b[3 + i])) + 1e-30;
                    if (b[i] < it.lo[i] - tol || b[3 + i] > it.hi[i] + tol) return crt_set_error(nullptr, CRT_ERR_INVALID, "bvh check: child box outside its parent's box");
                }
            if (n.child[c] >= 0) {
                Item ch;
                ch.node = n.child[c];
                for (int i = 0; i < 3; i++) { ch.lo[i] = b[i]; ch.hi[i] = b[3 + i]; }
                st.push_back(ch);
            }
            else if (np > 0) {
                const uint32_t code = (uint32_t)~n.child[c];
                const uint32_t slot = code & CRT_LEAF_INDEX_MASK;
                if (slot >= np) return crt_set_error(nullptr, CRT_ERR_INVALID, "bvh check: leaf slot out of range");
                if (np >= 2 || c == 0) seen[slot]++;
                const float* pr = &P.prims[4 * (size_t)slot];
                if ((code >> CRT_LEAF_TYPE_SHIFT) == CRT_SPHERE) {
                    const float r = fabsf(pr[3]);
                    for (int i = 0; i < 3; i++)
                        if ((double)pr[i] - r < b[i] || (double)pr[i] + r > b[3 + i]) return crt_set_error(nullptr, CRT_ERR_INVALID, "bvh check: sphere outside its leaf box");
                }
            }
        }
    }
    for (size_t i = 0; i < np; i++)
        if (seen[i] != 1) return crt_set_error(nullptr, CRT_ERR_INVALID, "bvh check: a primitive is referenced by no leaf or by several");
    return CRT_OK;
}

crt_status crt_bvh_get_stats(const crt_context* ctx, crt_bvh_stats* out)
{
    if (!ctx || !out) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_bvh_get_stats: NULL argument");
    if (!ctx->have_scene) return crt_set_error(nullptr, CRT_ERR_STATE, "crt_bvh_get_stats: no scene uploaded");
    *out = ctx->stats;
    return CRT_OK;
}

// ------------------------------------------------------------------------------------------
// camera
// ------------------------------------------------------------------------------------------
static void derive_camera(crt_context* ctx)
{
    const crt_input_struct& in = ctx->inputs;
    DevCamera& C = ctx->cam;
    C.ox = in.origin_x; C.oy = in.origin_y; C.oz = in.origin_z;
    C.fx = in.orientation_x; C.fy = in.orientation_y; C.fz = in.orientation_z;
    C.ux = in.up_x; C.uy = in.up_y; C.uz = in.up_z;
    // rightV = Normalize(Cross(upV, forwardV)), Kernel.cu:133 with Math.cuh:157-161,225-229; host IEEE binary32
    volatile float cx = C.uy * C.fz - C.uz * C.fy;
    volatile float cy = -(C.ux * C.fz - C.uz * C.fx);
    volatile float cz = C.ux * C.fy - C.uy * C.fx;
    volatile float dd = cx * cx + cy * cy + cz * cz;
    volatile float inv = 1 / sqrtf(dd);
    C.rx = inv * cx; C.ry = inv * cy; C.rz = inv * cz;
    C.far_plane = in.far_plane; C.near_plane = in.near_plane; C.fov = in.fov;
    volatile float k = 1.0f / in.fov;
    C.far_k = k * 10.0f;
    for (int i = 0; i < 3; i++) { C.bg0[i] = in.background_start[i]; C.bg1[i] = in.background_end[i]; }
    C.lens_radius = ctx->aperture * 0.5f;
    C.focus_dist = ctx->focus_dist;
}

crt_status crt_set_camera(crt_context* ctx, const crt_input_struct* inputs)
{
    if (!ctx || !inputs) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_set_camera: NULL argument");
    ctx->inputs = *inputs;
    derive_camera(ctx);
    ctx->have_cam = true;
    return CRT_OK;
}

crt_status crt_set_lens(crt_context* ctx, float aperture, float focus_dist)
{
    if (!ctx) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_set_lens: ctx is NULL");
    if (!(aperture >= 0.f) || !(focus_dist > 0.f)) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_set_lens: aperture must be >= 0 and focus_dist > 0");
    ctx->aperture = aperture;
    ctx->focus_dist = focus_dist;
    if (ctx->have_cam) derive_camera(ctx);
    return CRT_OK;
}

// ------------------------------------------------------------------------------------------
// framebuffer
// ------------------------------------------------------------------------------------------
crt_status crt_framebuffer_resize(crt_context* ctx, uint32_t width, uint32_t height)
{
    if (!ctx) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_framebuffer_resize: ctx is NULL");
    if (width == 0 || height == 0 || (uint64_t)width * height > 0x7FFFFFFFull)
        return crt_set_error(ctx, CRT_ERR_INVALID, "crt_framebuffer_resize: bad size");
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    const bool external = ctx->d_accum && ctx->d_accum != ctx->d_accum_own;
    if (width != ctx->width || height != ctx->height || !ctx->d_rgba) {
        cudaFree(ctx->d_accum_own); ctx->d_accum_own = nullptr;
        cudaFree(ctx->d_accum_f); ctx->d_accum_f = nullptr;
        cudaFree(ctx->d_rgba); ctx->d_rgba = nullptr;
        cudaFree(ctx->d_ids); ctx->d_ids = nullptr;
        size_t n = (size_t)width * height;
        CU(cudaMalloc(&ctx->d_accum_own, n * 4 * sizeof(long long)));
        CU(cudaMalloc(&ctx->d_accum_f, n * sizeof(float4)));
        CU(cudaMalloc(&ctx->d_rgba, n * 4));
        CU(cudaMalloc(&ctx->d_ids, n * 4));
        ctx->width = width;
        ctx->height = height;
        if (!external) ctx->d_accum = ctx->d_accum_own;
        else ctx->d_accum = ctx->d_accum_own; // an external buffer of the old size is no longer valid
    }
    return crt_accum_clear(ctx);
}

crt_status crt_accum_clear(crt_context* ctx)
{
    if (!ctx) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_accum_clear: ctx is NULL");
    if (!ctx->d_accum) return crt_set_error(ctx, CRT_ERR_STATE, "crt_accum_clear: no framebuffer");
    CU(cudaSetDevice(ctx->device));
    uint32_t n = ctx->width * ctx->height * 2u; // two 16-byte halves per pixel
    k_clear<<<(n + 255) / 256, 256, 0, ctx->stream>>>(reinterpret_cast<longlong2*>(ctx->d_accum), n);
    ctx->launches++;
    CU(cudaGetLastError());
    return CRT_OK;
}

crt_status crt_set_accum_device_buffer(crt_context* ctx, void* device_int64x4)
{
    if (!ctx) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_set_accum_device_buffer: ctx is NULL");
    if (!ctx->d_accum_own) return crt_set_error(ctx, CRT_ERR_STATE, "crt_set_accum_device_buffer: call crt_framebuffer_resize first");
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    if (((uintptr_t)device_int64x4 & 31u) != 0)
        return crt_set_error(ctx, CRT_ERR_INVALID, "crt_set_accum_device_buffer: the buffer must be 32-byte aligned");
    ctx->d_accum = device_int64x4 ? (long long*)device_int64x4 : ctx->d_accum_own;
    return CRT_OK;
}

void* crt_accum_device_ptr(crt_context* ctx)
{
    return ctx ? (void*)ctx->d_accum : nullptr;
}

} // extern "C"

// ------------------------------------------------------------------------------------------
// render
// ------------------------------------------------------------------------------------------
// Decides how much of the scene each CTA stages in shared memory. Everything if it fits next to
// `blocks_per_sm` resident CTAs; otherwise the breadth-first prefix of the node array (the hot top levels).
static size_t plan_smem(const crt_context* ctx, int blocks_per_sm, DevScene& S, bool& all)
{
    long budget = ctx->smem_per_sm / std::max(blocks_per_sm, 1) - 1024;
    budget = std::min<long>(budget, ctx->smem_optin);
    if (ctx->smem_budget_override >= 0) budget = std::min<long>(ctx->smem_budget_override, ctx->smem_optin);
    budget = std::max<long>(budget, 0);
    const long node_b = 64; // the megakernel keeps the unpadded word-major node table (trace.cuh)
    const long full = (long)S.n_nodes * node_b + (long)S.n_prims * 48;
    if (full <= budget) {
        S.smem_nodes = S.n_nodes; S.smem_prims = S.n_prims; S.smem_mats = S.n_prims;
        all = true;
        return (size_t)full;
    }
    all = false;
    long rem = budget;
    S.smem_nodes = (int)std::min<long>(S.n_nodes, rem / node_b); rem -= (long)S.smem_nodes * node_b;
    S.smem_prims = (int)std::min<long>(S.n_prims, rem / 16); rem -= (long)S.smem_prims * 16;
    S.smem_mats = (int)std::min<long>(S.n_prims, rem / 32); rem -= (long)S.smem_mats * 32;
    return (size_t)(budget - rem);
}

template <bool EXACT, bool COUNT, bool ALL>
static cudaError_t launch_mega(const RenderArgs& A, dim3 grid, size_t smem, cudaStream_t st)
{
    cudaError_t e = cudaFuncSetAttribute(k_render_mega<EXACT, COUNT, ALL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_render_mega<EXACT, COUNT, ALL><<<grid, CRT_BLOCK_THREADS, smem, st>>>(A);
    return cudaGetLastError();
}

static cudaError_t dispatch_mega(const RenderArgs& A, dim3 grid, size_t smem, cudaStream_t st, bool exact, bool count, bool all)
{
    if (exact) {
        if (count) return all ? launch_mega<true, true, true>(A, grid, smem, st) : launch_mega<true, true, false>(A, grid, smem, st);
        return all ? launch_mega<true, false, true>(A, grid, smem, st) : launch_mega<true, false, false>(A, grid, smem, st);
    }
    if (count) return all ? launch_mega<false, true, true>(A, grid, smem, st) : launch_mega<false, true, false>(A, grid, smem, st);
    return all ? launch_mega<false, false, true>(A, grid, smem, st) : launch_mega<false, false, false>(A, grid, smem, st);
}

static int mega_blocks_per_sm(bool exact, bool count)
{
    // occupancy of the variant with no dynamic shared memory; used only to size the smem budget
    int n = 0;
    cudaError_t e;
    if (exact) e = count ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_render_mega<true, true, true>, CRT_BLOCK_THREADS, 0)
                         : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_render_mega<true, false, true>, CRT_BLOCK_THREADS, 0);
    else e = count ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_render_mega<false, true, true>, CRT_BLOCK_THREADS, 0)
                   : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_render_mega<false, false, true>, CRT_BLOCK_THREADS, 0);
    if (e != cudaSuccess || n < 1) n = 4;
    return n;
}

// ------------------------------------------------------------------------------------------
// wavefront driver
// ------------------------------------------------------------------------------------------
#define CRT_MAX_DEVICES 64

template <int THRESH, bool ALL> static cudaError_t launch_extend(const WfArgs& W, int grid, size_t smem, cudaStream_t st)
{
    // per instantiation and per device (function attributes belong to a device's context); the attribute is sticky
    static size_t configured[CRT_MAX_DEVICES] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    size_t& configured_smem = configured[dev & (CRT_MAX_DEVICES - 1)];
    if (smem > configured_smem) {
        cudaError_t e = cudaFuncSetAttribute(k_wf_extend<THRESH, ALL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured_smem = smem;
    }
    k_wf_extend<THRESH, ALL><<<grid, WF_EXT_THREADS, smem, st>>>(W);
    return cudaGetLastError();
}

static cudaError_t dispatch_extend(const WfArgs& W, int grid, size_t smem, cudaStream_t st, int thresh, bool all)
{
#define CRT_EXT_CASE(T) case T: return all ? launch_extend<T, true>(W, grid, smem, st) : launch_extend<T, false>(W, grid, smem, st)
    switch (thresh) {
        CRT_EXT_CASE(8);
        CRT_EXT_CASE(16);
        CRT_EXT_CASE(20);
        CRT_EXT_CASE(28);
        CRT_EXT_CASE(32);
    default:
        CRT_EXT_CASE(24);
    }
#undef CRT_EXT_CASE
}

// Renders A.spp_count samples per covered pixel with the wavefront pipeline. Synchronises the stream once per
// `wf_batch_iters` iterations to read the active count (8 bytes, pinned).
static crt_status render_wavefront(crt_context* ctx, RenderArgs A)
{
    const uint64_t P = (uint64_t)A.cover_w * A.cover_h;
    if (P == 0 || A.spp_count == 0) return CRT_OK;
    const size_t capacity = (size_t)std::min<uint64_t>(ctx->wf_slots, P * A.spp_count);
    CU(wavefront_reserve(ctx->wf, std::max<size_t>(capacity, 1024)));
    WavefrontState& w = ctx->wf;

    WfArgs W;
    memset(&W, 0, sizeof(W));
    W.r0 = w.r0; W.r1 = w.r1; W.thr = w.thr; W.px = w.px;
    W.c = w.c;
    W.n_pixels_cover = (uint32_t)P;
    W.capacity = (int)capacity;

    bool all = false;
    DevScene ext_scene = A.scene;
    // one CTA of 1024 threads per SM: the whole opt-in shared memory is available to the one scene copy
    long budget = ctx->smem_optin - 1024; // static shared memory of the kernel + allocation granularity
    if (ctx->smem_budget_override >= 0) budget = std::min<long>(budget, ctx->smem_budget_override);
    size_t smem;
    {
        const long full = (long)ext_scene.n_nodes * 16 * CRT_SMEM_NODE_F4 + (long)ext_scene.n_prims * 16;
        if (full <= budget) {
            ext_scene.smem_nodes = ext_scene.n_nodes; ext_scene.smem_prims = ext_scene.n_prims; ext_scene.smem_mats = 0;
            all = true;
            smem = (size_t)full;
        }
        else {
            budget = std::min<long>(budget, ctx->pipe_prefix_bytes); // top of the tree only; the rest of the SM's memory stays L1
            long rem = budget;
            ext_scene.smem_nodes = (int)std::min<long>(ext_scene.n_nodes, rem / (16 * CRT_SMEM_NODE_F4)); rem -= (long)ext_scene.smem_nodes * 16 * CRT_SMEM_NODE_F4;
            ext_scene.smem_prims = (int)std::min<long>(ext_scene.n_prims, rem / 16); rem -= (long)ext_scene.smem_prims * 16;
            ext_scene.smem_mats = 0;
            smem = (size_t)(budget - rem);
        }
    }
    const int ext_grid = ctx->sm_count;
    const int shade_grid = (int)std::min<size_t>((size_t)ctx->sm_count * 8, (capacity + WF_SHADE_THREADS - 1) / WF_SHADE_THREADS);

    // passes: the work counter is 32 bits
    const uint32_t stride = A.spp_stride ? A.spp_stride : 1u;
    const uint32_t spp_per_pass = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(A.spp_count, 0xF0000000ull / P));
    for (uint32_t done = 0; done < A.spp_count; done += spp_per_pass) {
        const uint32_t spp = std::min(spp_per_pass, A.spp_count - done);
        W.R = A;
        W.R.spp_begin = A.spp_begin + done * stride;
        W.R.spp_count = spp;
        W.R.spp_stride = stride;
        const uint32_t work_total = (uint32_t)(P * spp);
        const uint32_t first = (uint32_t)std::min<uint64_t>(capacity, work_total);
        k_wf_generate<<<((uint32_t)capacity + WF_SHADE_THREADS - 1) / WF_SHADE_THREADS, WF_SHADE_THREADS, 0, ctx->stream>>>(W, work_total);
        ctx->launches++;
        CU(cudaGetLastError());
        int iter = 0;
        unsigned long long first_paths = first;
        while (true) {
            for (int it = 0; it < ctx->wf_batch_iters; it++, iter++) {
                W.iter = iter;
                W.R.scene = ext_scene;
                CU(dispatch_extend(W, ext_grid, smem, ctx->stream, ctx->wf_thresh, all));
                W.R.scene = A.scene; // shade reads the tables from global memory
                W.R.scene.smem_nodes = W.R.scene.smem_prims = W.R.scene.smem_mats = 0;
                k_wf_shade<<<shade_grid, WF_SHADE_THREADS, 0, ctx->stream>>>(W);
                CU(cudaGetLastError());
                ctx->launches += 2;
            }
            k_wf_fold_counters<<<1, 1, 0, ctx->stream>>>(w.c, A.counters, first_paths, (iter - 1) & 1, w.d_alive);
            ctx->launches++;
            first_paths = 0;
            CU(cudaMemcpyAsync(w.h_alive, w.d_alive, 2 * sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
            CU(cudaStreamSynchronize(ctx->stream));
            if (w.h_alive[0] == 0u && w.h_alive[1] == 0u) break; // no live path and no unclaimed (pixel, sample) item
        }
        const uint32_t npx = A.cover_w * A.cover_h;
        k_wf_add_count<<<(npx + 255) / 256, 256, 0, ctx->stream>>>(A.accum, A.width, A.cover_w, A.cover_h, (long long)spp, A.tile_rows, A.tile_rank, A.tile_world);
        ctx->launches++;
        CU(cudaGetLastError());
    }
    return CRT_OK;
}

// ------------------------------------------------------------------------------------------
// pipelined persistent megakernel driver
// ------------------------------------------------------------------------------------------
__global__ void k_pipe_init(WfShard* shards, uint32_t work_total)
{
    const uint32_t i = threadIdx.x;
    if (i >= WF_SHARDS) return;
    const uint32_t per = work_total / WF_SHARDS, extra = work_total % WF_SHARDS;
    const uint32_t b = i * per + min(i, extra);
    shards[i].head = b;
    shards[i].end = b + per + (i < extra ? 1u : 0u);
}

template <int NSLOTS, bool ALL, bool SPH> static cudaError_t launch_pipe(const PipeArgs& W, int grid, size_t smem, cudaStream_t st)
{
    static size_t configured[CRT_MAX_DEVICES] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    size_t& configured_smem = configured[dev & (CRT_MAX_DEVICES - 1)];
    if (smem > configured_smem) {
        cudaError_t e = cudaFuncSetAttribute(k_render_pipe<NSLOTS, ALL, SPH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured_smem = smem;
    }
    k_render_pipe<NSLOTS, ALL, SPH><<<grid, PIPE_THREADS, smem, st>>>(W);
    return cudaGetLastError();
}

// the two pool sizes the planner picks have kernels with the size compiled in; anything else (CRT_PIPE_SLOTS) runs the generic one
static cudaError_t dispatch_pipe(const PipeArgs& W, int grid, size_t smem, cudaStream_t st, bool all)
{
    const bool sph = all && W.R.scene.spheres_only; // leaf visits without the type dispatch need the staged primitive table
#define CRT_PIPE_CASE(N) case N: return !all ? launch_pipe<N, false, false>(W, grid, smem, st) : sph ? launch_pipe<N, true, true>(W, grid, smem, st) : launch_pipe<N, true, false>(W, grid, smem, st)
    switch (W.n_slots) {
        CRT_PIPE_CASE(1024);
        CRT_PIPE_CASE(2048);
    default:
        CRT_PIPE_CASE(0);
    }
#undef CRT_PIPE_CASE
}

static crt_status render_pipeline(crt_context* ctx, RenderArgs A)
{
    const uint64_t P = (uint64_t)A.cover_w * A.cover_h;
    if (P == 0 || A.spp_count == 0) return CRT_OK;
    // shared memory: [scene prefix][pool of N slots: 56 B of path state + 7 16-bit ring entries]
    const long per_slot = 3 * 16 + 8 + 2 * PQ_COUNT;
    const long budget = (ctx->smem_budget_override >= 0 ? std::min<long>(ctx->smem_budget_override, ctx->smem_optin) : (long)ctx->smem_optin) - 1024;
    const long node_b = 16 * CRT_PAIR_NODE_F4; // a staged prefix slot of the global path (trace.cuh crt_pair_node)
    // whole scene staged: the PAIR layout of trace.cuh, 48-byte nodes + a dummy node, 16-bit references
    const long full = ((long)A.scene.n_nodes + 1) * 16 * CRT_PAIR_NODE_F4 + (long)A.scene.n_prims * 48;
    const bool pair_ok = A.scene.n_prims <= CRT_PAIR_MAX_PRIMS && CRT_PAIR_NODE_F4 * ((long)A.scene.n_nodes + 1) < CRT_PAIR_LEAF;
    int n_slots = ctx->pipe_slots > 0 ? ctx->pipe_slots : (pair_ok && full + 2048 * per_slot <= budget ? 2048 : 1024);
    while (n_slots & (n_slots - 1)) n_slots &= n_slots - 1; // power of two
    n_slots = std::min(std::max(n_slots, 64), 32768); // ring entries are 16-bit slot ids, 0xffff = empty
    if ((long)n_slots * per_slot > budget) return crt_set_error(ctx, CRT_ERR_INVALID, "render_pipeline: CRT_PIPE_SLOTS does not fit in shared memory");
    bool all = false;
    long rem = budget - (long)n_slots * per_slot;
    long budget_used = budget;
    if (pair_ok && full <= rem) {
        A.scene.smem_nodes = A.scene.n_nodes; A.scene.smem_prims = A.scene.n_prims; A.scene.smem_mats = A.scene.n_prims;
        all = true;
        rem -= full;
    }
    else {
        // The scene does not fit: stage only the top of the tree and leave the rest of the SM's 256 KB to L1, which serves
        // the deeper levels better than a longer prefix (measured: 96 KB of shared memory beats 227 KB by 4 % on the 1M-sphere
        // scene and 7 % on the 10k glass/metal scene).
        rem = std::min<long>(rem, ctx->pipe_prefix_bytes);
        budget_used = (long)n_slots * per_slot + rem;
        // slot 0 of the staged table is the dummy node of node_visit_pairg
        A.scene.smem_nodes = (int)std::max<long>(0, std::min<long>(A.scene.n_nodes, rem / node_b - 1)); rem -= ((long)A.scene.smem_nodes + 1) * node_b;
        A.scene.smem_prims = (int)std::min<long>(A.scene.n_prims, rem / 16); rem -= (long)A.scene.smem_prims * 16;
        A.scene.smem_mats = (int)std::min<long>(A.scene.n_prims, rem / 32); rem -= (long)A.scene.smem_mats * 32;
    }
    const size_t smem = (size_t)(budget_used - rem);
    if (!all) { // the nodes behind the staged prefix are read in the child-interleaved layout
        const size_t nn = (size_t)A.scene.n_nodes;
        if (ctx->cap_pair < nn || !ctx->d_nodes_pair) {
            cudaFree(ctx->d_nodes_pair); ctx->d_nodes_pair = nullptr; ctx->cap_pair = 0;
            CU(cudaMalloc(&ctx->d_nodes_pair, std::max<size_t>(nn, 1) * 64));
            ctx->cap_pair = nn;
            ctx->pair_gen = 0;
        }
        if (ctx->pair_gen != ctx->nodes_gen && nn) {
            k_pair_nodes<<<(unsigned)((nn + 255) / 256), 256, 0, ctx->stream>>>(A.scene.nodes, ctx->d_nodes_pair, (int)nn);
            CU(cudaGetLastError());
            ctx->launches++;
            ctx->pair_gen = ctx->nodes_gen;
        }
        A.scene.nodes_pair = ctx->d_nodes_pair;
    }

    PipeArgs W;
    memset(&W, 0, sizeof(W));
    W.shards = ctx->d_shards;
    W.stat = ctx->d_counters;
    W.n_pixels_cover = (uint32_t)P;
    W.inv_pixels = ~0ull / P + 1ull; // pipe_item: division by multiplication (d == 1 is special-cased there)
    W.inv_cover_w = ~0ull / A.cover_w + 1ull;
    W.n_slots = n_slots;
    // When a warp leaves its traversal loop for a look at the queues: fewer than `thresh` lanes busy, `max_rounds` rounds at the
    // latest. Scene entirely in shared memory: the kernel is issue-bound and the best pair depends on how long a traversal is
    // (profiles/r02_th_refill_sweep.txt, 12..32 x 1..32 on nine scenes, after idle lanes began to keep the rays they shade).
    // RTIOW (nine node visits per ray = three rounds): (20, 4) is +3.7 % over round 1's (14, 16), but costs 4..8 % on glass /
    // metal grids of ANY size and on diffuse grids from 1 000 spheres up, whose traversals are longer than four rounds. (16, 6)
    // is the pair that loses nowhere (+0.8 % RTIOW, +2..3 % on 400..800 spheres, -0.2..-0.7 % on 1 000..1 600) and is the static
    // choice; where crt_render measures (AUTO), it times (20, 4) as well and keeps the faster pair (cal_thresh / cal_rounds).
    // Scene mostly behind L1: latency-bound, idle lanes cost nothing and every refill costs instructions, so a warp runs its batch
    // of rays to completion (THRESH 1: +9 % on the 1 M-sphere scene, +7 % on the 10 k glass/metal one; re-checked with the
    // 44-byte nodes: still the best).
    const bool cal = all && ctx->cal_thresh > 0 && ctx->cal_rounds > 0;
    W.thresh = ctx->pipe_thresh > 0 ? std::min(ctx->pipe_thresh, 32) : cal ? ctx->cal_thresh : (all ? 16 : 1);
    W.max_rounds = ctx->pipe_rounds > 0 ? ctx->pipe_rounds : cal ? ctx->cal_rounds : (all ? 6 : 64);
    const uint32_t stride = A.spp_stride ? A.spp_stride : 1u;
    const uint32_t spp_per_pass = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(A.spp_count, 0xF0000000ull / P));
    for (uint32_t done = 0; done < A.spp_count; done += spp_per_pass) {
        const uint32_t spp = std::min(spp_per_pass, A.spp_count - done);
        W.R = A;
        W.R.spp_begin = A.spp_begin + done * stride;
        W.R.spp_count = spp;
        W.R.spp_stride = stride;
        // Three launches per pass. Folding the two small ones into the render kernel (work counters reset by the last CTA to exit,
        // sample counts added at kernel start) was built and measured in round 2: the C1 step did not move (0.490 vs 0.493 ms: that
        // frame is bound by the fill and drain of the path pool, not by launches), and the __threadfence() the exit ticket needs
        // made ptxas turn every accumulation RED of the kernel into an ATOM with a discarded result: -3.4 % on C2. Dropped.
        k_pipe_init<<<1, WF_SHARDS, 0, ctx->stream>>>(ctx->d_shards, (uint32_t)(P * spp));
        CU(cudaGetLastError());
        CU(dispatch_pipe(W, ctx->sm_count, smem, ctx->stream, all));
        const uint32_t npx = A.cover_w * A.cover_h;
        k_wf_add_count<<<(npx + 255) / 256, 256, 0, ctx->stream>>>(A.accum, A.width, A.cover_w, A.cover_h, (long long)spp, A.tile_rows, A.tile_rank, A.tile_world);
        CU(cudaGetLastError());
        ctx->launches += 3;
    }
    return CRT_OK;
}

// Decodes crt_render_params::row_tiles (CRT_ROW_TILES) and returns how many of the `rows` image rows belong to the shard:
// slabs of tile_rows rows are dealt out round robin, slab t to rank t % world; the last slab may be short.
static uint32_t crt_row_tiles_local_rows(uint32_t row_tiles, uint32_t rows, uint32_t* tile_rows, uint32_t* tile_rank, uint32_t* tile_world,
                                         const char** why)
{
    const uint32_t rank = row_tiles & 0xffu, world = (row_tiles >> 8) & 0xffu, tr = row_tiles >> 16;
    *why = nullptr;
    if (world == 0 || rank >= world || tr == 0) {
        *why = "row_tiles: need 0 <= rank < world <= 255 and rows_per_tile >= 1 (CRT_ROW_TILES)";
        return 0;
    }
    *tile_rows = tr; *tile_rank = rank; *tile_world = world;
    const uint32_t full = rows / tr, rem = rows % tr;
    return (full / world + (rank < full % world ? 1u : 0u)) * tr + (full % world == rank ? rem : 0u);
}

static crt_status render_mega(crt_context* ctx, RenderArgs A, bool exact, bool count)
{
    bool all = false;
    size_t smem = plan_smem(ctx, mega_blocks_per_sm(exact, count), A.scene, all);
    dim3 grid(((A.cover_w + 15) / 16) * ((A.cover_h + 15) / 16));
    const bool pair_ok = A.scene.n_prims <= CRT_PAIR_MAX_PRIMS && CRT_PAIR_NODE_F4 * ((long)A.scene.n_nodes + 1) < CRT_PAIR_LEAF;
    if (all && !exact && !count && pair_ok && !getenv("CRT_MEGA_NO_PAIR")) {
        // whole scene staged: the PAIR layout of trace.cuh (48-byte nodes + a dummy node), never larger than the planned table + 48 B
        smem = ((size_t)A.scene.n_nodes + 1) * 16 * CRT_PAIR_NODE_F4 + (size_t)A.scene.n_prims * 48;
        if (A.scene.spheres_only) {
            CU(cudaFuncSetAttribute(k_render_mega<false, false, true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            k_render_mega<false, false, true, 2><<<grid, CRT_BLOCK_THREADS, smem, ctx->stream>>>(A);
        }
        else {
            CU(cudaFuncSetAttribute(k_render_mega<false, false, true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            k_render_mega<false, false, true, 1><<<grid, CRT_BLOCK_THREADS, smem, ctx->stream>>>(A);
        }
        CU(cudaGetLastError());
    }
    else CU(dispatch_mega(A, grid, smem, ctx->stream, exact, count, all));
    ctx->launches++;
    return CRT_OK;
}

extern "C" {

uint32_t crt_row_tiles_rows(uint32_t row_tiles, uint32_t height)
{
    if (!row_tiles) return height;
    uint32_t tr, rank, world;
    const char* why = nullptr;
    const uint32_t rows = crt_row_tiles_local_rows(row_tiles, height, &tr, &rank, &world, &why);
    return why ? 0u : rows;
}

crt_status crt_render(crt_context* ctx, const crt_render_params* p)
{
    NvtxRange nvtx_range("crt_render");
    if (!ctx || !p) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_render: NULL argument");
    if (!ctx->have_scene) return crt_set_error(ctx, CRT_ERR_STATE, "crt_render: no scene uploaded");
    if (!ctx->have_cam) return crt_set_error(ctx, CRT_ERR_STATE, "crt_render: no camera set");
    if (!ctx->d_accum) return crt_set_error(ctx, CRT_ERR_STATE, "crt_render: no framebuffer (crt_framebuffer_resize)");
    if (p->variant < CRT_VARIANT_AUTO || p->variant > CRT_VARIANT_PIPELINE) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_render: bad variant");
    if (p->jitter != CRT_JITTER_RANDOM && p->jitter != CRT_JITTER_CENTER) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_render: bad jitter mode");
    CU(cudaSetDevice(ctx->device));

    RenderArgs A;
    memset(&A, 0, sizeof(A));
    A.scene = ctx->scene;
    A.cam = ctx->cam;
    A.accum = ctx->d_accum;
    A.counters = ctx->d_counters;
    A.width = ctx->width; A.height = ctx->height;
    A.cover_w = ctx->width; A.cover_h = ctx->height;
    if (p->flags & CRT_FLAG_TRUNCATED_GRID) { A.cover_w = ctx->width / 16 * 16; A.cover_h = ctx->height / 16 * 16; }
    if (p->row_tiles) { // this call renders one row-tile shard: cover_h becomes its local row count
        const char* why = nullptr;
        A.cover_h = crt_row_tiles_local_rows(p->row_tiles, A.cover_h, &A.tile_rows, &A.tile_rank, &A.tile_world, &why);
        if (why) return crt_set_error(ctx, CRT_ERR_INVALID, why);
    }
    A.spp_begin = p->spp_begin; A.spp_count = p->spp_count; A.spp_stride = p->spp_stride ? p->spp_stride : 1;
    A.max_depth = p->max_depth;
    A.seed_lo = (uint32_t)p->seed; A.seed_hi = (uint32_t)(p->seed >> 32);
    crt_philox_round_keys(A.seed_lo, A.seed_hi, A.philox_keys);
    A.jitter = p->jitter;
    const bool exact = (p->flags & CRT_FLAG_EXACT_MATH) != 0, count = (p->flags & CRT_FLAG_COUNT_TESTS) != 0;

    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    if (p->max_depth == 0 && p->spp_count > 0 && A.cover_w > 0 && A.cover_h > 0) {
        // color() with max_depth 0 (Kernel.cu:39: the bounce loop runs zero times; :79 returns black) traces nothing: every
        // sample is black. The samples still count: the frame resolves to opaque black, as the reference's does.
        const uint32_t npx = A.cover_w * A.cover_h;
        k_wf_add_count<<<(npx + 255) / 256, 256, 0, ctx->stream>>>(A.accum, A.width, A.cover_w, A.cover_h, (long long)p->spp_count, A.tile_rows, A.tile_rank, A.tile_world);
        ctx->launches++;
        CU(cudaGetLastError());
        CU(cudaEventRecord(ctx->ev1, ctx->stream));
        ctx->timed = true;
        return CRT_OK;
    }
    if (p->spp_count > 0 && A.cover_w > 0 && A.cover_h > 0) {
        // AUTO, chosen by measurement (DESIGN.md "variant choice"): the pipelined persistent kernel wins as soon as traversal
        // dominates (RTIOW 486 prims: 13.0 vs 11.5 Grays/s; 10k prims: 7.5 vs 5.4); below a few hundred primitives its
        // queue handling costs more than the divergence it removes (67 prims: 16.4 vs 21.1) and the megakernel is used.
        int variant = p->variant;
        ctx->cal_thresh = ctx->cal_rounds = 0; // the static refill pair unless the measurement below says otherwise
        if (variant == CRT_VARIANT_AUTO) {
            variant = (exact || count || ctx->scene.n_prims < 256) ? CRT_VARIANT_MEGAKERNEL : CRT_VARIANT_PIPELINE;
            // Where both kernels hold the whole scene in shared memory the better one depends on the scene AND the view (DESIGN.md
            // 3.1: the pipeline runs at 17-23 Grays/s whatever it is given, the megakernel at 12-64), and every variant writes the
            // same bits. So the first samples of the frame are rendered by both, timed, and the rest goes to the winner; the choice
            // is kept while tree, frame size and depth stay what they are. One host synchronisation per such change.
            const bool both = !exact && !count && ctx->scene.n_prims >= 1 && ctx->scene.n_prims <= CRT_PAIR_MAX_PRIMS;
            cudaStreamCaptureStatus capturing = cudaStreamCaptureStatusNone;
            if (both && ctx->auto_calibrate) cudaStreamIsCapturing(ctx->stream, &capturing); // measuring synchronises: not under graph capture
            if (both && ctx->auto_calibrate && capturing == cudaStreamCaptureStatusNone) {
                uint64_t key = ctx->nodes_gen * 0x9E3779B97F4A7C15ull;
                for (uint64_t v : {(uint64_t)A.cover_w, (uint64_t)A.cover_h, (uint64_t)A.max_depth, (uint64_t)p->row_tiles, (uint64_t)A.jitter})
                    key = (key ^ v) * 0x100000001B3ull;
                key |= 1ull;
                const uint32_t c = std::min<uint32_t>(8u, A.spp_count / 16u); // samples each kernel is timed on
                // ... provided they are a launch like the one that follows: on a small frame a few samples measure the fixed cost of a
                // launch (the pipelined kernel's pool fill and drain), not its rate (400x225: 1 spp 0.14 vs 0.27 ms, 16 spp 1.55 vs 0.46 ms)
                const bool representative = (uint64_t)A.cover_w * A.cover_h * c >= 1500000ull;
                if (ctx->auto_key == key) {
                    variant = ctx->auto_variant;
                    ctx->cal_thresh = ctx->auto_thresh;
                    ctx->cal_rounds = ctx->auto_rounds;
                }
                else if (c >= 1u && representative) {
                    for (int k = 0; k < 4; k++)
                        if (!ctx->ev_cal[k]) CU(cudaEventCreate(&ctx->ev_cal[k]));
                    RenderArgs W = A; // one untimed sample per kernel first (module load, attribute set-up), then c timed samples per candidate
                    uint32_t at = A.spp_begin;
                    auto part = [&](int v, uint32_t n, int thresh, int rounds) -> crt_status {
                        W.spp_begin = at; W.spp_count = n;
                        at += n * A.spp_stride;
                        ctx->cal_thresh = thresh; ctx->cal_rounds = rounds;
                        return v == CRT_VARIANT_MEGAKERNEL ? render_mega(ctx, W, false, false) : render_pipeline(ctx, W);
                    };
                    crt_status cs;
                    if ((cs = part(CRT_VARIANT_MEGAKERNEL, 1, 0, 0)) != CRT_OK || (cs = part(CRT_VARIANT_PIPELINE, 1, 0, 0)) != CRT_OK) return cs;
                    CU(cudaEventRecord(ctx->ev_cal[0], ctx->stream));
                    if ((cs = part(CRT_VARIANT_MEGAKERNEL, c, 0, 0)) != CRT_OK) return cs;
                    CU(cudaEventRecord(ctx->ev_cal[1], ctx->stream));
                    if ((cs = part(CRT_VARIANT_PIPELINE, c, 0, 0)) != CRT_OK) return cs; // the static refill pair
                    CU(cudaEventRecord(ctx->ev_cal[2], ctx->stream));
                    if ((cs = part(CRT_VARIANT_PIPELINE, c, 20, 4)) != CRT_OK) return cs; // the pair of short traversals (render_pipeline)
                    CU(cudaEventRecord(ctx->ev_cal[3], ctx->stream));
                    CU(cudaEventSynchronize(ctx->ev_cal[3]));
                    float t_mega = 0.f, t_pipe = 0.f, t_pipe_short = 0.f;
                    CU(cudaEventElapsedTime(&t_mega, ctx->ev_cal[0], ctx->ev_cal[1]));
                    CU(cudaEventElapsedTime(&t_pipe, ctx->ev_cal[1], ctx->ev_cal[2]));
                    CU(cudaEventElapsedTime(&t_pipe_short, ctx->ev_cal[2], ctx->ev_cal[3]));
                    const bool short_wins = t_pipe_short < t_pipe;
                    variant = std::min(t_pipe, t_pipe_short) < t_mega ? CRT_VARIANT_PIPELINE : CRT_VARIANT_MEGAKERNEL;
                    ctx->auto_key = key;
                    ctx->auto_variant = variant;
                    ctx->auto_thresh = ctx->cal_thresh = short_wins ? 20 : 0;
                    ctx->auto_rounds = ctx->cal_rounds = short_wins ? 4 : 0;
                    if (getenv("CRT_VERBOSE"))
                        fprintf(stderr, "[crt] AUTO: megakernel %.3f ms, pipeline %.3f ms, pipeline (20, 4) %.3f ms for %u samples of %ux%u -> %s%s\n", t_mega, t_pipe,
                                t_pipe_short, c, A.cover_w, A.cover_h, variant == CRT_VARIANT_PIPELINE ? "pipeline" : "megakernel",
                                variant == CRT_VARIANT_PIPELINE && short_wins ? " (20, 4)" : "");
                    A.spp_begin = at;
                    A.spp_count -= 2u + 3u * c;
                }
            }
        }
        ctx->last_variant = variant;
        if (variant == CRT_VARIANT_WAVEFRONT || variant == CRT_VARIANT_PIPELINE) {
            if (exact || count) return crt_set_error(ctx, CRT_ERR_UNSUPPORTED, "crt_render: EXACT_MATH / COUNT_TESTS are megakernel-only");
            crt_status ws = variant == CRT_VARIANT_WAVEFRONT ? render_wavefront(ctx, A) : render_pipeline(ctx, A);
            if (ws != CRT_OK) return ws;
            CU(cudaEventRecord(ctx->ev1, ctx->stream));
            ctx->timed = true;
            return CRT_OK;
        }
        crt_status ms_ = render_mega(ctx, A, exact, count);
        if (ms_ != CRT_OK) return ms_;
    }
    CU(cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->timed = true;
    return CRT_OK;
}

int32_t crt_last_variant(crt_context* ctx)
{
    return ctx ? ctx->last_variant : 0;
}

crt_status crt_last_render_ms(crt_context* ctx, float* out_ms)
{
    if (!ctx || !out_ms) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_last_render_ms: NULL argument");
    if (!ctx->timed) return crt_set_error(ctx, CRT_ERR_STATE, "crt_last_render_ms: nothing rendered yet");
    CU(cudaSetDevice(ctx->device));
    CU(cudaEventSynchronize(ctx->ev1));
    CU(cudaEventElapsedTime(out_ms, ctx->ev0, ctx->ev1));
    return CRT_OK;
}

crt_status crt_resolve_device(crt_context* ctx, void* device_out_rgba8)
{
    NvtxRange nvtx_range("crt_resolve_device");
    if (!ctx || !device_out_rgba8) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_resolve_device: NULL argument");
    if (!ctx->d_accum) return crt_set_error(ctx, CRT_ERR_STATE, "crt_resolve_device: no framebuffer");
    CU(cudaSetDevice(ctx->device));
    uint32_t n = ctx->width * ctx->height;
    k_resolve<<<(n + 255) / 256, 256, 0, ctx->stream>>>(ctx->d_accum, (uint32_t*)device_out_rgba8, ctx->width, ctx->width, ctx->height, 0u, 0u, 0u);
    ctx->launches++;
    CU(cudaGetLastError());
    return CRT_OK;
}

crt_status crt_resolve_device_truncated(crt_context* ctx, void* device_out_rgba8)
{
    NvtxRange nvtx_range("crt_resolve_device_truncated");
    if (!ctx || !device_out_rgba8) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_resolve_device_truncated: NULL argument");
    if (!ctx->d_accum) return crt_set_error(ctx, CRT_ERR_STATE, "crt_resolve_device_truncated: no framebuffer");
    CU(cudaSetDevice(ctx->device));
    const uint32_t cw = ctx->width / 16 * 16, ch = ctx->height / 16 * 16; // Kernel.cu:183-184: grid = (W / 16, H / 16) blocks of 16 x 16
    const uint32_t n = cw * ch;
    if (n == 0) return CRT_OK;
    k_resolve<<<(n + 255) / 256, 256, 0, ctx->stream>>>(ctx->d_accum, (uint32_t*)device_out_rgba8, ctx->width, cw, ch, 0u, 0u, 0u);
    ctx->launches++;
    CU(cudaGetLastError());
    return CRT_OK;
}

crt_status crt_resolve_rows(crt_context* ctx, uint32_t row_tiles, void* device_out_rgba8)
{
    NvtxRange nvtx_range("crt_resolve_rows");
    if (!ctx || !device_out_rgba8) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_resolve_rows: NULL argument");
    if (!ctx->d_accum) return crt_set_error(ctx, CRT_ERR_STATE, "crt_resolve_rows: no framebuffer");
    uint32_t tr = 0, rank = 0, world = 0, rows = ctx->height;
    if (row_tiles) {
        const char* why = nullptr;
        rows = crt_row_tiles_local_rows(row_tiles, ctx->height, &tr, &rank, &world, &why);
        if (why) return crt_set_error(ctx, CRT_ERR_INVALID, why);
    }
    const uint32_t n = ctx->width * rows;
    if (n == 0) return CRT_OK;
    CU(cudaSetDevice(ctx->device));
    k_resolve<<<(n + 255) / 256, 256, 0, ctx->stream>>>(ctx->d_accum, (uint32_t*)device_out_rgba8, ctx->width, ctx->width, rows, tr, rank, world);
    ctx->launches++;
    CU(cudaGetLastError());
    return CRT_OK;
}

crt_status crt_resolve(crt_context* ctx, uint32_t* out_rgba8)
{
    if (!ctx || !out_rgba8) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_resolve: NULL argument");
    crt_status s = crt_resolve_device(ctx, ctx->d_rgba);
    if (s != CRT_OK) return s;
    CU(cudaMemcpyAsync(out_rgba8, ctx->d_rgba, (size_t)ctx->width * ctx->height * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return CRT_OK;
}

crt_status crt_read_rgba(crt_context* ctx, uint32_t* out_rgba8)
{
    NvtxRange nvtx_range("crt_read_rgba");
    if (!ctx || !out_rgba8) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_read_rgba: NULL argument");
    if (!ctx->d_rgba) return crt_set_error(ctx, CRT_ERR_STATE, "crt_read_rgba: no framebuffer");
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpyAsync(out_rgba8, ctx->d_rgba, (size_t)ctx->width * ctx->height * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return CRT_OK;
}

void* crt_rgba_device_ptr(crt_context* ctx)
{
    return ctx ? (void*)ctx->d_rgba : nullptr;
}

// ---- peer memory: the multi-GPU close-out without a reduce-to-root (kernels.cuh: k_resolve_peers) ----
crt_status crt_ipc_export(crt_context* ctx, int32_t which, uint8_t out_handle[CRT_IPC_HANDLE_BYTES])
{
    static_assert(sizeof(cudaIpcMemHandle_t) == CRT_IPC_HANDLE_BYTES, "cudaIpcMemHandle_t is 64 bytes");
    if (!ctx || !out_handle) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_ipc_export: NULL argument");
    void* p = nullptr;
    if (which == CRT_BUFFER_ACCUM) {
        // only the library's own allocation is the base of a cudaMalloc block, which is what an IPC handle names
        if (ctx->d_accum != ctx->d_accum_own)
            return crt_set_error(ctx, CRT_ERR_STATE, "crt_ipc_export: a caller-owned accumulation buffer is set; export that one yourself");
        p = ctx->d_accum_own;
    }
    else if (which == CRT_BUFFER_RGBA) p = ctx->d_rgba;
    else return crt_set_error(ctx, CRT_ERR_INVALID, "crt_ipc_export: which must be CRT_BUFFER_ACCUM or CRT_BUFFER_RGBA");
    if (!p) return crt_set_error(ctx, CRT_ERR_STATE, "crt_ipc_export: no framebuffer (crt_framebuffer_resize)");
    CU(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, p));
    memcpy(out_handle, &h, CRT_IPC_HANDLE_BYTES);
    return CRT_OK;
}

crt_status crt_ipc_open(crt_context* ctx, const uint8_t handle[CRT_IPC_HANDLE_BYTES], void** out_device_ptr)
{
    if (!ctx || !handle || !out_device_ptr) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_ipc_open: NULL argument");
    *out_device_ptr = nullptr;
    CU(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, CRT_IPC_HANDLE_BYTES);
    void* p = nullptr;
    CU(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess)); // maps the exporter's block; enables peer access if needed
    ctx->ipc_open.push_back(p);
    *out_device_ptr = p;
    return CRT_OK;
}

crt_status crt_ipc_close(crt_context* ctx, void* device_ptr)
{
    if (!ctx || !device_ptr) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_ipc_close: NULL argument");
    for (size_t i = 0; i < ctx->ipc_open.size(); i++) {
        if (ctx->ipc_open[i] != device_ptr) continue;
        CU(cudaSetDevice(ctx->device));
        CU(cudaStreamSynchronize(ctx->stream));
        ctx->ipc_open.erase(ctx->ipc_open.begin() + (long)i);
        CU(cudaIpcCloseMemHandle(device_ptr));
        return CRT_OK;
    }
    return crt_set_error(ctx, CRT_ERR_INVALID, "crt_ipc_close: not a pointer returned by crt_ipc_open on this context");
}

crt_status crt_peer_enable(crt_context* ctx, int32_t peer_device)
{
    if (!ctx) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_peer_enable: ctx is NULL");
    if (peer_device == ctx->device) return CRT_OK;
    CU(cudaSetDevice(ctx->device));
    int can = 0;
    CU(cudaDeviceCanAccessPeer(&can, ctx->device, peer_device));
    if (!can) return crt_set_error(ctx, CRT_ERR_UNSUPPORTED, "crt_peer_enable: this device cannot address the memory of peer_device");
    const cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return CRT_OK; } // idempotent
    if (e != cudaSuccess) return cuda_fail(ctx, e, "cudaDeviceEnablePeerAccess");
    return CRT_OK;
}

crt_status crt_resolve_peers(crt_context* ctx, const void* const* peer_accum, int32_t n_peers, uint32_t row_begin, uint32_t row_end,
                             void* device_out_rgba8, void* device_out_sum)
{
    NvtxRange nvtx_range("crt_resolve_peers (close-out over peer memory)");
    if (!ctx || !peer_accum || !device_out_rgba8) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_resolve_peers: NULL argument");
    if (!ctx->d_accum) return crt_set_error(ctx, CRT_ERR_STATE, "crt_resolve_peers: no framebuffer");
    if (n_peers < 1 || n_peers > CRT_MAX_PEERS) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_resolve_peers: 1 <= n_peers <= 16");
    if (row_begin > row_end || row_end > ctx->height) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_resolve_peers: bad row range");
    PeerAccum P;
    memset(&P, 0, sizeof(P));
    P.n = n_peers;
    for (int g = 0; g < n_peers; g++) {
        if (!peer_accum[g] || ((uintptr_t)peer_accum[g] & 31u) != 0)
            return crt_set_error(ctx, CRT_ERR_INVALID, "crt_resolve_peers: every peer buffer must be a 32-byte aligned device pointer");
        P.p[g] = (const long long*)peer_accum[g];
    }
    if (((uintptr_t)device_out_sum & 31u) != 0) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_resolve_peers: the sum buffer must be 32-byte aligned");
    const uint32_t n = (row_end - row_begin) * ctx->width;
    if (n == 0) return CRT_OK;
    CU(cudaSetDevice(ctx->device));
    const dim3 grid((n + 255) / 256);
    uint32_t* out = (uint32_t*)device_out_rgba8;
    long long* sum = (long long*)device_out_sum;
    const uint32_t first = row_begin * ctx->width;
    if (n_peers <= 2) k_resolve_peers<2><<<grid, 256, 0, ctx->stream>>>(P, out, sum, first, n);
    else if (n_peers <= 4) k_resolve_peers<4><<<grid, 256, 0, ctx->stream>>>(P, out, sum, first, n);
    else if (n_peers <= 8) k_resolve_peers<8><<<grid, 256, 0, ctx->stream>>>(P, out, sum, first, n);
    else k_resolve_peers<CRT_MAX_PEERS><<<grid, 256, 0, ctx->stream>>>(P, out, sum, first, n);
    ctx->launches++;
    CU(cudaGetLastError());
    return CRT_OK;
}

crt_status crt_read_accum(crt_context* ctx, float* out_float4)
{
    if (!ctx || !out_float4) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_read_accum: NULL argument");
    if (!ctx->d_accum) return crt_set_error(ctx, CRT_ERR_STATE, "crt_read_accum: no framebuffer");
    CU(cudaSetDevice(ctx->device));
    const uint32_t n = ctx->width * ctx->height;
    k_accum_to_float<<<(n + 255) / 256, 256, 0, ctx->stream>>>(ctx->d_accum, ctx->d_accum_f, n);
    ctx->launches++;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(out_float4, ctx->d_accum_f, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return CRT_OK;
}

crt_status crt_primary_ids(crt_context* ctx, int32_t jitter, uint64_t seed, int32_t* out_ids)
{
    if (!ctx || !out_ids) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_primary_ids: NULL argument");
    if (!ctx->have_scene) return crt_set_error(ctx, CRT_ERR_STATE, "crt_primary_ids: no scene uploaded");
    if (!ctx->have_cam) return crt_set_error(ctx, CRT_ERR_STATE, "crt_primary_ids: no camera set");
    if (!ctx->d_ids) return crt_set_error(ctx, CRT_ERR_STATE, "crt_primary_ids: no framebuffer (crt_framebuffer_resize)");
    if (jitter != CRT_JITTER_RANDOM && jitter != CRT_JITTER_CENTER) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_primary_ids: bad jitter mode");
    CU(cudaSetDevice(ctx->device));
    IdArgs A;
    memset(&A, 0, sizeof(A));
    A.scene = ctx->scene;
    A.cam = ctx->cam;
    A.out_ids = ctx->d_ids;
    A.width = ctx->width; A.height = ctx->height;
    A.seed_lo = (uint32_t)seed; A.seed_hi = (uint32_t)(seed >> 32);
    A.jitter = jitter;
    bool all = false;
    size_t smem = plan_smem(ctx, 4, A.scene, all);
    dim3 grid(((A.width + 15) / 16) * ((A.height + 15) / 16));
    if (all) {
        CU(cudaFuncSetAttribute(k_primary_ids<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_primary_ids<true><<<grid, CRT_BLOCK_THREADS, smem, ctx->stream>>>(A);
    }
    else {
        CU(cudaFuncSetAttribute(k_primary_ids<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_primary_ids<false><<<grid, CRT_BLOCK_THREADS, smem, ctx->stream>>>(A);
    }
    ctx->launches++;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(out_ids, ctx->d_ids, (size_t)ctx->width * ctx->height * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return CRT_OK;
}

crt_status crt_get_counters(crt_context* ctx, crt_counters* out)
{
    if (!ctx || !out) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_get_counters: NULL argument");
    CU(cudaSetDevice(ctx->device));
    static_assert(sizeof(DevCounters) == 128, "one counter shard per 128 bytes");
    std::vector<DevCounters> h(CRT_COUNTER_SHARDS);
    CU(cudaMemcpyAsync(h.data(), ctx->d_counters, h.size() * sizeof(DevCounters), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    out->paths = out->rays = out->box_tests = out->prim_tests = 0;
    for (const DevCounters& c : h) { out->paths += c.paths; out->rays += c.rays; out->box_tests += c.box_tests; out->prim_tests += c.prim_tests; }
    out->kernel_launches = ctx->launches;
    return CRT_OK;
}

crt_status crt_reset_counters(crt_context* ctx)
{
    if (!ctx) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_reset_counters: ctx is NULL");
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemsetAsync(ctx->d_counters, 0, CRT_COUNTER_SHARDS * sizeof(DevCounters), ctx->stream));
    ctx->launches = 0;
    return CRT_OK;
}

crt_status crt_write_ppm(const char* path, const uint32_t* rgba8, uint32_t width, uint32_t height, int flip_rows)
{
    if (!path || !rgba8 || !width || !height) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_write_ppm: bad argument");
    FILE* f = fopen(path, "wb");
    if (!f) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_write_ppm: cannot open output file");
    fprintf(f, "P6\n%u %u\n255\n", width, height);
    std::vector<unsigned char> row((size_t)width * 3);
    for (uint32_t r = 0; r < height; r++) {
        const uint32_t* src = rgba8 + (size_t)(flip_rows ? height - 1 - r : r) * width;
        for (uint32_t x = 0; x < width; x++) {
            row[3 * x] = src[x] & 255; row[3 * x + 1] = (src[x] >> 8) & 255; row[3 * x + 2] = (src[x] >> 16) & 255;
        }
        fwrite(row.data(), 1, row.size(), f);
    }
    fclose(f);
    return CRT_OK;
}

// PNG (ISO/IEC 15948), 8-bit RGB, filter 0 on every row, the zlib stream made of STORED deflate blocks: readable by every
// decoder, no compressor needed. The reference links stb_image_write for this (Utils/RawStbImage.h:8-9) and never calls it.
crt_status crt_write_png(const char* path, const uint32_t* rgba8, uint32_t width, uint32_t height, int flip_rows)
{
    if (!path || !rgba8 || !width || !height) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_write_png: bad argument");
    static uint32_t crc_table[256];
    static std::once_flag once;
    std::call_once(once, [] {
        for (uint32_t n = 0; n < 256; n++) {
            uint32_t c = n;
            for (int k = 0; k < 8; k++) c = (c & 1u) ? 0xedb88320u ^ (c >> 1) : c >> 1;
            crc_table[n] = c;
        }
    });
    auto be32 = [](unsigned char* p, uint32_t v) { p[0] = (unsigned char)(v >> 24); p[1] = (unsigned char)(v >> 16); p[2] = (unsigned char)(v >> 8); p[3] = (unsigned char)v; };
    // raw scanlines: filter byte + RGB
    const size_t stride = 1 + (size_t)width * 3;
    std::vector<unsigned char> raw(stride * height);
    for (uint32_t r = 0; r < height; r++) {
        const uint32_t* src = rgba8 + (size_t)(flip_rows ? height - 1 - r : r) * width;
        unsigned char* row = raw.data() + stride * r;
        row[0] = 0;
        for (uint32_t x = 0; x < width; x++) {
            row[1 + 3 * x] = src[x] & 255; row[2 + 3 * x] = (src[x] >> 8) & 255; row[3 + 3 * x] = (src[x] >> 16) & 255;
        }
    }
    // zlib: header, stored blocks of at most 65535 bytes, Adler-32 of the raw data
    std::vector<unsigned char> z;
    z.reserve(raw.size() + raw.size() / 65535 * 5 + 16);
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, b = 0;
    for (size_t off = 0; off < raw.size();) {
        const size_t n = std::min<size_t>(65535, raw.size() - off);
        z.push_back(off + n == raw.size() ? 1 : 0);
        z.push_back((unsigned char)(n & 255)); z.push_back((unsigned char)(n >> 8));
        z.push_back((unsigned char)(~n & 255)); z.push_back((unsigned char)((~n >> 8) & 255));
        z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
        for (size_t i = 0; i < n;) { // Adler-32 in runs short enough not to overflow 32 bits
            const size_t run = std::min<size_t>(5552, n - i);
            for (size_t k = 0; k < run; k++) { a += raw[off + i + k]; b += a; }
            a %= 65521u; b %= 65521u;
            i += run;
        }
        off += n;
    }
    unsigned char ad[4];
    be32(ad, b << 16 | a);
    z.insert(z.end(), ad, ad + 4);
    FILE* f = fopen(path, "wb");
    if (!f) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_write_png: cannot open output file");
    auto chunk = [&](const char type[4], const unsigned char* data, size_t n) {
        unsigned char head[8];
        be32(head, (uint32_t)n);
        memcpy(head + 4, type, 4);
        uint32_t c = 0xffffffffu;
        for (int i = 4; i < 8; i++) c = crc_table[(c ^ head[i]) & 255] ^ (c >> 8);
        for (size_t i = 0; i < n; i++) c = crc_table[(c ^ data[i]) & 255] ^ (c >> 8);
        unsigned char tail[4];
        be32(tail, c ^ 0xffffffffu);
        fwrite(head, 1, 8, f);
        if (n) fwrite(data, 1, n, f);
        fwrite(tail, 1, 4, f);
    };
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    fwrite(sig, 1, 8, f);
    unsigned char ihdr[13];
    be32(ihdr, width); be32(ihdr + 4, height);
    ihdr[8] = 8; ihdr[9] = 2; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0; // 8 bit, truecolour, deflate, adaptive filtering, no interlace
    chunk("IHDR", ihdr, 13);
    chunk("IDAT", z.data(), z.size());
    chunk("IEND", nullptr, 0);
    const bool ok = ferror(f) == 0;
    fclose(f);
    return ok ? CRT_OK : crt_set_error(nullptr, CRT_ERR_INVALID, "crt_write_png: write error");
}

} // extern "C"

// ------------------------------------------------------------------------------------------
// crt_read_accum_static: crt_read_accum for callers that cannot allocate (code under CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc). No counterpart in the reference.
//
// crt_read_accum converts the fixed-point buffer into a staging array that crt_framebuffer_resize cudaMalloc'ed (d_accum_f) and
// copies that to the host. Here the staging array is a file-scope __device__ array of CRT_STATIC_ACCUM_MAX_PIXELS float4
// (4 MiB; every GPU that loads libcrt.so holds it until the process exits), and the call uses no cudaMalloc, cudaFree or
// allocating library routine.
//   * In HOST code the name g_static_accum_stage is a host-side stand-in for the symbol, NOT a device address: a kernel given
//     it faults, or - where the GPU can reach host memory - silently reads and writes the wrong memory. Its device address is
//     asked for once per device, with that device current, by cudaGetSymbolAddress, and that pointer is what the kernel gets.
//   * Every call shares the one array, and the array keeps the last call's contents. Nothing in it has to start at zero: the
//     kernel overwrites entries [0, n_pixels) before the copy reads exactly that range. Two calls that OVERLAP on one GPU
//     (different streams, or different host threads) race on the array; calls on one stream are ordered and safe.
// ------------------------------------------------------------------------------------------
static __device__ __align__(256) float4 g_static_accum_stage[CRT_STATIC_ACCUM_MAX_PIXELS];

// the variant of k_accum_to_float (kernels.cuh) that writes into the static staging array
__global__ void k_accum_to_float_static(const long long* __restrict__ accum, float4* __restrict__ stage, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const longlong2* px = reinterpret_cast<const longlong2*>(accum + 4 * (size_t)i);
    const longlong2 rg = px[0], bn = px[1];
    stage[i] = make_float4(crt_from_fixed(rg.x), crt_from_fixed(rg.y), crt_from_fixed(bn.x), (float)bn.y);
}

// Device address of the staging array on `device` (which must be current): looked up once per device.
static cudaError_t static_accum_stage_ptr(int device, float4** out)
{
    static std::mutex mu;
    static float4* ptr[CRT_MAX_DEVICES] = {};
    if (device < 0 || device >= CRT_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[device]) {
        float4* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_static_accum_stage);
        if (e != cudaSuccess) return e;
        ptr[device] = p;
    }
    *out = ptr[device];
    return cudaSuccess;
}

extern "C" crt_status crt_read_accum_static(crt_context* ctx, const void* device_accum_int64x4, uint32_t n_pixels, float* out_float4)
{
    if (!ctx || !device_accum_int64x4 || !out_float4) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_read_accum_static: NULL argument");
    if (n_pixels > (uint32_t)CRT_STATIC_ACCUM_MAX_PIXELS)
        return crt_set_error(ctx, CRT_ERR_INVALID, "crt_read_accum_static: n_pixels exceeds CRT_STATIC_ACCUM_MAX_PIXELS (262144), the size of the "
                                                  "static staging array; use crt_read_accum, or read the buffer in pieces");
    if (n_pixels == 0) return CRT_OK;
    if (((uintptr_t)device_accum_int64x4 & 15u) != 0) return crt_set_error(ctx, CRT_ERR_INVALID, "crt_read_accum_static: the buffer must be 16-byte aligned");
    CU(cudaSetDevice(ctx->device));
    float4* stage = nullptr;
    CU(static_accum_stage_ptr(ctx->device, &stage)); // NOT g_static_accum_stage itself: that name is a host-side stand-in here
    k_accum_to_float_static<<<(n_pixels + 255) / 256, 256, 0, ctx->stream>>>((const long long*)device_accum_int64x4, stage, n_pixels);
    ctx->launches++;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(out_float4, stage, (size_t)n_pixels * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return CRT_OK;
}
