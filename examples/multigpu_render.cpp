/*
 * multigpu_render.cpp — native multi-GPU host over the C ABI (include/crt.h), no Python: ONE process, one host thread and
 * one context per GPU. The reference has no multi-GPU path; this is what a C++ host like CudaLayer would do with the library on
 * an NVSwitch box (bench.py does the same with one PROCESS per GPU, cudaraytracer_b200/multigpu.py).
 *
 *   multigpu_render [n_gpus] [mode: tiles|spp|nccl] [scene 1..5] [width] [height] [spp] [max_depth] [out.png]
 *
 * nccl:  north_star's close-out, from C++: GPU g renders samples g, g + G, ... of every pixel into its int64 accumulation
 *        buffer; ONE ncclReduce(ncclInt64, ncclSum) brings the G buffers to GPU 0, which resolves. Integer sums are exact and
 *        order-free, so the frame has the bits of the one-GPU frame. (NCCL comes from the system libnccl; only this example
 *        links it, libcrt.so does not.)
 * tiles: GPU g renders ALL samples of the 16-row slabs g, g + G, ... (crt_render_params::row_tiles) and resolves them straight
 *        into GPU 0's frame (crt_resolve_rows with a peer pointer): nothing to sum, 4 bytes per pixel over NVLink.
 * spp:   GPU g renders samples g, g + G, ... of every pixel; then every GPU sums, for its block of rows, the buffers of all
 *        GPUs through peer pointers and resolves them into GPU 0's frame (crt_resolve_peers): exact (int64), the same bits.
 * Either way the frame is bit-identical to what one GPU renders (Philox is keyed by pixel / sample / bounce).
 */
#include "crt.h"

#include <cuda_runtime.h>
#include <nccl.h>

#include <atomic>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

struct Barrier // std::barrier without <barrier>: every GPU thread arrives, then all leave
{
    explicit Barrier(int n) : n_(n) {}
    void wait()
    {
        std::unique_lock<std::mutex> lk(mu_);
        const int gen = gen_;
        if (++count_ == n_) { count_ = 0; gen_++; cv_.notify_all(); }
        else cv_.wait(lk, [&] { return gen != gen_; });
    }
    std::mutex mu_;
    std::condition_variable cv_;
    int n_, count_ = 0, gen_ = 0;
};

struct Shared
{
    int G, kind;
    bool tiles, nccl;
    std::vector<ncclComm_t> comm;
    std::vector<cudaStream_t> stream;
    uint32_t W, H, spp, depth;
    std::vector<crt_context*> ctx;
    std::vector<void*> accum, rgba; // device pointers of every context's own buffers
    std::atomic<int> failed{0};
    std::string first_error;
    std::mutex err_mu;
    Barrier* bar;
};

bool ok(Shared& S, int g, crt_status st, const char* what)
{
    if (st == CRT_OK) return true;
    std::lock_guard<std::mutex> lk(S.err_mu);
    if (!S.failed.exchange(1)) S.first_error = std::string("GPU ") + std::to_string(g) + ": " + what + ": " + crt_last_error_string(S.ctx[g]);
    return false;
}

void gpu_thread(Shared& S, int g)
{
#define STEP(call) (S.failed.load() || ok(S, g, (call), #call))
    crt_context* ctx = nullptr;
    crt_scene_builder* sb = nullptr;
    crt_input_struct cam;
    STEP(crt_context_create(g, &ctx));
    S.ctx[g] = ctx;
    if (S.nccl && ctx) { // the render kernels and the reduce share one stream per GPU: stream order is the only synchronisation
        cudaSetDevice(g);
        if (cudaStreamCreateWithFlags(&S.stream[g], cudaStreamNonBlocking) != cudaSuccess) ok(S, g, CRT_ERR_CUDA, "cudaStreamCreate");
        else STEP(crt_set_stream(ctx, S.stream[g]));
    }
    STEP(crt_scene_generate(S.kind, 1984, 0, &sb));
    if (sb) {
        STEP(crt_scene_upload(ctx, crt_scene_builder_desc(sb))); // the scene is replicated: every GPU builds / uploads its copy
        STEP(crt_scene_builder_camera(sb, S.W, S.H, &cam));
        STEP(crt_set_camera(ctx, &cam));
    }
    STEP(crt_framebuffer_resize(ctx, S.W, S.H));
    if (ctx) { S.accum[g] = crt_accum_device_ptr(ctx); S.rgba[g] = crt_rgba_device_ptr(ctx); }
    S.bar->wait(); // every context exists, every pointer is published
    for (int o = 0; o < S.G && !S.failed.load() && !S.nccl; o++)
        if (o != g && (o == 0 || !S.tiles)) STEP(crt_peer_enable(ctx, o)); // tiles: only GPU 0's frame is addressed remotely

    crt_render_params p;
    memset(&p, 0, sizeof(p));
    p.max_depth = S.depth; p.seed = 1984; p.variant = CRT_VARIANT_AUTO; p.jitter = CRT_JITTER_RANDOM;
    if (S.nccl) {
        p.spp_begin = (uint32_t)g; p.spp_stride = (uint32_t)S.G;
        p.spp_count = S.spp > (uint32_t)g ? (S.spp - (uint32_t)g + (uint32_t)S.G - 1) / (uint32_t)S.G : 0;
        STEP(crt_render(ctx, &p));
        // the single NCCL reduce of north_star: G x (W x H x 4) int64 -> GPU 0, in place there, on the render stream
        if (!S.failed.load()) {
            const size_t count = (size_t)S.W * S.H * 4;
            cudaSetDevice(g);
            const ncclResult_t nr = ncclReduce(S.accum[g], S.accum[0], count, ncclInt64, ncclSum, 0, S.comm[g], S.stream[g]);
            if (nr != ncclSuccess) { ok(S, g, CRT_ERR_CUDA, ncclGetErrorString(nr)); }
        }
        if (g == 0) STEP(crt_resolve_device(ctx, S.rgba[0]));
        STEP(crt_synchronize(ctx));
        S.bar->wait();
    }
    else if (S.tiles) {
        p.spp_begin = 0; p.spp_count = S.spp; p.spp_stride = 1;
        p.row_tiles = CRT_ROW_TILES(g, S.G, 16);
        STEP(crt_render(ctx, &p));
        STEP(crt_resolve_rows(ctx, p.row_tiles, S.rgba[0]));
        STEP(crt_synchronize(ctx));
        S.bar->wait(); // every slab has landed in GPU 0's frame
    }
    else {
        p.spp_begin = (uint32_t)g; p.spp_stride = (uint32_t)S.G;
        p.spp_count = S.spp > (uint32_t)g ? (S.spp - (uint32_t)g + (uint32_t)S.G - 1) / (uint32_t)S.G : 0;
        STEP(crt_render(ctx, &p));
        STEP(crt_synchronize(ctx));
        S.bar->wait(); // every GPU's samples are in its buffer
        const uint32_t r0 = (uint32_t)((uint64_t)g * S.H / S.G), r1 = (uint32_t)((uint64_t)(g + 1) * S.H / S.G);
        STEP(crt_resolve_peers(ctx, S.accum.data(), S.G, r0, r1, S.rgba[0], nullptr));
        STEP(crt_synchronize(ctx));
        S.bar->wait(); // every block of rows has landed in GPU 0's frame
    }
    if (sb) crt_scene_builder_free(sb);
#undef STEP
}

} // namespace

int main(int argc, char** argv)
{
    Shared S;
    S.G = argc > 1 ? atoi(argv[1]) : 1;
    S.nccl = argc > 2 && !strcmp(argv[2], "nccl");
    S.tiles = !S.nccl && !(argc > 2 && !strcmp(argv[2], "spp"));
    S.kind = argc > 3 ? atoi(argv[3]) : CRT_SCENE_RTIOW;
    S.W = argc > 4 ? (uint32_t)atoi(argv[4]) : 400; S.H = argc > 5 ? (uint32_t)atoi(argv[5]) : 225;
    S.spp = argc > 6 ? (uint32_t)atoi(argv[6]) : 16; S.depth = argc > 7 ? (uint32_t)atoi(argv[7]) : 50;
    const char* out = argc > 8 ? argv[8] : "frame.png";
    if (S.G < 1 || S.G > 16) { fprintf(stderr, "n_gpus must be 1..16\n"); return 2; }
    S.ctx.assign((size_t)S.G, nullptr); S.accum.assign((size_t)S.G, nullptr); S.rgba.assign((size_t)S.G, nullptr);
    S.comm.assign((size_t)S.G, nullptr); S.stream.assign((size_t)S.G, nullptr);
    if (S.nccl) { // one communicator per GPU of this process
        std::vector<int> devs((size_t)S.G);
        for (int g = 0; g < S.G; g++) devs[(size_t)g] = g;
        const ncclResult_t nr = ncclCommInitAll(S.comm.data(), S.G, devs.data());
        if (nr != ncclSuccess) { fprintf(stderr, "ncclCommInitAll: %s\n", ncclGetErrorString(nr)); return 1; }
    }
    Barrier bar(S.G);
    S.bar = &bar;
    std::vector<std::thread> pool;
    for (int g = 1; g < S.G; g++) pool.emplace_back(gpu_thread, std::ref(S), g);
    gpu_thread(S, 0);
    for (std::thread& t : pool) t.join();
    int rc = 0;
    if (S.failed.load()) { fprintf(stderr, "%s\n", S.first_error.c_str()); rc = 1; }
    else {
        std::vector<uint32_t> frame((size_t)S.W * S.H);
        unsigned long long rays = 0;
        for (int g = 0; g < S.G; g++) { crt_counters c; if (crt_get_counters(S.ctx[g], &c) == CRT_OK) rays += c.rays; }
        if (crt_read_rgba(S.ctx[0], frame.data()) != CRT_OK || crt_write_png(out, frame.data(), S.W, S.H, 1) != CRT_OK) {
            fprintf(stderr, "%s\n", crt_last_error_string(S.ctx[0]));
            rc = 1;
        }
        else printf("%d GPU(s), %s, %ux%u, %u spp, depth %u: %llu rays -> %s\n", S.G,
                    S.nccl ? "spp sharding + one ncclReduce(int64, sum)" : S.tiles ? "tile sharding" : "spp sharding + peer-memory close-out", S.W, S.H,
                    S.spp, S.depth, rays, out);
    }
    for (crt_context* c : S.ctx) crt_context_destroy(c);
    for (int g = 0; g < S.G; g++) {
        if (S.comm[(size_t)g]) ncclCommDestroy(S.comm[(size_t)g]);
        if (S.stream[(size_t)g]) { cudaSetDevice(g); cudaStreamDestroy(S.stream[(size_t)g]); }
    }
    return rc;
}
