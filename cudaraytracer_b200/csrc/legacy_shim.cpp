// legacy_shim.cpp — LaunchKernel / LaunchRandInit / LaunchRenderInit (Kernel.cu:178-204) on top of the
// native C ABI. See include/crt_legacy.h for the contract.

#include "crt_legacy.h"
#include "crt_internal.h"

#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

namespace {

struct Legacy
{
    crt_context* ctx = nullptr;
    int device = -1;
    uint64_t scene_hash = 0;
    bool have_scene = false;
    uint32_t width = 0, height = 0;
    uint32_t frame = 0; // advances the Philox sample index so successive frames carry fresh noise,
                        // as the reference's persistent per-pixel XORWOW state does (Kernel.cu:123,149)
    uint32_t next_sample = 0;
};

Legacy g;
std::mutex g_mu;

// Change detector of the flattened scene, run on every LaunchKernel call: eight bytes per multiply (a byte-wise FNV is one
// dependent multiply per byte, ~1 GB/s - 6 ms per frame on a 60 k-sphere scene, more than the frame itself at 1 spp).
uint64_t fnv(uint64_t h, const void* p, size_t n)
{
    const unsigned char* b = (const unsigned char*)p;
    size_t i = 0;
    for (; i + 8 <= n; i += 8) {
        uint64_t w;
        memcpy(&w, b + i, 8);
        h = (h ^ w) * 0x9E3779B97F4A7C15ull;
        h ^= h >> 29;
    }
    for (; i < n; i++) { h ^= b[i]; h *= 0x100000001B3ull; }
    return h;
}

uint64_t hash_flat(const CrtFlatScene& f)
{
    uint64_t h = 0xCBF29CE484222325ull;
    h = fnv(h, f.hittables.data(), f.hittables.size() * sizeof(crt_hittable));
    h = fnv(h, f.materials.data(), f.materials.size() * sizeof(crt_material));
    h = fnv(h, f.textures.data(), f.textures.size() * sizeof(crt_texture));
    for (const crt_image& im : f.images) {
        h = fnv(h, &im.pixels, sizeof(im.pixels));
        h = fnv(h, &im.width, sizeof(int32_t) * 2);
        // sample the texel buffer sparsely: a reloaded image at the same address is still noticed
        if (im.pixels && im.width > 0 && im.height > 0) {
            size_t n = (size_t)im.width * im.height * 3, step = n / 64 + 1;
            for (size_t k = 0; k < n; k += step) h = fnv(h, im.pixels + k, 1);
        }
    }
    return h;
}

bool ensure_context()
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
    if (g.ctx && g.device == dev) return true;
    if (g.ctx) { crt_context_destroy(g.ctx); g = Legacy(); }
    if (crt_context_create(dev, &g.ctx) != CRT_OK) {
        fprintf(stderr, "[crt] LaunchKernel: %s\n", crt_last_error_string(nullptr));
        g.ctx = nullptr;
        return false;
    }
    g.device = dev;
    return true;
}

} // namespace

extern "C" {

void LaunchKernel(unsigned int* pos, unsigned int image_width, unsigned int image_height, const unsigned int samples_per_pixel,
                  const unsigned int max_depth, void* world, void* /*d_rand_state*/, crt_input_struct inputs)
{
    std::lock_guard<std::mutex> lk(g_mu);
    if (!pos || !world || image_width == 0 || image_height == 0) {
        fprintf(stderr, "[crt] LaunchKernel: bad argument\n");
        return;
    }
    if (!ensure_context()) return;
    crt_context* ctx = g.ctx;

    // The caller mutates the scene through the same unified-memory pointers between launches
    // (CudaLayer.cpp:493-494 etc.): re-flatten every frame (cheap pointer walk) and rebuild only on change.
    CrtFlatScene flat;
    std::string err;
    if (!crt_flatten_reference_world(world, &flat, &err)) {
        fprintf(stderr, "[crt] LaunchKernel: %s\n", err.c_str());
        return;
    }
    uint64_t h = hash_flat(flat);
    if (!g.have_scene || h != g.scene_hash) {
        crt_scene_desc d = flat.desc();
        if (crt_scene_upload(ctx, &d) != CRT_OK) {
            fprintf(stderr, "[crt] LaunchKernel: %s\n", crt_last_error_string(ctx));
            return;
        }
        g.scene_hash = h;
        g.have_scene = true;
    }
    if (g.width != image_width || g.height != image_height) {
        if (crt_framebuffer_resize(ctx, image_width, image_height) != CRT_OK) {
            fprintf(stderr, "[crt] LaunchKernel: %s\n", crt_last_error_string(ctx));
            return;
        }
        g.width = image_width;
        g.height = image_height;
    }
    else {
        crt_accum_clear(ctx); // the reference has no accumulation across frames (Kernel.cu:128)
    }
    crt_set_camera(ctx, &inputs);

    crt_render_params p;
    memset(&p, 0, sizeof(p));
    p.spp_begin = g.next_sample;
    p.spp_count = samples_per_pixel;
    p.spp_stride = 1;
    p.max_depth = max_depth;
    p.seed = 1984; // Kernel.cu:175
    p.variant = CRT_VARIANT_AUTO;
    p.jitter = CRT_JITTER_RANDOM;
    const char* trunc = getenv("CRT_LEGACY_TRUNCATED_GRID");
    p.flags = (trunc && atoi(trunc)) ? CRT_FLAG_TRUNCATED_GRID : 0;
    // truncated grid: the pixels the reference's launch does not cover keep what the caller's buffer held
    if (crt_render(ctx, &p) != CRT_OK || (p.flags ? crt_resolve_device_truncated(ctx, pos) : crt_resolve_device(ctx, pos)) != CRT_OK || crt_synchronize(ctx) != CRT_OK) {
        fprintf(stderr, "[crt] LaunchKernel: %s\n", crt_last_error_string(ctx));
        return;
    }
    g.next_sample += samples_per_pixel;
    g.frame++;
}

void LaunchRandInit(void* /*d_rand_state2*/)
{
    // RandInit seeds one state that no kernel reads (CudaLayer.cpp:374-375 never passes it on).
}

void LaunchRenderInit(crt_dim3 /*grid*/, crt_dim3 /*block*/, unsigned int /*window_width*/, unsigned int /*window_height*/,
                      void* /*d_rand_state*/)
{
    // RenderInit re-seeds every pixel with 1984 + pixel_index (Kernel.cu:175): restart the sample counter.
    std::lock_guard<std::mutex> lk(g_mu);
    g.next_sample = 0;
    g.frame = 0;
}

crt_status crt_legacy_counters(crt_counters* out)
{
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g.ctx) return crt_set_error(nullptr, CRT_ERR_STATE, "crt_legacy_counters: LaunchKernel has not run");
    return crt_get_counters(g.ctx, out);
}

void crt_legacy_shutdown(void)
{
    std::lock_guard<std::mutex> lk(g_mu);
    if (g.ctx) crt_context_destroy(g.ctx);
    g = Legacy();
}

} // extern "C"
