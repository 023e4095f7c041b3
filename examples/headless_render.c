/*
 * headless_render.c — plain-C host driving the render path through the C ABI (include/crt.h) and writing the frame to a PPM
 * file: the headless replacement of CudaRayTracer's window loop (Core/Application.cpp:44-62 -> CudaLayer::RunCudaUpdate,
 * Cuda/CudaLayer.cpp:372-387, whose GL blit has no display to go to on this box).
 *
 *   headless_render [scene 1..5] [width] [height] [spp] [max_depth] [passes] [out.ppm]
 *
 * Progressive: `passes` calls of crt_render add spp samples each to the accumulation buffer (sample indices continue),
 * the frame is resolved after every pass, the last one is written.
 */
#include "crt.h"

#include <stdio.h>
#include <stdlib.h>

#define CK(call)                                                                                   \
    do {                                                                                           \
        crt_status s__ = (call);                                                                   \
        if (s__ != CRT_OK) {                                                                       \
            fprintf(stderr, "%s failed (%d): %s\n", #call, (int)s__, crt_last_error_string(ctx)); \
            return 1;                                                                              \
        }                                                                                          \
    } while (0)

int main(int argc, char** argv)
{
    int kind = argc > 1 ? atoi(argv[1]) : CRT_SCENE_RTIOW;
    uint32_t W = argc > 2 ? (uint32_t)atoi(argv[2]) : 400, H = argc > 3 ? (uint32_t)atoi(argv[3]) : 225;
    uint32_t spp = argc > 4 ? (uint32_t)atoi(argv[4]) : 16, depth = argc > 5 ? (uint32_t)atoi(argv[5]) : 50;
    uint32_t passes = argc > 6 ? (uint32_t)atoi(argv[6]) : 1;
    const char* out = argc > 7 ? argv[7] : "frame.ppm";

    crt_context* ctx = NULL;
    CK(crt_context_create(0, &ctx));
    crt_scene_builder* sb = NULL;
    CK(crt_scene_generate(kind, 1984, 0, &sb));
    CK(crt_scene_upload(ctx, crt_scene_builder_desc(sb)));
    crt_input_struct cam;
    CK(crt_scene_builder_camera(sb, W, H, &cam));
    CK(crt_set_camera(ctx, &cam));
    CK(crt_framebuffer_resize(ctx, W, H));

    uint32_t* frame = (uint32_t*)malloc((size_t)W * H * 4);
    if (!frame) return 2;
    crt_render_params p = {0};
    p.spp_count = spp; p.spp_stride = 1; p.max_depth = depth; p.seed = 1984; p.variant = CRT_VARIANT_AUTO;
    double total_ms = 0.0;
    for (uint32_t k = 0; k < passes; k++) {
        p.spp_begin = k * spp;
        CK(crt_render(ctx, &p));
        CK(crt_resolve(ctx, frame));
        float ms = 0.f;
        CK(crt_last_render_ms(ctx, &ms));
        total_ms += ms;
    }
    crt_counters c;
    CK(crt_get_counters(ctx, &c));
    CK(crt_write_ppm(out, frame, W, H, 1)); /* row 0 is the bottom of the view (CudaLayer.cpp:402) */
    printf("%ux%u, %u x %u spp, depth %u: %.2f ms on the GPU, %llu rays, %.1f Mrays/s -> %s\n", W, H, passes, spp, depth, total_ms,
           (unsigned long long)c.rays, total_ms > 0 ? c.rays / total_ms / 1e3 : 0.0, out);
    free(frame);
    crt_scene_builder_free(sb);
    crt_context_destroy(ctx);
    return 0;
}
