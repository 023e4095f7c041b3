/*
 * crt_legacy.h — the reference's own three C entry points, exported unchanged by libcrt.so so that
 * CudaRayTracer's host (Cuda/CudaLayer.cpp:12-19 declares them, :93-101 and :372-375 call them) links
 * against this library instead of its Kernel.cu without a source change.
 *
 *   Kernel.cu:178-191  LaunchKernel      -> flatten `world` (re-done only when the scene graph changed),
 *                                           render spp samples with the B200 path, resolve into `pos`,
 *                                           synchronise (the reference ends with cudaDeviceSynchronize()).
 *   Kernel.cu:193-197  LaunchRandInit    -> no-op (its state is never read, CudaLayer.cpp:374-375)
 *   Kernel.cu:199-204  LaunchRenderInit  -> no-op apart from restarting the frame counter: Philox is
 *                                           counter-based, there is no per-pixel state to seed.
 *
 * Types are spelled as layout-compatible C PODs: `crt_dim3` for dim3 (3 x unsigned), `crt_input_struct`
 * for InputStruct (SharedStructs.h:3-24, passed BY VALUE), opaque pointers for Hittable* / curandState*.
 * Errors: like the reference the launchers return void; failures are printed to stderr and the frame
 * is left untouched (the reference would have exited through checkCudaErrors).
 */
#ifndef CRT_LEGACY_H
#define CRT_LEGACY_H

#include "crt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct crt_dim3 { unsigned int x, y, z; } crt_dim3;

void LaunchKernel(unsigned int* pos, unsigned int image_width, unsigned int image_height,
                  const unsigned int samples_per_pixel, const unsigned int max_depth, void* world /* Hittable* */,
                  void* d_rand_state /* curandState*, unused */, crt_input_struct inputs);
void LaunchRandInit(void* d_rand_state2 /* curandState*, unused */);
void LaunchRenderInit(crt_dim3 grid, crt_dim3 block, unsigned int window_width, unsigned int window_height,
                      void* d_rand_state /* curandState*, unused */);

/* Counters of the hidden context the three launchers share (for harnesses that time the drop-in). */
crt_status crt_legacy_counters(crt_counters* out);
void       crt_legacy_shutdown(void);

#ifdef __cplusplus
}
#endif
#endif
