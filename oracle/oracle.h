/*
 * oracle.h — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement, in plain C, of the reference's render path (Trippasch/CudaRayTracer, paths relative
 * to CudaRayTracer/src). Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library; the product (cudaraytracer_b200/libcrt.so) never does.
 *
 * Pinning: the reference has no tests or golden vectors. This oracle is pinned against
 * the reference's OWN headers compiled for the host (oracle/refhost.cpp -> oracle/_ref/libcrt_refhost.so):
 * tests/test_oracle_vs_refhost.py demands bit-identical RGBA8 images, FP32 means, primary-hit IDs and
 * ray counts where /root/reference is present, and tests/golden/ holds vectors generated from that
 * build (tests/golden/make_golden.py) for boxes where it is not.
 */
#ifndef CRT_ORACLE_H
#define CRT_ORACLE_H

#include "../include/crt.h"

#ifdef __cplusplus
extern "C" {
#endif

enum { ORACLE_BVH_REFERENCE = 0, /* BVHNode ctor + BVHNode::Hit as written (Hittable.cuh:303-439) */
       ORACLE_BRUTE_FORCE = 1 }; /* closest hit over the whole list, for scenes deeper than the reference's stack[16] */

/* Same contract as refhost_render (oracle/refhost.cpp). Returns 0 on success. */
int oracle_render(const crt_scene_desc* scene, const crt_input_struct* cam, uint32_t width, uint32_t height, uint32_t spp,
                  uint32_t max_depth, int jitter_center, int n_threads, int accel, uint32_t* out_rgba8, float* out_mean,
                  int32_t* out_ids, uint64_t* out_rays);

/* Seed base of RenderInit (Kernel.cu:175 uses 1984). Only for noise-floor measurements between two oracle runs. */
void oracle_set_seed_base(uint64_t seed);

/* Only pixels with x % sx == 0 and y % sy == 0 are computed (same rays as in the full frame); the rest of the output arrays is
 * left alone. For checks at full benchmark resolution. 1, 1 restores the whole frame. */
void oracle_set_pixel_stride(uint32_t sx, uint32_t sy);

/* Thin lens of the new path (crt_set_lens): aperture = lens diameter, 0 = the reference's pinhole (default). */
void oracle_set_lens(float aperture, float focus_dist);

/* Kernel.cu:151-157 + 12-19 applied to an accumulation buffer (float4 per pixel: rgb sum, sample count). */
void oracle_resolve(const float* accum_float4, uint32_t n_pixels, uint32_t* out_rgba8);

/* XORWOW known-answer hook: n uniforms from curand_init(seed, 0, 0). */
void oracle_curand_uniform(uint64_t seed, int n, float* out);

/* Sphere::Hit (Hittable.cuh:80-110): out = t, p(3), normal(3), u, v. */
int oracle_sphere_hit(const float center[3], float radius, const float ro[3], const float rd[3], float t_min, float t_max,
                      float* out);
/* AABB::Hit (AABB.cuh:30-50) */
int oracle_aabb_hit(const float lo[3], const float hi[3], const float ro[3], const float rd[3], float t_min, float t_max);

/* Philox4x32-10 (Salmon et al. 2011), the RNG of the NEW path, restated for known-answer tests. */
void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);

#ifdef __cplusplus
}
#endif
#endif
