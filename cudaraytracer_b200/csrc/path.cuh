// path.cuh — the per-sample arithmetic every render kernel shares: start of a camera path and what happens to a path after
// its closest-hit query (Kernel.cu:139-146 and :40-79). Pure device functions with no thread-index or shared-memory
// dependence, so that the three kernel variants compute the same FP32 radiance for a (pixel, sample) - and so that
// tests/sim/path_sim.cpp can compile them for the host.
#pragma once

#include "shade.cuh"

// Camera path start for (pixel, sample): Philox block (pixel, sample, 0, CAMERA) -> jitter + lens (Kernel.cu:139-146).
template <bool EXACT>
__device__ __forceinline__ void gen_primary(const RenderArgs& A, int x, int y, uint32_t pixel, uint32_t sample, float3& o, float3& d)
{
    uint4 r = crt_philox4x32_10k(make_uint4(pixel, sample, 0u, CRT_STREAM_CAMERA), A.philox_keys);
    float j1 = 0.5f, j2 = 0.5f;
    if (A.jitter == CRT_JITTER_RANDOM) { j1 = crt_uniform(r.x); j2 = crt_uniform(r.y); }
    camera_ray<EXACT>(A.cam, A.width, A.height, x, y, j1, j2, o, d);
    if (A.cam.lens_radius > 0.f) apply_lens(A.cam, crt_uniform(r.z), crt_uniform(r.w), o, d);
}

// What happens to a path after its closest-hit query (Kernel.cu:40-77). Returns true if the path continues with the
// scattered ray (o, d, thr updated, bounce incremented); false if it ended, with its radiance in `add`.
template <bool SMEM_ALL>
__device__ __forceinline__ bool scatter_step(const RenderArgs& A, const SceneView& S, Hit h, uint32_t pixel, uint32_t sample,
                                             uint32_t& bounce, float3& o, float3& d, float3& thr, float3& add)
{
    if (h.leaf < 0) { // miss: background (Kernel.cu:41-44)
        add = thr * background(A.cam, d);
        return false;
    }
    uint4 r = crt_philox4x32_10k(make_uint4(pixel, sample, bounce, CRT_STREAM_SCATTER), A.philox_keys);
    float4 xi = make_float4(crt_uniform(r.x), crt_uniform(r.y), crt_uniform(r.z), crt_uniform(r.w));
    bool alive = shade_hit<SMEM_ALL>(S, h, o, d, thr, add, xi);
    bounce++;
    if (alive && bounce >= A.max_depth) { // exceeded recursion: black (Kernel.cu:79)
        alive = false;
        add = f3(0.f, 0.f, 0.f);
    }
    return alive;
}
