// shade.cuh — camera ray generation, hit-point reconstruction, material scatter, textures, background.
//
// Replaces (reference, CudaRayTracer/src): Kernel.cu:130-146 (camera), Kernel.cu:40-77 (miss / material switch),
// Material.cuh:43-62 (Lambertian), :75-94 (Metal), :106-145 (Dielectric), :164-176 (DiffuseLight),
// Texture.cuh:42-45,58-67,83-105 (Constant / Checker / Image), Hittable.cuh:91-95,119-125 (sphere point,
// outward normal, UV), Hittable.cuh:23-27 (SetFaceNormal, rects only), Math.cuh:252-260 (RandomInUnitSphere).
// RNG: the reference draws from a per-pixel XORWOW state; here every (pixel, sample, bounce) owns one
// Philox4x32-10 block of four uniforms, so the streams differ and image parity is statistical.
#pragma once

#include "trace.cuh"

#define CRT_PI 3.141592654f /* Math.cuh:9 */

// Camera ray of Kernel.cu:139-146 for pixel (x, y) with jitter (j1, j2) in (0,1].
template <bool EXACT>
__device__ __forceinline__ void camera_ray(const DevCamera& C, uint32_t width, uint32_t height, int x, int y, float j1, float j2,
                                           float3& origin, float3& dir)
{
    const float cx = (float)width / 2.0f, cy = (float)height / 2.0f; // Kernel.cu:135 (exact in binary32)
    const float fw = (float)width;
    if (EXACT) {
        float u = div_rn(add_rn(sub_rn((float)x, cx), j1), fw);
        float v = div_rn(add_rn(sub_rn(cy, (float)y), j2), fw);
        // distFromCenter = (u * rightV) + (v * upV)
        float3 dist = f3(add_rn(mul_rn(u, C.rx), mul_rn(v, C.ux)), add_rn(mul_rn(u, C.ry), mul_rn(v, C.uy)),
                         add_rn(mul_rn(u, C.rz), mul_rn(v, C.uz)));
        // startPos = (near * dist) + origin + (fov * forwardV)
        float3 start = f3(add_rn(add_rn(mul_rn(C.near_plane, dist.x), C.ox), mul_rn(C.fov, C.fx)),
                          add_rn(add_rn(mul_rn(C.near_plane, dist.y), C.oy), mul_rn(C.fov, C.fy)),
                          add_rn(add_rn(mul_rn(C.near_plane, dist.z), C.oz), mul_rn(C.fov, C.fz)));
        // secondPlanePos = (far * dist) + (far_k * forwardV) + origin
        float3 second = f3(add_rn(add_rn(mul_rn(C.far_plane, dist.x), mul_rn(C.far_k, C.fx)), C.ox),
                           add_rn(add_rn(mul_rn(C.far_plane, dist.y), mul_rn(C.far_k, C.fy)), C.oy),
                           add_rn(add_rn(mul_rn(C.far_plane, dist.z), mul_rn(C.far_k, C.fz)), C.oz));
        float3 dv = f3(sub_rn(second.x, start.x), sub_rn(second.y, start.y), sub_rn(second.z, start.z));
        float inv = div_rn(1.0f, sqrt_rn(dot3_rn(dv, dv))); // Normalize, Math.cuh:225-229
        origin = start;
        dir = f3(mul_rn(inv, dv.x), mul_rn(inv, dv.y), mul_rn(inv, dv.z));
    }
    else {
        float inv_w = fast_rcp(fw);
        float u = (((float)x - cx) + j1) * inv_w;
        float v = ((cy - (float)y) + j2) * inv_w;
        float3 dist = f3(fmaf(u, C.rx, v * C.ux), fmaf(u, C.ry, v * C.uy), fmaf(u, C.rz, v * C.uz));
        float3 start = f3(fmaf(C.fov, C.fx, fmaf(C.near_plane, dist.x, C.ox)), fmaf(C.fov, C.fy, fmaf(C.near_plane, dist.y, C.oy)),
                          fmaf(C.fov, C.fz, fmaf(C.near_plane, dist.z, C.oz)));
        float3 second = f3(fmaf(C.far_plane, dist.x, fmaf(C.far_k, C.fx, C.ox)), fmaf(C.far_plane, dist.y, fmaf(C.far_k, C.fy, C.oy)),
                           fmaf(C.far_plane, dist.z, fmaf(C.far_k, C.fz, C.oz)));
        float3 dv = second - start;
        float inv = rsqrtf(dot3(dv, dv));
        origin = start;
        dir = inv * dv;
    }
}

// Thin-lens extension (not in the reference): move the origin on the lens disk spanned by (right, up)
// and re-aim at the point where the pinhole ray meets the focus plane. (l1, l2) uniform in (0,1].
__device__ __forceinline__ void apply_lens(const DevCamera& C, float l1, float l2, float3& origin, float3& dir)
{
    // concentric-free polar disk sample: r = sqrt(l1), phi = 2 pi l2
    float r = C.lens_radius * fast_sqrt(l1);
    float s, c;
    __sincosf(2.0f * CRT_PI * l2, &s, &c);
    float3 right = f3(C.rx, C.ry, C.rz), up = f3(C.ux, C.uy, C.uz), fwd = f3(C.fx, C.fy, C.fz);
    float3 off = madd3(r * c, right, (r * s) * up);
    float tf = fast_div(C.focus_dist, dot3(dir, fwd));
    float3 focus = madd3(tf, dir, origin);
    origin = origin + off;
    float3 dv = focus - origin;
    dir = rsqrtf(dot3(dv, dv)) * dv;
}

// Uniform point in the unit ball from three uniforms: same distribution as the reference's rejection
// loop (Math.cuh:252-260), fixed cost, no divergence.
__device__ __forceinline__ float3 ball_sample(float u0, float u1, float u2)
{
    float z = fmaf(-2.0f, u0, 1.0f);
    float s = fast_sqrt(fmaxf(0.0f, fmaf(-z, z, 1.0f)));
    float sn, cs;
    __sincosf(2.0f * CRT_PI * u1, &sn, &cs);
    float r = exp2f(__log2f(u2) * (1.0f / 3.0f)); // u2^(1/3): MUFU.LG2 + MUFU.EX2 (u2 in (0,1], so log2 is finite)
    return f3(r * s * cs, r * s * sn, r * z);
}

// Image::value, Texture.cuh:83-105 (texels padded to RGBA8).
__device__ __forceinline__ float3 image_value(const DevImage* images, int image_plus1, float u, float v)
{
    if (image_plus1 == 0)
        return f3(0.f, 1.f, 1.f);
    DevImage im = images[image_plus1 - 1];
    // Clamp of Math.cuh:307-310 is (x < a) ? a : ((x > b) ? b : x): a NaN passes THROUGH it (fminf / fmaxf would drop it). At a
    // sphere pole |n.y| can round above 1, acos gives NaN, v = 1 - NaN stays NaN and the GPU's float -> int conversion makes it
    // row 0 (Texture.cuh:88-89): the same texel the reference kernel and the oracle (f2i_gpu) fetch there.
    u = (u < 0.0f) ? 0.0f : ((u > 1.0f) ? 1.0f : u);
    v = (v < 0.0f) ? 0.0f : ((v > 1.0f) ? 1.0f : v);
    v = 1.0f - v;
    int i = crt_f2i(u * (float)im.width);
    int j = crt_f2i(v * (float)im.height);
    if (i >= im.width) i = im.width - 1;
    if (j >= im.height) j = im.height - 1;
    uchar4 p = __ldg(im.texels + (size_t)j * im.width + i);
    const float k = 1.0f / 255.0f;
    return f3(k * (float)p.x, k * (float)p.y, k * (float)p.z);
}

// Background on a miss, Kernel.cu:41-44.
__device__ __forceinline__ float3 background(const DevCamera& C, float3 d)
{
    float uy = d.y * rsqrtf(dot3(d, d));
    float t = 0.5f * (uy + 1.0f);
    float s = 1.0f - t;
    return f3(fmaf(s, C.bg0[0], t * C.bg1[0]), fmaf(s, C.bg0[1], t * C.bg1[1]), fmaf(s, C.bg0[2], t * C.bg1[2]));
}

// One surface interaction. In: ray (o, d), hit (t, leaf code), four uniforms. Out: scattered ray in (o, d),
// throughput multiplied by the attenuation; returns true if the path continues. When it ends, `emitted`
// holds what the path adds (already multiplied by the throughput): the light's emission, or zero.
template <bool SMEM_ALL>
__device__ __forceinline__ bool shade_hit(const SceneView& S, Hit h, float3& o, float3& d, float3& thr, float3& emitted, float4 xi)
{
    const int slot = h.leaf & CRT_LEAF_INDEX_MASK;
    const int type = (h.leaf >> CRT_LEAF_TYPE_SHIFT) & 3;
    float4 m0, m1;
    load_mat<SMEM_ALL>(S, slot, m0, m1);
    const uint32_t bits = __float_as_uint(m1.w);
    const int mat_type = bits & 15, tex_type = (bits >> 4) & 15, image_plus1 = bits >> 8;

    float3 p = madd3(h.t, d, o); // Ray::PointAtParameter
    float3 n;
    float tu = 0.f, tv = 0.f;
    float4 pr = load_prim<SMEM_ALL>(S, slot);
    if (type == CRT_SPHERE) {
        float inv_r = fast_rcp(pr.w);
        n = inv_r * (p - f3(pr.x, pr.y, pr.z)); // always outward; front_face is never set for spheres
        if (tex_type == CRT_IMAGE) {            // GetSphereUV, Hittable.cuh:119-125
            float theta = acosf(-n.y);
            float phi = atan2f(-n.z, n.x) + CRT_PI;
            tu = phi / (2 * CRT_PI);
            tv = theta / CRT_PI;
        }
    }
    else {
        // outward normal along the rect's axis, flipped to face the ray (SetFaceNormal)
        float3 outward = type == CRT_XYRECT ? f3(0.f, 0.f, 1.f) : type == CRT_XZRECT ? f3(0.f, 1.f, 0.f) : f3(1.f, 0.f, 0.f);
        n = dot3(d, outward) < 0.f ? outward : -outward;
        float pa = type == CRT_YZRECT ? p.y : p.x;
        float pb = type == CRT_XYRECT ? p.y : p.z;
        tu = fast_div(pa - pr.x, pr.y - pr.x);
        tv = fast_div(pb - pr.z, pr.w - pr.z);
    }

    // texture value (dielectric records carry (1,1,1) in m0)
    float3 albedo = f3(m0.x, m0.y, m0.z);
    if (tex_type == CRT_CHECKER) { // Texture.cuh:58-67
        float sines = __sinf(10.f * p.x) * __sinf(10.f * p.y) * __sinf(10.f * p.z);
        if (!(sines < 0.f)) albedo = f3(m1.x, m1.y, m1.z);
    }
    else if (tex_type == CRT_IMAGE) {
        albedo = image_value(S.images, image_plus1, tu, tv);
    }

    emitted = f3(0.f, 0.f, 0.f);
    if (mat_type == CRT_LAMBERTIAN) {
        float3 b = ball_sample(xi.x, xi.y, xi.z);
        d = n + b; // target - p with target = p + normal + random (Material.cuh:46-47)
        o = p;
        thr = albedo * thr;
        return true;
    }
    if (mat_type == CRT_METAL) {
        float3 ud = rsqrtf(dot3(d, d)) * d;
        float3 refl = madd3(-2.0f * dot3(ud, n), n, ud);
        float fuzz = m0.w;
        float3 nd = refl;
        if (fuzz > 0.f) nd = madd3(fuzz, ball_sample(xi.x, xi.y, xi.z), refl);
        o = p;
        d = nd;
        thr = albedo * thr;
        return dot3(nd, n) > 0.f; // absorbed otherwise: the path returns black (Kernel.cu:58-60)
    }
    if (mat_type == CRT_DIELECTRIC) { // Material.cuh:106-136
        const float ir = m0.w;
        float dn = dot3(d, n);
        float inv_len = rsqrtf(dot3(d, d));
        float3 refl = madd3(-2.0f * dn, n, d); // Reflect of the UN-normalised direction
        float3 outn;
        float ratio, cosine;
        if (dn > 0.f) {
            outn = -n;
            ratio = ir;
            cosine = dn * inv_len;
            cosine = fast_sqrt(fmaf(-(ir * ir), fmaf(-cosine, cosine, 1.0f), 1.0f)); // NaN when negative: only used if Refract succeeds
        }
        else {
            outn = n;
            ratio = fast_rcp(ir);
            cosine = -dn * inv_len;
        }
        // Refract, Math.cuh:292-304
        float3 uv = inv_len * d;
        float dt = dot3(uv, outn);
        float disc = fmaf(-(ratio * ratio), fmaf(-dt, dt, 1.0f), 1.0f);
        float prob = 1.0f;
        float3 refr = refl;
        if (disc > 0.f) {
            refr = madd3(-fast_sqrt(disc), outn, ratio * madd3(-dt, outn, uv));
            float r0 = fast_div(1.0f - ir, 1.0f + ir);
            r0 = r0 * r0;
            float c1 = 1.0f - cosine;
            float c2 = c1 * c1;
            prob = fmaf(1.0f - r0, c2 * c2 * c1, r0); // pow(1 - cosine, 5)
        }
        o = p;
        d = (xi.w < prob) ? refl : refr;
        return true; // attenuation (1,1,1)
    }
    // DiffuseLight: emitted = light_intensity * texture; the path ends (Kernel.cu:67-69)
    emitted = thr * (m0.w * albedo);
    return false;
}
