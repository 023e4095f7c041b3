/*
 * oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE. See oracle.h for the rules and the pinning story.
 *
 * Plain-C restatement of the reference's per-pixel path tracer, operation for operation, in IEEE
 * binary32 without FMA contraction (built with -ffp-contract=off), so that it is bit-identical to the
 * reference's own headers compiled for the host (oracle/refhost.cpp). Every function cites the
 * reference lines it follows (paths relative to CudaRayTracer/src).
 */
#include "oracle.h"

#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#define PI_F 3.141592654f /* Math.cuh:9 */

/* ---------------------------------------------------------------- Vec3, Math.cuh:16-229 */
typedef struct { float e[3]; } v3;

static inline v3 V3(float a, float b, float c) { v3 r; r.e[0] = a; r.e[1] = b; r.e[2] = c; return r; }
static inline v3 vadd(v3 a, v3 b) { return V3(a.e[0] + b.e[0], a.e[1] + b.e[1], a.e[2] + b.e[2]); }  /* :134 */
static inline v3 vsub(v3 a, v3 b) { return V3(a.e[0] - b.e[0], a.e[1] - b.e[1], a.e[2] - b.e[2]); }  /* :139 */
static inline v3 vmul(v3 a, v3 b) { return V3(a.e[0] * b.e[0], a.e[1] * b.e[1], a.e[2] * b.e[2]); }  /* :127 */
static inline v3 smul(float t, v3 v) { return V3(t * v.e[0], t * v.e[1], t * v.e[2]); }              /* :137,:147 */
static inline v3 sdiv(v3 v, float t) { return V3(v.e[0] / t, v.e[1] / t, v.e[2] / t); }              /* :142 */
static inline v3 vneg(v3 a) { return V3(-a.e[0], -a.e[1], -a.e[2]); }                                /* :56 */
static inline float dot(v3 a, v3 b) { return a.e[0] * b.e[0] + a.e[1] * b.e[1] + a.e[2] * b.e[2]; } /* :152 */
static inline v3 cross(v3 a, v3 b)                                                                  /* :157-161 */
{
    return V3((a.e[1] * b.e[2] - a.e[2] * b.e[1]), (-(a.e[0] * b.e[2] - a.e[2] * b.e[0])), (a.e[0] * b.e[1] - a.e[1] * b.e[0]));
}
static inline float vlength(v3 v) { return sqrtf(v.e[0] * v.e[0] + v.e[1] * v.e[1] + v.e[2] * v.e[2]); } /* :77 */
static inline v3 unit_vector(v3 v) { return sdiv(v, vlength(v)); }                                       /* :220 */
static inline v3 normalize(v3 v) { float inv = 1 / sqrtf(dot(v, v)); return smul(inv, v); }              /* :225 */
static inline float clampf(float x, float a, float b) { return (x < a) ? a : ((x > b) ? b : x); }        /* :307 */

/* ---------------------------------------------------------------- cuRAND XORWOW host path
 * curand_kernel.h: _curand_init_scratch (seed scramble; subsequence 0 and offset 0 skip nothing),
 * curand(curandStateXORWOW_t*), curand_uniform.h:69-72 _curand_uniform. Call sites: Kernel.cu:139-140,175;
 * Math.cuh:231-234; Material.cuh:131. */
typedef struct { unsigned int d, v[5]; } xorwow;

static void xorwow_init(xorwow* s, unsigned long long seed)
{
    unsigned int s0 = ((unsigned int)seed) ^ 0xaad26b49U;
    unsigned int s1 = (unsigned int)(seed >> 32) ^ 0xf7dcefddU;
    unsigned int t0 = 1099087573U * s0;
    unsigned int t1 = 2591861531U * s1;
    s->d = 6615241U + t1 + t0;
    s->v[0] = 123456789U + t0;
    s->v[1] = 362436069U ^ t0;
    s->v[2] = 521288629U + t1;
    s->v[3] = 88675123U ^ t1;
    s->v[4] = 5783321U + t0;
}

static inline unsigned int xorwow_next(xorwow* s)
{
    unsigned int t = (s->v[0] ^ (s->v[0] >> 2));
    s->v[0] = s->v[1];
    s->v[1] = s->v[2];
    s->v[2] = s->v[3];
    s->v[3] = s->v[4];
    s->v[4] = (s->v[4] ^ (s->v[4] << 4)) ^ (t ^ (t << 1));
    s->d += 362437U;
    return s->v[4] + s->d;
}

static inline float xorwow_uniform(xorwow* s)
{
    unsigned int x = xorwow_next(s);
    return x * 2.3283064e-10f + (2.3283064e-10f / 2.0f);
}

/* Math.cuh:231-234 Random(): Vec3(u(), u(), u()). The three calls are function arguments, whose
 * evaluation order C++ leaves unspecified; g++ (the compiler of oracle/_ref) evaluates them last to
 * first, so e[2] is drawn first. tests/test_oracle_vs_refhost.py pins this choice. */
static inline v3 random_vec(xorwow* s)
{
    v3 r;
    r.e[2] = xorwow_uniform(s);
    r.e[1] = xorwow_uniform(s);
    r.e[0] = xorwow_uniform(s);
    return r;
}

/* Math.cuh:252-260 */
static inline v3 random_in_unit_sphere(xorwow* s)
{
    v3 p;
    do {
        p = vsub(smul(2.0f, random_vec(s)), V3(1.0f, 1.0f, 1.0f));
    } while ((p.e[0] * p.e[0] + p.e[1] * p.e[1] + p.e[2] * p.e[2]) >= 1.0f);
    return p;
}

/* Math.cuh:287-304 */
static inline v3 reflect(v3 v, v3 n) { return vsub(v, smul(2.0f * dot(v, n), n)); }

static inline int refract(v3 v, v3 n, float eta, v3* refracted)
{
    v3 uv = unit_vector(v);
    float dt = dot(uv, n);
    float discriminant = 1.0f - eta * eta * (1 - dt * dt);
    if (discriminant > 0) {
        *refracted = vsub(smul(eta, vsub(uv, smul(dt, n))), smul(sqrtf(discriminant), n));
        return 1;
    }
    return 0;
}

/* ---------------------------------------------------------------- scene records (flat, from crt.h) */
typedef struct { v3 orig, dir; } ray;

typedef struct
{
    v3 p, normal;
    int obj; /* stands for mat_ptr: index of the hittable whose material was hit */
    float t, u, v;
    int front_face;
} hit_record; /* Hittable.cuh:14-28 */

typedef struct bvh_node
{
    v3 bmin, bmax;
    /* child: >= 0 hittable index (leaf object), -1 null, <= -2: inner node ~(index+... ) see CHILD_NODE */
    int left, right;
    int left_is_node, right_is_node;
} bvh_node;

typedef struct
{
    const crt_scene_desc* sd;
    bvh_node* nodes;
    int n_nodes, cap_nodes;
    int accel;
    int* active; /* brute force: indices of active hittables in list order */
    int n_active;
} oscene;

static inline v3 ray_at(const ray* r, float t) { return vadd(r->orig, smul(t, r->dir)); } /* Ray.cuh:26-29 */

/* Hittable.cuh:119-125 */
static inline void sphere_uv(v3 p, float* u, float* v)
{
    float theta = acosf(-p.e[1]);
    float phi = atan2f(-p.e[2], p.e[0]) + PI_F;
    *u = phi / (2 * PI_F);
    *v = theta / PI_F;
}

/* Sphere::Hit, Hittable.cuh:80-110 */
static int sphere_hit(const crt_hittable* s, int index, const ray* r, float t_min, float t_max, hit_record* rec)
{
    v3 center = V3(s->cx, s->cy, s->cz);
    float radius = s->a;
    v3 oc = vsub(r->orig, center);
    float a = dot(r->dir, r->dir);
    float b = dot(oc, r->dir);
    float c = dot(oc, oc) - radius * radius;
    float discriminant = b * b - a * c;
    if (discriminant > 0) {
        float temp = (-b - sqrtf(discriminant)) / a;
        if (temp < t_max && temp > t_min) {
            rec->t = temp;
            rec->p = ray_at(r, rec->t);
            rec->normal = sdiv(vsub(rec->p, center), radius);
            sphere_uv(rec->normal, &rec->u, &rec->v);
            rec->obj = index;
            return 1;
        }
        temp = (-b + sqrtf(discriminant)) / a;
        if (temp < t_max && temp > t_min) {
            rec->t = temp;
            rec->p = ray_at(r, rec->t);
            rec->normal = sdiv(vsub(rec->p, center), radius);
            sphere_uv(rec->normal, &rec->u, &rec->v);
            rec->obj = index;
            return 1;
        }
    }
    return 0;
}

/* HitRecord::SetFaceNormal, Hittable.cuh:23-27 */
static inline void set_face_normal(hit_record* rec, const ray* r, v3 outward)
{
    rec->front_face = dot(r->dir, outward) < 0;
    rec->normal = rec->front_face ? outward : vneg(outward);
}

/* XYRect / XZRect / YZRect::Hit, Hittable.cuh:139-168, 195-224, 251-280.
 * ia, ib = in-plane axes, ik = normal axis. wa / wb = the extent used along ia / ib:
 * XY: x<-width, y<-height; XZ: x<-width, z<-height; YZ: y<-height, z<-width. */
static int rect_hit(const crt_hittable* h, int index, const ray* r, float t_min, float t_max, hit_record* rec)
{
    int ia, ib, ik;
    float wa, wb;
    float c[3] = {h->cx, h->cy, h->cz};
    if (h->type == CRT_XYRECT) { ia = 0; ib = 1; ik = 2; wa = h->a; wb = h->b; }
    else if (h->type == CRT_XZRECT) { ia = 0; ib = 2; ik = 1; wa = h->a; wb = h->b; }
    else { ia = 1; ib = 2; ik = 0; wa = h->b; wb = h->a; }
    float a0 = c[ia] - (wa / 2), a1 = c[ia] + (wa / 2);
    float b0 = c[ib] - (wb / 2), b1 = c[ib] + (wb / 2);
    float k = c[ik];
    float inv = 1.0f / r->dir.e[ik];
    float t = (k - r->orig.e[ik]) * inv;
    if (t < t_min || t > t_max)
        return 0;
    float pa = r->orig.e[ia] + t * r->dir.e[ia];
    float pb = r->orig.e[ib] + t * r->dir.e[ib];
    if (pa < a0 || pa > a1 || pb < b0 || pb > b1)
        return 0;
    rec->u = (pa - a0) / (a1 - a0);
    rec->v = (pb - b0) / (b1 - b0);
    rec->t = t;
    v3 outward = V3(ik == 0 ? 1.0f : 0.0f, ik == 1 ? 1.0f : 0.0f, ik == 2 ? 1.0f : 0.0f);
    set_face_normal(rec, r, outward);
    rec->obj = index;
    rec->p = ray_at(r, t);
    return 1;
}

/* BVHNode::PerformHit, Hittable.cuh:470-485 */
static inline int perform_hit(const oscene* S, int obj, const ray* r, float t_min, float t_max, hit_record* rec)
{
    const crt_hittable* h = &S->sd->hittables[obj];
    if (h->type == CRT_SPHERE)
        return sphere_hit(h, obj, r, t_min, t_max, rec);
    return rect_hit(h, obj, r, t_min, t_max, rec);
}

/* AABB::Hit, AABB.cuh:30-50 */
static int aabb_hit(v3 bmin, v3 bmax, const ray* r, float t_min, float t_max)
{
    for (int a = 0; a < 3; a++) {
        float invD = 1.0f / r->dir.e[a];
        float t0 = (bmin.e[a] - r->orig.e[a]) * invD;
        float t1 = (bmax.e[a] - r->orig.e[a]) * invD;
        if (invD < 0.0f) {
            float temp = t0;
            t0 = t1;
            t1 = temp;
        }
        t_min = t0 > t_min ? t0 : t_min;
        t_max = t1 < t_max ? t1 : t_max;
        if (t_max <= t_min)
            return 0;
    }
    return 1;
}

/* ---------------------------------------------------------------- BVH: BVHNode ctor, Hittable.cuh:303-385
 * Median split of the LIST ORDER (after dropping inactive objects and a stable sort by type), leaves of
 * one (left == right) or two objects. Children are encoded as: is_node ? node index : hittable index (-1 = null). */
static void prim_box(const crt_hittable* h, v3* lo, v3* hi)
{
    float c[3] = {h->cx, h->cy, h->cz};
    if (h->type == CRT_SPHERE) { /* Hittable.cuh:112-116 */
        *lo = vsub(V3(c[0], c[1], c[2]), V3(h->a, h->a, h->a));
        *hi = vadd(V3(c[0], c[1], c[2]), V3(h->a, h->a, h->a));
        return;
    }
    int ia, ib, ik;
    float wa, wb;
    if (h->type == CRT_XYRECT) { ia = 0; ib = 1; ik = 2; wa = h->a; wb = h->b; }
    else if (h->type == CRT_XZRECT) { ia = 0; ib = 2; ik = 1; wa = h->a; wb = h->b; }
    else { ia = 1; ib = 2; ik = 0; wa = h->b; wb = h->a; }
    lo->e[ia] = c[ia] - (wa / 2); hi->e[ia] = c[ia] + (wa / 2);
    lo->e[ib] = c[ib] - (wb / 2); hi->e[ib] = c[ib] + (wb / 2);
    lo->e[ik] = c[ik] - 0.0001f;  hi->e[ik] = c[ik] + 0.0001f;
}

static int new_node(oscene* S)
{
    if (S->n_nodes == S->cap_nodes) {
        S->cap_nodes = S->cap_nodes ? S->cap_nodes * 2 : 64;
        S->nodes = (bvh_node*)realloc(S->nodes, (size_t)S->cap_nodes * sizeof(bvh_node));
    }
    memset(&S->nodes[S->n_nodes], 0, sizeof(bvh_node));
    return S->n_nodes++;
}

/* objs[0..n) = the (already active-filtered) slice the reference would pass as [start, end). */
static void build_node(oscene* S, int me, int* objs, int n)
{
    const crt_hittable* H = S->sd->hittables;
    /* thrust::sort by type on the host is a stable merge sort; insertion sort keeps it stable here. */
    for (int i = 1; i < n; i++) {
        int key = objs[i], j = i - 1;
        while (j >= 0 && H[objs[j]].type > H[key].type) { objs[j + 1] = objs[j]; j--; }
        objs[j + 1] = key;
    }
    bvh_node nd;
    memset(&nd, 0, sizeof(nd));
    nd.left = nd.right = -1;
    if (n == 0) { /* :316-319: no active objects; box stays AABB() = zeros */
        S->nodes[me] = nd;
        return;
    }
    if (n == 1) {
        nd.left = nd.right = objs[0];
    }
    else if (n == 2) {
        nd.left = objs[0];
        nd.right = objs[1];
    }
    else {
        int first_type = H[objs[0]].type, mid = 0;
        while (mid < n && H[objs[mid]].type == first_type) mid++; /* partition of a type-sorted range */
        if (mid == 0 || mid == n) mid = n / 2;                      /* :340-342 */
        int l = new_node(S), r = new_node(S);
        nd.left = l; nd.left_is_node = 1;
        nd.right = r; nd.right_is_node = 1;
        build_node(S, l, objs, mid);
        build_node(S, r, objs + mid, n - mid);
    }
    v3 bl = V3(0, 0, 0), bh = V3(0, 0, 0), rl = V3(0, 0, 0), rh = V3(0, 0, 0);
    if (nd.left_is_node) { bl = S->nodes[nd.left].bmin; bh = S->nodes[nd.left].bmax; }
    else prim_box(&H[nd.left], &bl, &bh);
    if (nd.right_is_node) { rl = S->nodes[nd.right].bmin; rh = S->nodes[nd.right].bmax; }
    else prim_box(&H[nd.right], &rl, &rh);
    /* SurroundingBox, AABB.cuh:53-62 */
    nd.bmin = V3(fminf(bl.e[0], rl.e[0]), fminf(bl.e[1], rl.e[1]), fminf(bl.e[2], rl.e[2]));
    nd.bmax = V3(fmaxf(bh.e[0], rh.e[0]), fmaxf(bh.e[1], rh.e[1]), fmaxf(bh.e[2], rh.e[2]));
    S->nodes[me] = nd;
}

/* BVHNode::Hit, Hittable.cuh:387-439. The reference's stack holds 16 entries with no overflow check;
 * 64 here changes nothing for trees the reference can traverse and keeps deeper ones defined. */
static int bvh_hit(const oscene* S, const ray* r, float t_min, float t_max, hit_record* rec)
{
    const bvh_node* N = S->nodes;
    if (!aabb_hit(N[0].bmin, N[0].bmax, r, t_min, t_max))
        return 0;
    struct { int node; float t_min, t_max; } stack[64], current;
    int top = -1;
    stack[++top].node = 0; stack[top].t_min = t_min; stack[top].t_max = t_max;
    int hit_something = 0;
    while (top >= 0) {
        current.node = stack[top].node; current.t_min = stack[top].t_min; current.t_max = stack[top].t_max;
        top--;
        const bvh_node* n = &N[current.node];
        if (!aabb_hit(n->bmin, n->bmax, r, current.t_min, current.t_max))
            continue;
        if (n->left >= 0 && n->left_is_node) {
            ++top;
            stack[top].node = n->left; stack[top].t_min = current.t_min; stack[top].t_max = hit_something ? rec->t : current.t_max;
        }
        else if (n->left >= 0) {
            hit_something |= perform_hit(S, n->left, r, current.t_min, hit_something ? rec->t : current.t_max, rec);
        }
        if (n->right >= 0 && n->right_is_node) {
            ++top;
            stack[top].node = n->right; stack[top].t_min = current.t_min; stack[top].t_max = hit_something ? rec->t : current.t_max;
        }
        else if (n->right >= 0) {
            hit_something |= perform_hit(S, n->right, r, current.t_min, hit_something ? rec->t : current.t_max, rec);
        }
    }
    return hit_something;
}

static int brute_hit(const oscene* S, const ray* r, float t_min, float t_max, hit_record* rec)
{
    int hit_something = 0;
    for (int i = 0; i < S->n_active; i++)
        hit_something |= perform_hit(S, S->active[i], r, t_min, hit_something ? rec->t : t_max, rec);
    return hit_something;
}

static inline int world_hit(const oscene* S, const ray* r, float t_min, float t_max, hit_record* rec)
{
    return S->accel == ORACLE_BRUTE_FORCE ? brute_hit(S, r, t_min, t_max, rec) : bvh_hit(S, r, t_min, t_max, rec);
}

/* ---------------------------------------------------------------- textures, Texture.cuh:42-105 */
/* float -> int as the GPU performs it (cvt.rzi.s32.f32: NaN -> 0, saturating). Image::value (Texture.cuh:88-89) converts
 * u * width and v * height with a plain cast; GetSphereUV (Hittable.cuh:119-125) yields NaN at a sphere's poles when
 * |normal.y| rounds above 1 (acos), and the cast of NaN is undefined on the host: the reference's headers compiled for x86
 * index the texel array with INT_MIN there and crash (seen on the textured scene from 4 spp up). The kernel this oracle
 * stands for is the CUDA build, so the conversion follows the GPU. */
static int f2i_gpu(float x)
{
    if (x != x) return 0;
    if (x >= 2147483648.0f) return 2147483647;
    if (x <= -2147483648.0f) return (-2147483647 - 1);
    return (int)x;
}

static v3 texture_value(const oscene* S, const crt_texture* t, float u, float v, v3 p)
{
    if (t->type == CRT_CONSTANT)
        return V3(t->color[0], t->color[1], t->color[2]);
    if (t->type == CRT_CHECKER) {
        float sines = sinf(10 * p.e[0]) * sinf(10 * p.e[1]) * sinf(10 * p.e[2]);
        if (sines < 0)
            return V3(t->odd[0], t->odd[1], t->odd[2]);
        return V3(t->even[0], t->even[1], t->even[2]);
    }
    /* Image::value */
    const crt_image* im = (t->image >= 0 && (uint64_t)t->image < S->sd->n_images) ? &S->sd->images[t->image] : 0;
    if (!im || !im->pixels)
        return V3(0, 1, 1);
    u = clampf(u, 0.0f, 1.0f);
    v = 1.0f - clampf(v, 0.0f, 1.0f);
    int i = f2i_gpu(u * im->width);
    int j = f2i_gpu(v * im->height);
    if (i >= im->width) i = im->width - 1;
    if (j >= im->height) j = im->height - 1;
    const float color_scale = 1.0f / 255.0f;
    const uint8_t* pixel = im->pixels + (size_t)j * (3 * im->width) + (size_t)i * 3;
    return V3(color_scale * pixel[0], color_scale * pixel[1], color_scale * pixel[2]);
}

/* ---------------------------------------------------------------- color(), Kernel.cu:30-80 with the
 * Scatter/Emitted bodies of Material.cuh:43-62, 75-94, 106-145, 164-176 inlined at their call sites. */
static v3 color(const oscene* S, ray r, int max_depth, xorwow* rng, const crt_input_struct* in, uint64_t* rays)
{
    ray cur_ray = r;
    v3 cur_attenuation = V3(1.0f, 1.0f, 1.0f);
    v3 black = V3(0.0f, 0.0f, 0.0f);
    v3 bg0 = V3(in->background_start[0], in->background_start[1], in->background_start[2]);
    v3 bg1 = V3(in->background_end[0], in->background_end[1], in->background_end[2]);
    hit_record rec;
    memset(&rec, 0, sizeof(rec));

    for (int i = 0; i < max_depth; i++) {
        (*rays)++;
        if (!world_hit(S, &cur_ray, 0.001f, FLT_MAX, &rec)) {
            v3 unit_direction = unit_vector(cur_ray.dir);
            float t = 0.5f * (unit_direction.e[1] + 1.0f);
            v3 c = vadd(smul(1.0f - t, bg0), smul(t, bg1));
            return vmul(cur_attenuation, c);
        }
        const crt_material* m = &S->sd->materials[S->sd->hittables[rec.obj].material];
        const crt_texture* tex = m->type != CRT_DIELECTRIC ? &S->sd->textures[m->texture] : 0;
        v3 emitted = V3(0.0f, 0.0f, 0.0f);
        ray scattered;
        v3 attenuation;
        switch (m->type) {
        case CRT_LAMBERTIAN: {
            v3 target = vadd(vadd(rec.p, rec.normal), random_in_unit_sphere(rng));
            scattered.orig = rec.p;
            scattered.dir = vsub(target, rec.p);
            attenuation = texture_value(S, tex, rec.u, rec.v, rec.p);
            break;
        }
        case CRT_METAL: {
            float fuzz = m->param < 1 ? m->param : 1; /* Material.cuh:71 */
            v3 reflected = reflect(unit_vector(cur_ray.dir), rec.normal);
            scattered.orig = rec.p;
            scattered.dir = vadd(reflected, smul(fuzz, random_in_unit_sphere(rng)));
            attenuation = texture_value(S, tex, rec.u, rec.v, rec.p);
            if (!(dot(scattered.dir, rec.normal) > 0))
                return vmul(emitted, cur_attenuation);
            break;
        }
        case CRT_DIELECTRIC: {
            float ir = m->param;
            v3 outward_normal;
            v3 reflected = reflect(cur_ray.dir, rec.normal);
            float ni_over_nt;
            attenuation = V3(1.0f, 1.0f, 1.0f);
            v3 refracted = V3(0.0f, 0.0f, 0.0f);
            float reflect_prob;
            float cosine;
            if (dot(cur_ray.dir, rec.normal) > 0.0f) {
                outward_normal = vneg(rec.normal);
                ni_over_nt = ir;
                cosine = dot(cur_ray.dir, rec.normal) / vlength(cur_ray.dir);
                cosine = sqrtf(1.0f - ir * ir * (1 - cosine * cosine));
            }
            else {
                outward_normal = rec.normal;
                ni_over_nt = 1.0f / ir;
                cosine = -dot(cur_ray.dir, rec.normal) / vlength(cur_ray.dir);
            }
            if (refract(cur_ray.dir, outward_normal, ni_over_nt, &refracted)) {
                float r0 = (1.0f - ir) / (1.0f + ir); /* Reflectance, Material.cuh:139-145 */
                r0 = r0 * r0;
                reflect_prob = r0 + (1.0f - r0) * powf((1.0f - cosine), 5.0f);
            }
            else
                reflect_prob = 1.0f;
            scattered.orig = rec.p;
            if (xorwow_uniform(rng) < reflect_prob)
                scattered.dir = reflected;
            else
                scattered.dir = refracted;
            break;
        }
        case CRT_DIFFUSELIGHT: {
            v3 tv = texture_value(S, tex, rec.u, rec.v, rec.p);
            emitted = smul((float)m->light_intensity, tv); /* int * Vec3 -> operator*(float, Vec3) */
            return vmul(emitted, cur_attenuation);
        }
        default:
            return black;
        }
        cur_attenuation = vmul(attenuation, cur_attenuation);
        cur_ray = scattered;
    }
    return black;
}

/* ---------------------------------------------------------------- Kernel body, Kernel.cu:113-157 */
/* The reference seeds with the constant 1984 (Kernel.cu:175). Tests move it only to measure the
 * Monte-Carlo noise floor between two independent oracle runs; 1984 is the default. */
static unsigned long long g_seed_base = 1984;
void oracle_set_seed_base(uint64_t seed) { g_seed_base = seed; }

/* Pixel subset for checks at full benchmark resolution (bench.py `parity`, tests of the 4K / 8K workloads): only pixels with
 * x % sx == 0 and y % sy == 0 are computed, with exactly the rays they have in the full frame; the others keep what the
 * caller put into the output arrays. 1, 1 (the default) is the whole frame. */
static uint32_t g_stride_x = 1, g_stride_y = 1;
void oracle_set_pixel_stride(uint32_t sx, uint32_t sy)
{
    g_stride_x = sx ? sx : 1;
    g_stride_y = sy ? sy : 1;
}

/* Thin-lens extension of the NEW path (crt_set_lens; north_star asks for defocus, the reference renders a pinhole and only
 * carries an unused RandomInUnitDisk, Math.cuh:276-285). Restated here so that the lens has a checker: lens point uniform on
 * the disk of radius aperture / 2 spanned by (right, up) - the polar map of two uniforms has the distribution of the
 * reference's rejection sampler -, aimed at the point where the pinhole ray meets the plane at focus_dist along forward.
 * aperture == 0 (the default) draws nothing and is the reference camera bit for bit. */
static float g_lens_radius = 0.0f, g_focus_dist = 1.0f;
void oracle_set_lens(float aperture, float focus_dist)
{
    g_lens_radius = aperture > 0.0f ? 0.5f * aperture : 0.0f;
    g_focus_dist = focus_dist > 0.0f ? focus_dist : 1.0f;
}

typedef struct
{
    const oscene* S;
    const crt_input_struct* in;
    uint32_t width, height, spp, max_depth;
    int jitter_center;
    uint32_t* out_rgba8;
    float* out_mean;
    int32_t* out_ids;
    uint32_t next_row; /* guarded by mu */
    uint64_t rays;
    pthread_mutex_t mu;
} job;

/* Kernel.cu:12-19 */
static inline int rgb_to_int(float r, float g, float b)
{
    r = clampf(r, 0.0f, 255.0f);
    g = clampf(g, 0.0f, 255.0f);
    b = clampf(b, 0.0f, 255.0f);
    float a = 255.0f;
    return ((int)(a) << 24) | ((int)(b) << 16) | ((int)(g) << 8) | (int)(r);
}

static inline ray camera_ray(const crt_input_struct* in, v3 rightV, uint32_t width, uint32_t height, int x, int y, float j1, float j2)
{
    v3 origin = V3(in->origin_x, in->origin_y, in->origin_z);
    v3 forwardV = V3(in->orientation_x, in->orientation_y, in->orientation_z);
    v3 upV = V3(in->up_x, in->up_y, in->up_z);
    v3 center = V3(width / 2.0f, height / 2.0f, 0.0f);
    float u = (float)((x - center.e[0]) + j1) / (float)(width);
    float v = (float)((center.e[1] - y) + j2) / (float)(width);
    v3 distFromCenter = vadd(smul(u, rightV), smul(v, upV));
    v3 startPos = vadd(vadd(smul(in->near_plane, distFromCenter), origin), smul(in->fov, forwardV));
    v3 secondPlanePos = vadd(vadd(smul(in->far_plane, distFromCenter), smul((1.0f / in->fov * 10.0f), forwardV)), origin);
    v3 dirVector = normalize(vsub(secondPlanePos, startPos));
    ray r;
    r.orig = startPos;
    r.dir = dirVector;
    return r;
}

static void pixel(job* J, int x, int y, uint64_t* rays)
{
    const crt_input_struct* in = J->in;
    const uint32_t width = J->width, height = J->height;
    unsigned int pixel_index = (y * width + x);
    xorwow rng;
    xorwow_init(&rng, g_seed_base + pixel_index); /* RenderInit, Kernel.cu:175: 1984 + pixel_index */

    v3 forwardV = V3(in->orientation_x, in->orientation_y, in->orientation_z);
    v3 upV = V3(in->up_x, in->up_y, in->up_z);
    v3 rightV = normalize(cross(upV, forwardV));

    if (J->out_ids) {
        ray r = camera_ray(in, rightV, width, height, x, y, 0.5f, 0.5f);
        hit_record rec;
        memset(&rec, 0, sizeof(rec));
        J->out_ids[pixel_index] = world_hit(J->S, &r, 0.001f, FLT_MAX, &rec) ? rec.obj : -1;
    }
    if (!J->out_rgba8 && !J->out_mean)
        return;

    v3 col = V3(0.0f, 0.0f, 0.0f);
    for (uint32_t s = 0; s < J->spp; s++) {
        float j1 = J->jitter_center ? 0.5f : xorwow_uniform(&rng);
        float j2 = J->jitter_center ? 0.5f : xorwow_uniform(&rng);
        ray r = camera_ray(in, rightV, width, height, x, y, j1, j2);
        if (g_lens_radius > 0.0f) {
            float l1 = xorwow_uniform(&rng), l2 = xorwow_uniform(&rng);
            float rad = g_lens_radius * sqrtf(l1), phi = 2.0f * 3.141592654f * l2;
            v3 off = vadd(smul(rad * cosf(phi), rightV), smul(rad * sinf(phi), upV));
            float tf = g_focus_dist / dot(r.dir, forwardV);
            v3 focus = vadd(r.orig, smul(tf, r.dir));
            r.orig = vadd(r.orig, off);
            r.dir = normalize(vsub(focus, r.orig));
        }
        col = vadd(col, color(J->S, r, (int)J->max_depth, &rng, in, rays));
    }
    col = sdiv(col, (float)(J->spp));
    if (J->out_mean) {
        J->out_mean[3 * (size_t)pixel_index + 0] = col.e[0];
        J->out_mean[3 * (size_t)pixel_index + 1] = col.e[1];
        J->out_mean[3 * (size_t)pixel_index + 2] = col.e[2];
    }
    col.e[0] = 255.0f * sqrtf(col.e[0]);
    col.e[1] = 255.0f * sqrtf(col.e[1]);
    col.e[2] = 255.0f * sqrtf(col.e[2]);
    if (J->out_rgba8)
        J->out_rgba8[pixel_index] = (uint32_t)rgb_to_int(col.e[0], col.e[1], col.e[2]);
}

static void* worker(void* arg)
{
    job* J = (job*)arg;
    uint64_t rays = 0;
    for (;;) {
        pthread_mutex_lock(&J->mu);
        uint32_t y = J->next_row++;
        pthread_mutex_unlock(&J->mu);
        if (y >= J->height)
            break;
        if (y % g_stride_y)
            continue;
        for (uint32_t x = 0; x < J->width; x += g_stride_x)
            pixel(J, (int)x, (int)y, &rays);
    }
    pthread_mutex_lock(&J->mu);
    J->rays += rays;
    pthread_mutex_unlock(&J->mu);
    return 0;
}

int oracle_render(const crt_scene_desc* scene, const crt_input_struct* cam, uint32_t width, uint32_t height, uint32_t spp,
                  uint32_t max_depth, int jitter_center, int n_threads, int accel, uint32_t* out_rgba8, float* out_mean,
                  int32_t* out_ids, uint64_t* out_rays)
{
    if (!scene || !cam || !width || !height)
        return 1;
    for (uint64_t i = 0; i < scene->n_hittables; i++) {
        const crt_hittable* h = &scene->hittables[i];
        if (h->type < CRT_SPHERE || h->type > CRT_YZRECT || h->material < 0 || (uint64_t)h->material >= scene->n_materials)
            return 1;
        const crt_material* m = &scene->materials[h->material];
        if (m->type < CRT_LAMBERTIAN || m->type > CRT_DIFFUSELIGHT)
            return 1;
        if (m->type != CRT_DIELECTRIC && (m->texture < 0 || (uint64_t)m->texture >= scene->n_textures))
            return 1;
    }
    oscene S;
    memset(&S, 0, sizeof(S));
    S.sd = scene;
    S.accel = accel;
    S.active = (int*)malloc(sizeof(int) * (size_t)(scene->n_hittables ? scene->n_hittables : 1));
    for (uint64_t i = 0; i < scene->n_hittables; i++)
        if (scene->hittables[i].active)
            S.active[S.n_active++] = (int)i;
    if (accel == ORACLE_BVH_REFERENCE) {
        int* objs = (int*)malloc(sizeof(int) * (size_t)(S.n_active ? S.n_active : 1));
        memcpy(objs, S.active, sizeof(int) * (size_t)S.n_active);
        int root = new_node(&S);
        build_node(&S, root, objs, S.n_active);
        free(objs);
    }

    job J;
    memset(&J, 0, sizeof(J));
    J.S = &S; J.in = cam; J.width = width; J.height = height; J.spp = spp; J.max_depth = max_depth;
    J.jitter_center = jitter_center; J.out_rgba8 = out_rgba8; J.out_mean = out_mean; J.out_ids = out_ids;
    pthread_mutex_init(&J.mu, 0);
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 1024) n_threads = 1024;
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)n_threads);
    for (int i = 1; i < n_threads; i++)
        pthread_create(&th[i], 0, worker, &J);
    worker(&J);
    for (int i = 1; i < n_threads; i++)
        pthread_join(th[i], 0);
    free(th);
    pthread_mutex_destroy(&J.mu);
    if (out_rays)
        *out_rays = J.rays;
    free(S.nodes);
    free(S.active);
    return 0;
}

void oracle_resolve(const float* accum, uint32_t n_pixels, uint32_t* out_rgba8)
{
    for (uint32_t i = 0; i < n_pixels; i++) {
        const float* a = accum + 4 * (size_t)i;
        if (!(a[3] > 0.0f)) { /* no samples yet: opaque black (the reference always has spp >= 1) */
            out_rgba8[i] = 0xFF000000u;
            continue;
        }
        v3 col = sdiv(V3(a[0], a[1], a[2]), a[3]);
        col.e[0] = 255.0f * sqrtf(col.e[0]);
        col.e[1] = 255.0f * sqrtf(col.e[1]);
        col.e[2] = 255.0f * sqrtf(col.e[2]);
        out_rgba8[i] = (uint32_t)rgb_to_int(col.e[0], col.e[1], col.e[2]);
    }
}

void oracle_curand_uniform(uint64_t seed, int n, float* out)
{
    xorwow s;
    xorwow_init(&s, seed);
    for (int i = 0; i < n; i++)
        out[i] = xorwow_uniform(&s);
}

int oracle_sphere_hit(const float center[3], float radius, const float ro[3], const float rd[3], float t_min, float t_max,
                      float* out)
{
    crt_hittable s;
    memset(&s, 0, sizeof(s));
    s.type = CRT_SPHERE; s.cx = center[0]; s.cy = center[1]; s.cz = center[2]; s.a = radius;
    ray r;
    r.orig = V3(ro[0], ro[1], ro[2]);
    r.dir = V3(rd[0], rd[1], rd[2]);
    hit_record rec;
    memset(&rec, 0, sizeof(rec));
    int h = sphere_hit(&s, 0, &r, t_min, t_max, &rec);
    if (h) {
        out[0] = rec.t;
        out[1] = rec.p.e[0]; out[2] = rec.p.e[1]; out[3] = rec.p.e[2];
        out[4] = rec.normal.e[0]; out[5] = rec.normal.e[1]; out[6] = rec.normal.e[2];
        out[7] = rec.u; out[8] = rec.v;
    }
    return h;
}

int oracle_aabb_hit(const float lo[3], const float hi[3], const float ro[3], const float rd[3], float t_min, float t_max)
{
    ray r;
    r.orig = V3(ro[0], ro[1], ro[2]);
    r.dir = V3(rd[0], rd[1], rd[2]);
    return aabb_hit(V3(lo[0], lo[1], lo[2]), V3(hi[0], hi[1], hi[2]), &r, t_min, t_max);
}

/* Philox4x32-10: Salmon, Moraes, Dror, Shaw, "Parallel random numbers: as easy as 1, 2, 3", SC'11 (Random123
 * philox.h constants). Not part of the reference; restated here only to check the new path's RNG against the
 * published known-answer vectors. */
void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
{
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
