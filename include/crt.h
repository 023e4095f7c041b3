/*
 * crt.h — C ABI of the B200-native path tracer that replaces CudaRayTracer's render path.
 *
 * Boundary being replaced (reference = Trippasch/CudaRayTracer, paths relative to CudaRayTracer/src):
 *   - Cuda/Kernel.cu:178-204      extern "C" LaunchKernel / LaunchRandInit / LaunchRenderInit
 *     (re-declared by the caller in Cuda/CudaLayer.cpp:12-19). Those three symbols are exported
 *     unchanged by this library (see crt_legacy.h); everything in this header is the native
 *     surface they are built on.
 *   - Utils/SharedStructs.h:3-24  InputStruct (camera + background block passed by value)
 *   - Hittables/Hittable.cuh:30-38, Material.cuh:6-12, Texture.cuh:7-12  the enum values
 *
 * Conventions: plain pointers and sizes, no C++ or torch types. Every entry point returns a
 * crt_status (never exits the process; the reference's checkCudaErrors aborts, helper_cuda.h:599-612).
 * Pointers are HOST pointers unless the name says "device". Framebuffer row 0 is the BOTTOM of the
 * view, as in the reference (CudaLayer.cpp:402 draws it V-flipped).
 */
#ifndef CRT_H
#define CRT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CRT_ABI_VERSION 4

typedef enum crt_status {
    CRT_OK = 0,
    CRT_ERR_INVALID = 1,     /* bad argument */
    CRT_ERR_CUDA = 2,        /* a CUDA runtime call failed; see crt_last_error_string */
    CRT_ERR_NOMEM = 3,
    CRT_ERR_UNSUPPORTED = 4,
    CRT_ERR_STATE = 5        /* call order violated (e.g. render before scene upload) */
} crt_status;

/* Values equal the reference enums so a record can be copied across without translation. */
enum { CRT_SPHERE = 0, CRT_XYRECT = 1, CRT_XZRECT = 2, CRT_YZRECT = 3 };            /* Hittable.cuh:30-38 */
enum { CRT_LAMBERTIAN = 0, CRT_METAL = 1, CRT_DIELECTRIC = 2, CRT_DIFFUSELIGHT = 3 }; /* Material.cuh:6-12 */
enum { CRT_CONSTANT = 0, CRT_CHECKER = 1, CRT_IMAGE = 2 };                           /* Texture.cuh:7-12 */

/* One hittable of the scene list (reference: Hittable + Sphere / XYRect / XZRect / YZRect,
 * Hittable.cuh:47-67,69-126,128-294). 32 bytes. */
typedef struct crt_hittable {
    int32_t type;      /* CRT_SPHERE ... CRT_YZRECT */
    int32_t active;    /* Hittable::isActive; inactive entries are skipped by the BVH build (Hittable.cuh:311-320) */
    float cx, cy, cz;  /* Sphere::center / *Rect::center */
    float a;           /* Sphere::radius, or rect width  */
    float b;           /* unused for spheres, or rect height */
    int32_t material;  /* index into crt_scene_desc.materials */
} crt_hittable;

/* Material + Lambertian/Metal/Dielectric/DiffuseLight (Material.cuh:19-177). 16 bytes. */
typedef struct crt_material {
    int32_t type;            /* CRT_LAMBERTIAN ... */
    int32_t texture;         /* albedo texture index (ignored for dielectric) */
    float param;             /* Metal::fuzz (clamped to <=1 on upload, Material.cuh:71) | Dielectric::ir */
    int32_t light_intensity; /* DiffuseLight::light_intensity (an int in the reference, Material.cuh:152) */
} crt_material;

/* Texture + Constant/Checker/Image (Texture.cuh:18-109). A Checker's two Constant children are
 * folded into the record. 48 bytes. */
typedef struct crt_texture {
    int32_t type;     /* CRT_CONSTANT ... */
    int32_t image;    /* index into crt_scene_desc.images for CRT_IMAGE, else -1 */
    float color[3];   /* Constant::color */
    float odd[3];     /* Checker::odd->color  (picked when the sine product is < 0) */
    float even[3];    /* Checker::even->color */
    int32_t pad_;
} crt_texture;

/* Image (Texture.cuh:70-109): tightly packed RGB8, bytes_per_pixel is always 3 in the reference.
 * pixels == NULL reproduces the reference's "no data" cyan (Texture.cuh:85-86). */
typedef struct crt_image {
    const uint8_t* pixels;
    int32_t width, height;
} crt_image;

typedef struct crt_scene_desc {
    const crt_hittable* hittables; uint64_t n_hittables;
    const crt_material* materials; uint64_t n_materials;
    const crt_texture* textures;   uint64_t n_textures;
    const crt_image* images;       uint64_t n_images;
} crt_scene_desc;

/* Field-for-field InputStruct (SharedStructs.h:3-24). 72 bytes. fov is in RADIANS and, as in the
 * reference camera (Kernel.cu:130-146), slides the ray start along `orientation`. */
typedef struct crt_input_struct {
    float origin_x, origin_y, origin_z;
    float orientation_x, orientation_y, orientation_z;
    float up_x, up_y, up_z;   /* the caller passes normalize(cross(orientation, right)), CudaLayer.cpp:45-62 */
    float far_plane, near_plane, fov;
    float background_start[3];
    float background_end[3];
} crt_input_struct;

enum { CRT_VARIANT_AUTO = 0,
       CRT_VARIANT_MEGAKERNEL = 1, /* one thread = one pixel, all samples (the reference's schedule, Kernel.cu:137-148) */
       CRT_VARIANT_WAVEFRONT = 2,  /* generate -> extend -> shade kernels over an HBM pool of path slots */
       CRT_VARIANT_PIPELINE = 3 }; /* persistent CTAs: shared-memory path pool, material-binned queues */
enum { CRT_JITTER_RANDOM = 0,   /* Philox jitter in (0,1], like curand_uniform (Kernel.cu:139-140) */
       CRT_JITTER_CENTER = 1 }; /* xi1 = xi2 = 0.5: RNG-free ray generation for ID parity */
enum { CRT_FLAG_EXACT_MATH = 1u,      /* IEEE, unfused arithmetic in ray generation + primitive tests */
       CRT_FLAG_COUNT_TESTS = 2u,     /* also count box / primitive tests (slower; for the roofline figure) */
       CRT_FLAG_TRUNCATED_GRID = 4u   /* reproduce the reference's W/16 x H/16 launch grid (Kernel.cu:183-184) */ };

typedef struct crt_render_params {
    uint32_t spp_begin;   /* first sample index of this call (Philox counter word) */
    uint32_t spp_count;   /* samples per pixel added to the accumulation buffer by this call */
    uint32_t spp_stride;  /* sample index step: rank g of G renders spp_begin=g, spp_stride=G. 0 means 1 */
    uint32_t max_depth;   /* closest-hit queries per path at most (Kernel.cu:39) */
    uint64_t seed;        /* Philox key */
    int32_t variant;      /* CRT_VARIANT_* */
    int32_t jitter;       /* CRT_JITTER_* */
    uint32_t flags;       /* CRT_FLAG_* */
    uint32_t row_tiles;   /* 0 = every row. CRT_ROW_TILES(rank, world, rows_per_tile): render only the rows of the slabs (of
                           * rows_per_tile rows) number rank, rank + world, ... - image-tile sharding over GPUs: every pixel lives
                           * on one GPU, so the close-out needs no reduction (crt_resolve_rows into the root's frame) */
} crt_render_params;
#define CRT_ROW_TILES(rank, world, rows_per_tile) (((uint32_t)(rows_per_tile) << 16) | ((uint32_t)(world) << 8) | (uint32_t)(rank))
/* Host-only: how many of the `height` image rows the shard `row_tiles` holds (0 for a malformed value; height for 0). */
uint32_t crt_row_tiles_rows(uint32_t row_tiles, uint32_t height);

typedef struct crt_counters {
    uint64_t paths;        /* camera paths started */
    uint64_t rays;         /* closest-hit queries issued (primary + scattered, incl. the escaping one) */
    uint64_t box_tests;    /* only with CRT_FLAG_COUNT_TESTS */
    uint64_t prim_tests;   /* only with CRT_FLAG_COUNT_TESTS */
    uint64_t kernel_launches; /* launches of this library's kernels since the last reset */
} crt_counters;

typedef struct crt_bvh_stats {
    uint64_t n_prims, n_nodes, max_depth, smem_nodes;
    double sah_cost;
    double build_ms;
    uint64_t node_bytes, prim_bytes;
    uint64_t h2d_bytes; /* bytes copied host -> device by the last scene upload (nodes, primitives, materials, texels) */
    uint64_t builder;   /* CRT_BUILDER_HOST_SAH or _GPU_LBVH: which builder made the tree in use */
    uint64_t cache_hits; /* crt_scene_upload calls since the last real upload that found their scene already on the device
                            (then h2d_bytes is 0 and build_ms is the time of the content hash) */
} crt_bvh_stats;

/* Which builder crt_scene_upload / crt_scene_commit use. Both replace the BVHNode constructor of the reference
 * (Hittable.cuh:303-385), which runs on the host at start-up and after every GUI edit (CudaLayer.cpp:247-256,493-494).
 *   HOST_SAH  threaded binned-SAH builder (csrc/scene_pack.cpp): the best trees, 190 ms for 1 M spheres on 16 cores.
 *   GPU_LBVH  Morton-code linear BVH built entirely on the device (csrc/lbvh.cuh): a few ms for 1 M spheres.
 *   AUTO      GPU_LBVH from 65536 hittables up (CRT_GPU_BUILD_MIN_PRIMS), HOST_SAH below.
 * The environment variable CRT_BVH_BUILDER=host|lbvh overrides the default of new contexts. */
enum { CRT_BUILDER_AUTO = 0, CRT_BUILDER_HOST_SAH = 1, CRT_BUILDER_GPU_LBVH = 2 };
/* (a third device builder, PLOC clustering, was measured in rounds 1-2 and removed: DESIGN.md 3.3) */

typedef struct crt_context crt_context;

/* ---- context ---- */
crt_status crt_context_create(int device, crt_context** out);
void       crt_context_destroy(crt_context* ctx);
const char* crt_last_error_string(const crt_context* ctx); /* ctx may be NULL: process-wide last error */
int        crt_abi_version(void);
/* Run on an existing CUDA stream (cudaStream_t passed as void*); NULL = the context's own stream. */
crt_status crt_set_stream(crt_context* ctx, void* cuda_stream);
crt_status crt_synchronize(crt_context* ctx);

/* ---- scene (replaces CudaLayer::GenerateWorld + BVHNode ctor, CudaLayer.cpp:103-256, Hittable.cuh:303-385) ---- */
crt_status crt_set_bvh_builder(crt_context* ctx, int32_t builder);
/* Unchanged-scene cache (default on; CRT_SCENE_CACHE=0 turns it off for new contexts). crt_scene_upload hashes the description
 * it is given (hittables, materials, textures, texels: a threaded 64-bit content hash, ~3 ms for the 96 MB of a million spheres);
 * when that is the scene already on the device and no edit is pending, it returns without validating, building or copying -
 * the caller's frame loop can hand over its scene every frame, as CudaLayer.cpp:372-375 hands over its world pointer, and
 * pays for an upload only when something changed. crt_bvh_stats.cache_hits / h2d_bytes tell which happened. */
crt_status crt_set_scene_cache(crt_context* ctx, int32_t enabled);
/* CRT_ERR_INVALID (with the offending index in the message) for an ACTIVE hittable with an unknown type, a material /
 * texture / image index out of range, or a NaN / infinite coordinate or size: the reference would never hit such a primitive,
 * a BVH over it would be poisoned. Inactive hittables are not looked at. A negative sphere radius is legal here: |r| bounds
 * it and the normal flips (the hollow-sphere idiom). The reference gives such a sphere an inverted box (Hittable.cuh:112-116),
 * which AABB::Hit rejects (AABB.cuh:44-45), so there it is visible only through whatever sibling box happens to cover it -
 * tree-dependent behaviour that no other tree can reproduce; none of the reference's scenes or GUI ranges produce one. */
crt_status crt_scene_upload(crt_context* ctx, const crt_scene_desc* scene);
/* Walk the reference's own pointer graph: `world` is the Hittable of type BVHNODE that
 * LaunchKernel receives (Kernel.cu:178-181). Host-readable (managed or host) memory. */
crt_status crt_scene_from_reference(crt_context* ctx, const void* world);
/* ---- scene edits (replaces the reference's rebuild-after-every-edit: `bvh_node->Destroy()` + `new BVHNode(m_List, 0, n)`,
 * CudaLayer.cpp:493-494,967-968,1375-1377). Edits are staged on the host copy of the uploaded scene; crt_scene_commit applies
 * them. Moving / resizing a primitive, giving it another material, or editing a material keeps the tree and only REFITS the
 * boxes on the GPU; (de)activating a primitive (Hittable::isActive, CudaLayer.cpp:1372-1387) or changing its kind rebuilds
 * the tree on the host (image textures are not re-uploaded). *out_rebuilt tells which happened. ---- */
crt_status crt_scene_edit_hittable(crt_context* ctx, uint64_t index, const crt_hittable* h);
crt_status crt_scene_edit_material(crt_context* ctx, uint64_t index, const crt_material* m);
crt_status crt_scene_commit(crt_context* ctx, int force_rebuild, int* out_rebuilt);
crt_status crt_bvh_get_stats(const crt_context* ctx, crt_bvh_stats* out);
/* Host-only: build and pack the BVH of `scene` exactly as crt_scene_upload does (no GPU needed), check its invariants (every
 * primitive inside the boxes of all its ancestors, every active hittable referenced by exactly one leaf, child indices in
 * range) and report the statistics. CRT_ERR_INVALID with a message if an invariant is broken. */
crt_status crt_bvh_build_check(const crt_scene_desc* scene, crt_bvh_stats* out);

/* ---- camera ---- */
crt_status crt_set_camera(crt_context* ctx, const crt_input_struct* inputs);
/* Thin-lens extension (not in the reference; north_star asks for defocus). aperture = lens DIAMETER;
 * aperture 0 restores the reference camera bit for bit. */
crt_status crt_set_lens(crt_context* ctx, float aperture, float focus_dist);

/* ---- framebuffer: progressive accumulation buffer ----
 * Device layout: four int64 per pixel {R, G, B, samples}; R/G/B are the radiance SUM in Q31.32 fixed point (value * 2^32).
 * Every FP32 sample is added exactly, and integer addition is associative, so the buffer is bit-identical for any
 * schedule of the same (pixel, sample) set: megakernel or wavefront, any atomic order, any number of GPUs whose buffers
 * are summed (NCCL sum over int64). crt_read_accum converts to float4 {sum r, g, b, samples} (one rounding per value). */
crt_status crt_framebuffer_resize(crt_context* ctx, uint32_t width, uint32_t height); /* also clears */
crt_status crt_accum_clear(crt_context* ctx);
/* Use a caller-owned, 32-byte-aligned DEVICE buffer of width*height*4 int64 (e.g. a torch tensor that is then
 * NCCL-reduced). NULL = back to the internal buffer. */
crt_status crt_set_accum_device_buffer(crt_context* ctx, void* device_int64x4);
void*      crt_accum_device_ptr(crt_context* ctx);

/* ---- render (replaces LaunchKernel, Kernel.cu:102-158,178-191). The megakernel is asynchronous on the context stream;
 * the wavefront variant synchronises the stream internally (it reads its active-path count back between batches of
 * extend/shade iterations). Both variants add the SAME samples: Philox is keyed by (seed; pixel, sample, bounce). ---- */
crt_status crt_render(crt_context* ctx, const crt_render_params* params);
/* mean -> 255*sqrt -> clamp -> pack A<<24|B<<16|G<<8|R (Kernel.cu:12-19,151-157). out is HOST memory. Synchronises. */
crt_status crt_resolve(crt_context* ctx, uint32_t* out_rgba8);
crt_status crt_resolve_device(crt_context* ctx, void* device_out_rgba8); /* async */
/* The same for the pixels a CRT_FLAG_TRUNCATED_GRID render covers, (W / 16 * 16) x (H / 16 * 16): the other pixels of the
 * buffer are left untouched, as the reference's launch grid leaves them (Kernel.cu:183-184 with :117-119). Async. */
crt_status crt_resolve_device_truncated(crt_context* ctx, void* device_out_rgba8);
crt_status crt_read_accum(crt_context* ctx, float* out_float4);          /* width*height*4 floats, synchronises */
/* crt_read_accum for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc). NO COUNTERPART IN THE REFERENCE. Converts n_pixels entries of the
 * int64x4 accumulation buffer at device_accum_int64x4 (device memory of ctx's GPU, 16-byte aligned; e.g. crt_accum_device_ptr
 * or a caller-owned buffer, or a sub-range of one) to float4 {sum r, g, b, samples} in HOST memory at out_float4, exactly as
 * crt_read_accum does, and synchronises ctx's stream. The staging array is a static __device__ array inside libcrt.so of
 * CRT_STATIC_ACCUM_MAX_PIXELS float4 (4 MiB on every GPU that loads the library, until the process exits): the call uses no
 * cudaMalloc, cudaFree or allocating library routine. More pixels than that: CRT_ERR_INVALID (read the buffer in pieces).
 * All calls share the one array, which keeps the last call's contents (nothing in it needs clearing: entries [0, n_pixels)
 * are overwritten before they are read). Calls that OVERLAP on one GPU - different streams or host threads - race on it. */
#define CRT_STATIC_ACCUM_MAX_PIXELS 262144
crt_status crt_read_accum_static(crt_context* ctx, const void* device_accum_int64x4, uint32_t n_pixels, float* out_float4);
/* Resolve only the rows of a row-tile shard (row_tiles as in crt_render_params; 0 = all rows) from this context's accumulation
 * buffer into the frame at device_out_rgba8 (pixel 0 of a full frame; may be a peer mapping of another GPU's frame). Async. */
crt_status crt_resolve_rows(crt_context* ctx, uint32_t row_tiles, void* device_out_rgba8);
/* Copy the context's own RGBA8 frame (what crt_resolve or a peer's crt_resolve_peers wrote into it) to HOST memory. Synchronises. */
crt_status crt_read_rgba(crt_context* ctx, uint32_t* out_rgba8);
void*      crt_rgba_device_ptr(crt_context* ctx);

/* ---- multi-GPU close-out over peer memory (NVLink; one process per GPU). The reference has no multi-GPU path; this replaces
 * "reduce every rank's accumulation buffer to the root, resolve there" by ONE kernel per rank: for the rows a rank owns it
 * reads the buffers of all ranks through peer mappings, adds them (int64: exact, the same bits as the NCCL sum), resolves
 * (Kernel.cu:12-19,151-157) and stores the packed pixels straight into the root's frame. Per pixel 32 bytes per peer travel to
 * the owner of the row and 4 bytes to the root, instead of 32 bytes per peer into the root.
 *   crt_ipc_export   a 64-byte CUDA IPC handle of the context's OWN accumulation buffer or RGBA8 frame (send it to the other
 *                    processes by any means);
 *   crt_ipc_open     map a buffer exported by ANOTHER process on this context's device; crt_ipc_close / context destruction unmaps;
 *   crt_resolve_peers  rows [row_begin, row_end): peer_accum[g] are device pointers valid on this device (peer mappings or
 *                    local buffers, at most 16), device_out_rgba8 points at pixel 0 of the destination frame (may be a peer
 *                    mapping), device_out_sum (may be NULL) receives the summed rows in the accumulation layout. Asynchronous on
 *                    the context stream. The CALLER orders it after every rank's render and before the next clear (a
 *                    stream-ordered collective on the same stream, e.g. a one-element all-reduce, is enough).
 * crt_framebuffer_resize with a new size frees the exported buffers: peers must close their mappings first and the handles must
 * be exchanged again afterwards. ---- */
#define CRT_IPC_HANDLE_BYTES 64
enum { CRT_BUFFER_ACCUM = 0, CRT_BUFFER_RGBA = 1 };
/* All GPUs in ONE process (one context per device, e.g. one host thread per GPU; examples/multigpu_render.cpp): no IPC is
 * needed - let this context's device address peer_device's memory, then pass crt_accum_device_ptr / crt_rgba_device_ptr of
 * the context on peer_device to crt_resolve_peers / crt_resolve_rows here. Idempotent. CRT_ERR_UNSUPPORTED without a P2P path. */
crt_status crt_peer_enable(crt_context* ctx, int32_t peer_device);
crt_status crt_ipc_export(crt_context* ctx, int32_t which, uint8_t out_handle[CRT_IPC_HANDLE_BYTES]);
crt_status crt_ipc_open(crt_context* ctx, const uint8_t handle[CRT_IPC_HANDLE_BYTES], void** out_device_ptr);
crt_status crt_ipc_close(crt_context* ctx, void* device_ptr);
crt_status crt_resolve_peers(crt_context* ctx, const void* const* peer_accum, int32_t n_peers, uint32_t row_begin, uint32_t row_end,
                             void* device_out_rgba8, void* device_out_sum);

/* Closest-hit hittable index (index into the uploaded list, -1 = miss) of one primary ray per pixel.
 * Always exact math; jitter = CRT_JITTER_*. out is HOST memory, width*height int32. */
crt_status crt_primary_ids(crt_context* ctx, int32_t jitter, uint64_t seed, int32_t* out_ids);
crt_status crt_get_counters(crt_context* ctx, crt_counters* out);   /* synchronises */
crt_status crt_reset_counters(crt_context* ctx);
/* GPU milliseconds between the CUDA events that bracketed the last crt_render call (synchronises). */
crt_status crt_last_render_ms(crt_context* ctx, float* out_ms);
/* The kernel variant the last crt_render call ran (CRT_VARIANT_MEGAKERNEL / _WAVEFRONT / _PIPELINE; 0 before the first call).
 * With CRT_VARIANT_AUTO on a scene that both the megakernel and the pipelined kernel can stage completely (up to 2048
 * primitives), a call of at least 16 samples on a frame large enough for a few samples to be representative (1.5 M paths in
 * the timed part) renders its first samples with both (the pipelined kernel with each of its two refill-parameter pairs),
 * times them and gives the rest to the fastest (the variants write identical bits, so the frame does not depend on the choice); the choice is kept until the tree, the
 * frame size or the depth changes. CRT_AUTO_CALIBRATE=0 keeps the static rule (pipeline from 256 primitives up). The reference
 * has one kernel (Kernel.cu:102): no counterpart. */
int32_t crt_last_variant(crt_context* ctx);

/* ---- synthetic scenes of BASELINE.json, splitmix64-seeded; host-only, no GPU needed ---- */
enum { CRT_SCENE_RTIOW = 1, CRT_SCENE_TEXTURED = 2, CRT_SCENE_GRID_1M = 3, CRT_SCENE_STRESS_10K = 4,
       CRT_SCENE_REFERENCE_DEFAULT = 5 /* CudaLayer.cpp:103-256: XZRect checker ground + 16 spheres */ };
typedef struct crt_scene_builder crt_scene_builder;
crt_status crt_scene_generate(int32_t kind, uint64_t seed, uint64_t size_hint, crt_scene_builder** out);
const crt_scene_desc* crt_scene_builder_desc(const crt_scene_builder* b);
/* Default camera of the generated scene for an image of the given aspect. */
crt_status crt_scene_builder_camera(const crt_scene_builder* b, uint32_t width, uint32_t height, crt_input_struct* out);
void       crt_scene_builder_free(crt_scene_builder* b);
/* Reference host-side camera mapping (CudaLayer.cpp:45-62): position/orientation/up/fov(deg) -> InputStruct. */
crt_status crt_camera_look(const float position[3], const float orientation[3], const float world_up[3],
                           float fov_degrees, float near_plane, float far_plane,
                           const float background_start[3], const float background_end[3], crt_input_struct* out);

/* ---- scripted fly camera: the reference's interactive camera (Renderer/Camera.cpp:37-118, Camera.h:6-7,39-60) without a
 * window. One crt_flycam_step = one Camera::Inputs call (one frame): `keys` says which keys are held, (mouse_dx, mouse_dy) is
 * the cursor movement since the last frame while the right button is held (0,0 otherwise), `scroll` the wheel offset
 * (Camera::ProcessMouseScroll, :28-35). crt_flycam_inputs is the mapping CudaLayer.cpp:413-431 applies after every call. ---- */
enum { CRT_KEY_W = 1, CRT_KEY_A = 2, CRT_KEY_S = 4, CRT_KEY_D = 8, CRT_KEY_SPACE = 16, CRT_KEY_LCTRL = 32, CRT_KEY_LSHIFT = 64,
       CRT_KEY_C = 128 /* reset the position to (0, 2, 12) */ };
typedef struct crt_flycam {
    float position[3], orientation[3], up[3];
    float far_plane, near_plane, fov_degrees;
    float speed, sensitivity, yaw, pitch;
} crt_flycam;
crt_status crt_flycam_init(crt_flycam* cam, const float position[3]); /* Camera ctor + defaults: looks down -z, fov 45 */
/* returns 1 in *moved if the camera changed (the caller then clears the accumulation buffer), else 0 */
crt_status crt_flycam_step(crt_flycam* cam, uint32_t keys, float mouse_dx, float mouse_dy, float scroll, int* moved);
crt_status crt_flycam_inputs(const crt_flycam* cam, const float background_start[3], const float background_end[3],
                             crt_input_struct* out);

/* ---- headless framebuffer writer (replaces the GL blit, CudaLayer.cpp:379-386) ---- */
crt_status crt_write_ppm(const char* path, const uint32_t* rgba8, uint32_t width, uint32_t height, int flip_rows);
/* The same frame as an 8-bit RGB PNG (stored deflate blocks: no compression, readable everywhere). */
crt_status crt_write_png(const char* path, const uint32_t* rgba8, uint32_t width, uint32_t height, int flip_rows);
/* Binary PPM (P6, maxval 255) -> tightly packed RGB8, the layout Image expects (Texture.cuh:76; the reference decodes JPEGs with
 * stb_image, Utils/RawStbImage.h:11-22, which is third-party code this library does not carry). *out_pixels is malloc'ed;
 * release it with crt_free. */
crt_status crt_load_ppm(const char* path, uint8_t** out_pixels, int32_t* out_width, int32_t* out_height);
/* JPEG in memory -> tightly packed RGB8 (csrc/jpeg_decode.cpp). Replaces stbi_load(filename, &w, &h, &n, 0) of LoadImage
 * (Utils/RawStbImage.h:11-22, CudaLayer.cpp:896): baseline, extended-sequential and progressive Huffman JPEG, 8 bit, 1 or 3
 * components, any integer sampling ratio, restart intervals; reconstruction arithmetic (integer IDCT, triangle chroma
 * upsampling, fixed-point YCbCr matrix) bit-exact with the decoder the reference vendors. *out_file_channels (optional)
 * receives the file's component count (stb's n); the output is RGB8 in both cases, a grey image is replicated into the three
 * channels (the reference would read a 1-channel buffer with a 3-byte stride, Texture.cuh:76,102).
 * CRT_ERR_UNSUPPORTED: arithmetic coding, 12-bit, lossless/hierarchical, 4 components. Release with crt_free. */
crt_status crt_decode_jpeg(const uint8_t* data, uint64_t n_bytes, uint8_t** out_pixels, int32_t* out_width, int32_t* out_height,
                           int32_t* out_file_channels);
/* File loader for image textures: JPEG or binary PPM by magic number. */
crt_status crt_load_image(const char* path, uint8_t** out_pixels, int32_t* out_width, int32_t* out_height,
                          int32_t* out_file_channels);
void       crt_free(void* p);

#ifdef __cplusplus
}
#endif
#endif /* CRT_H */
