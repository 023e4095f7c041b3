// scenes.cpp — deterministic synthetic scenes of BASELINE.json and the reference's
// start-up scene, emitted as the flat crt_scene_desc of include/crt.h. Host-only code.
//
// The reference builds its scene with unseeded glibc rand() through the RND macro (Math.cuh:12,
// CudaLayer.cpp:200-219), which is not portable; every scene here draws from splitmix64 instead,
// xi = (x >> 40) * 2^-24 in [0,1), in the loop order documented per generator.

#include "crt.h"
#include "crt_internal.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

struct crt_scene_builder
{
    std::vector<crt_hittable> hittables;
    std::vector<crt_material> materials;
    std::vector<crt_texture> textures;
    std::vector<crt_image> images;
    std::vector<std::vector<uint8_t>> image_store;
    crt_scene_desc desc;
    // default camera (reference Camera fields, Renderer/Camera.h:39-45)
    float cam_pos[3] = {0.f, 2.f, 12.f};
    float cam_orient[3] = {0.f, 0.f, -1.f};
    float cam_fov_deg = 45.f;
    float cam_near = 0.1f, cam_far = 10.f;
    float bg0[3] = {1.f, 1.f, 1.f}, bg1[3] = {0.5f, 0.7f, 1.0f}; // CudaLayer.h:143-144
};

namespace {

struct SplitMix
{
    uint64_t s;
    explicit SplitMix(uint64_t seed) : s(seed) {}
    uint64_t next()
    {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    float xi() { return (float)(next() >> 40) * (1.0f / 16777216.0f); }
};

int AddConstant(crt_scene_builder* b, float r, float g, float bl)
{
    crt_texture t;
    memset(&t, 0, sizeof(t));
    t.type = CRT_CONSTANT;
    t.image = -1;
    t.color[0] = r; t.color[1] = g; t.color[2] = bl;
    b->textures.push_back(t);
    return (int)b->textures.size() - 1;
}

int AddChecker(crt_scene_builder* b, const float odd[3], const float even[3])
{
    crt_texture t;
    memset(&t, 0, sizeof(t));
    t.type = CRT_CHECKER;
    t.image = -1;
    for (int i = 0; i < 3; i++) { t.odd[i] = odd[i]; t.even[i] = even[i]; }
    b->textures.push_back(t);
    return (int)b->textures.size() - 1;
}

int AddImageTexture(crt_scene_builder* b, int image)
{
    crt_texture t;
    memset(&t, 0, sizeof(t));
    t.type = CRT_IMAGE;
    t.image = image;
    b->textures.push_back(t);
    return (int)b->textures.size() - 1;
}

int AddMaterial(crt_scene_builder* b, int type, int tex, float param, int intensity)
{
    crt_material m;
    m.type = type; m.texture = tex; m.param = param; m.light_intensity = intensity;
    b->materials.push_back(m);
    return (int)b->materials.size() - 1;
}

void AddHittable(crt_scene_builder* b, int type, float cx, float cy, float cz, float a, float bb, int mat)
{
    crt_hittable h;
    h.type = type; h.active = 1; h.cx = cx; h.cy = cy; h.cz = cz; h.a = a; h.b = bb; h.material = mat;
    b->hittables.push_back(h);
}

// Integer-only procedural RGB8 image (no libm, so identical on every box). Stands in for the reference's
// JPEG assets (assets/textures/8k_earth_nightmap.jpg 8192x4096x3 etc.), which are not on the GPU box.
int AddProceduralImage(crt_scene_builder* b, int w, int h, uint32_t salt)
{
    std::vector<uint8_t> px((size_t)w * h * 3);
    for (int y = 0; y < h; y++) {
        for (int x = 0; x < w; x++) {
            uint32_t bx = (uint32_t)x * 64u / (uint32_t)w, by = (uint32_t)y * 32u / (uint32_t)h;
            uint32_t hsh = (bx * 0x9E3779B1u) ^ (by * 0x85EBCA77u) ^ (salt * 0xC2B2AE3Du);
            hsh ^= hsh >> 15; hsh *= 0x2C1B3C6Du; hsh ^= hsh >> 12; hsh *= 0x297A2D39u; hsh ^= hsh >> 15;
            uint32_t fx = ((uint32_t)x * 1024u / (uint32_t)w) & 255u, fy = ((uint32_t)y * 512u / (uint32_t)h) & 255u;
            uint8_t* p = &px[((size_t)y * w + x) * 3];
            p[0] = (uint8_t)(48 + ((hsh & 0xFF) * 3 + fx) / 5);
            p[1] = (uint8_t)(48 + (((hsh >> 8) & 0xFF) * 3 + fy) / 5);
            p[2] = (uint8_t)(48 + (((hsh >> 16) & 0xFF) * 3 + ((fx ^ fy) & 255u)) / 5);
        }
    }
    b->image_store.push_back(std::move(px));
    crt_image im;
    im.pixels = b->image_store.back().data();
    im.width = w; im.height = h;
    b->images.push_back(im);
    return (int)b->images.size() - 1;
}

// The reference's own texture files when the caller points CRT_ASSET_DIR at its assets/textures directory (the JPEG decoder
// of jpeg_decode.cpp stands in for stbi_load, RawStbImage.h:11-22); -1 if the variable is unset or the file cannot be decoded.
int AddImageFromAssetDir(crt_scene_builder* b, const char* file)
{
    const char* dir = getenv("CRT_ASSET_DIR");
    if (!dir || !*dir) return -1;
    const std::string path = std::string(dir) + "/" + file;
    uint8_t* px = nullptr;
    int32_t w = 0, h = 0, n = 0;
    if (crt_load_image(path.c_str(), &px, &w, &h, &n) != CRT_OK) return -1;
    b->image_store.emplace_back(px, px + (size_t)w * h * 3);
    crt_free(px);
    crt_image im;
    im.pixels = b->image_store.back().data();
    im.width = w; im.height = h;
    b->images.push_back(im);
    return (int)b->images.size() - 1;
}

void LookAt(crt_scene_builder* b, float px, float py, float pz, float tx, float ty, float tz)
{
    float d[3] = {tx - px, ty - py, tz - pz};
    float inv = 1.0f / sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    b->cam_pos[0] = px; b->cam_pos[1] = py; b->cam_pos[2] = pz;
    for (int i = 0; i < 3; i++) b->cam_orient[i] = d[i] * inv;
}

// C1/C2. Draw order: for a in [-11,11) for b in [-11,11): m, cx-jitter, cz-jitter, then material draws.
void GenRtiow(crt_scene_builder* B, uint64_t seed)
{
    SplitMix R(seed);
    AddHittable(B, CRT_SPHERE, 0.f, -1000.f, 0.f, 1000.f, 0.f, AddMaterial(B, CRT_LAMBERTIAN, AddConstant(B, 0.5f, 0.5f, 0.5f), 0.f, 0));
    for (int a = -11; a < 11; a++) {
        for (int b = -11; b < 11; b++) {
            float m = R.xi();
            float cx = (float)a + 0.9f * R.xi(), cy = 0.2f, cz = (float)b + 0.9f * R.xi();
            float dx = cx - 4.0f, dy = cy - 0.2f, dz = cz;
            if (sqrtf(dx * dx + dy * dy + dz * dz) <= 0.9f)
                continue;
            int mat;
            if (m < 0.8f) {
                float r = R.xi() * R.xi(), g = R.xi() * R.xi(), bl = R.xi() * R.xi();
                mat = AddMaterial(B, CRT_LAMBERTIAN, AddConstant(B, r, g, bl), 0.f, 0);
            }
            else if (m < 0.95f) {
                float r = 0.5f * (1.0f + R.xi()), g = 0.5f * (1.0f + R.xi()), bl = 0.5f * (1.0f + R.xi());
                float fuzz = 0.5f * R.xi();
                mat = AddMaterial(B, CRT_METAL, AddConstant(B, r, g, bl), fuzz, 0);
            }
            else {
                mat = AddMaterial(B, CRT_DIELECTRIC, -1, 1.5f, 0);
            }
            AddHittable(B, CRT_SPHERE, cx, cy, cz, 0.2f, 0.f, mat);
        }
    }
    AddHittable(B, CRT_SPHERE, 0.f, 1.f, 0.f, 1.0f, 0.f, AddMaterial(B, CRT_DIELECTRIC, -1, 1.5f, 0));
    AddHittable(B, CRT_SPHERE, -4.f, 1.f, 0.f, 1.0f, 0.f, AddMaterial(B, CRT_LAMBERTIAN, AddConstant(B, 0.4f, 0.2f, 0.1f), 0.f, 0));
    AddHittable(B, CRT_SPHERE, 4.f, 1.f, 0.f, 1.0f, 0.f, AddMaterial(B, CRT_METAL, AddConstant(B, 0.7f, 0.6f, 0.5f), 0.0f, 0));
    LookAt(B, 13.f, 2.f, 3.f, 0.f, 0.f, 0.f);
    B->cam_fov_deg = 35.f;
}

// C3. size_hint = width of the big ("earth") image, default 8192 (height = width/2); the small image is
// 1241x1241 scaled by the same factor. 8x8 grid cycles Metal-Checker / Lambertian-Image / Lambertian-Constant / Dielectric.
void GenTextured(crt_scene_builder* B, uint64_t seed, uint64_t size_hint)
{
    SplitMix R(seed);
    int big_w = size_hint ? (int)size_hint : 8192;
    int small = (int)((int64_t)1241 * big_w / 8192);
    if (small < 8) small = 8;
    const float odd[3] = {0.2f, 0.3f, 0.1f}, even[3] = {0.9f, 0.9f, 0.9f}; // CudaLayer.cpp:157-160
    B->image_store.reserve(2);
    // SURVEY.md 8d names assets/textures/8k_earth_nightmap.jpg and UTH-logo-english.jpg; they are used when present
    // (CRT_ASSET_DIR, default size only), the procedural stand-ins of the same dimensions otherwise
    int img_big = size_hint ? -1 : AddImageFromAssetDir(B, "8k_earth_nightmap.jpg");
    if (img_big < 0) img_big = AddProceduralImage(B, big_w, big_w / 2, 1);
    int img_small = size_hint ? -1 : AddImageFromAssetDir(B, "UTH-logo-english.jpg");
    if (img_small < 0) img_small = AddProceduralImage(B, small, small, 2);
    AddHittable(B, CRT_SPHERE, 0.f, -1000.f, 0.f, 1000.f, 0.f, AddMaterial(B, CRT_LAMBERTIAN, AddChecker(B, odd, even), 0.f, 0));
    AddHittable(B, CRT_SPHERE, 0.f, 2.f, 0.f, 2.0f, 0.f, AddMaterial(B, CRT_LAMBERTIAN, AddImageTexture(B, img_big), 0.f, 0));
    int k = 0;
    for (int i = 0; i < 8; i++) {
        for (int j = 0; j < 8; j++, k++) {
            float cx = ((float)i - 3.5f) * 1.1f, cy = 0.3f, cz = 2.6f + (float)j * 0.9f;
            int mat;
            switch (k & 3) {
            case 0: mat = AddMaterial(B, CRT_METAL, AddChecker(B, odd, even), 0.25f * R.xi(), 0); break;
            case 1: mat = AddMaterial(B, CRT_LAMBERTIAN, AddImageTexture(B, img_small), 0.f, 0); break;
            case 2: {
                float r = R.xi(), g = R.xi(), bl = R.xi();
                mat = AddMaterial(B, CRT_LAMBERTIAN, AddConstant(B, r, g, bl), 0.f, 0);
                break;
            }
            default: mat = AddMaterial(B, CRT_DIELECTRIC, -1, 1.5f, 0); break;
            }
            AddHittable(B, CRT_SPHERE, cx, cy, cz, 0.3f, 0.f, mat);
        }
    }
    LookAt(B, 0.f, 3.f, 12.f, 0.f, 2.f, 0.f);
    B->cam_fov_deg = 45.f;
}

// C4. size_hint = grid side, default 1000. Draw order per cell (i outer, j inner): r, cx-jitter, cz-jitter, m, material draws.
void GenGrid(crt_scene_builder* B, uint64_t seed, uint64_t size_hint)
{
    SplitMix R(seed);
    int n = size_hint ? (int)size_hint : 1000;
    float half = (float)(n / 2);
    B->hittables.reserve((size_t)n * n + 1);
    B->materials.reserve((size_t)n * n + 1);
    B->textures.reserve((size_t)n * n + 1);
    AddHittable(B, CRT_SPHERE, 0.f, -1000.f, 0.f, 1000.f, 0.f, AddMaterial(B, CRT_LAMBERTIAN, AddConstant(B, 0.5f, 0.5f, 0.5f), 0.f, 0));
    for (int i = 0; i < n; i++) {
        for (int j = 0; j < n; j++) {
            float r = 0.15f + 0.2f * R.xi();
            float cx = ((float)i - half) + 0.8f * R.xi();
            float cz = ((float)j - half) + 0.8f * R.xi();
            float m = R.xi();
            int mat;
            if (m < 0.7f) {
                float cr = R.xi() * R.xi(), cg = R.xi() * R.xi(), cb = R.xi() * R.xi();
                mat = AddMaterial(B, CRT_LAMBERTIAN, AddConstant(B, cr, cg, cb), 0.f, 0);
            }
            else if (m < 0.9f) {
                float cr = 0.5f * (1.0f + R.xi()), cg = 0.5f * (1.0f + R.xi()), cb = 0.5f * (1.0f + R.xi());
                float fuzz = 0.5f * R.xi();
                mat = AddMaterial(B, CRT_METAL, AddConstant(B, cr, cg, cb), fuzz, 0);
            }
            else {
                mat = AddMaterial(B, CRT_DIELECTRIC, -1, 1.5f, 0);
            }
            AddHittable(B, CRT_SPHERE, cx, r, cz, r, 0.f, mat);
        }
    }
    float s = (float)n / 1000.0f;
    LookAt(B, 0.f, 40.f * s, 120.f * s, 0.f, 0.f, 0.f);
    B->cam_fov_deg = 45.f;
}

// C5. size_hint = grid side, default 100. 50 % Dielectric 1.5 / 50 % Metal (fuzz 0.3 xi). Draw order: m, then metal draws.
void GenStress(crt_scene_builder* B, uint64_t seed, uint64_t size_hint)
{
    SplitMix R(seed);
    int n = size_hint ? (int)size_hint : 100;
    float half = (float)(n / 2);
    AddHittable(B, CRT_SPHERE, 0.f, -1000.f, 0.f, 1000.f, 0.f, AddMaterial(B, CRT_LAMBERTIAN, AddConstant(B, 0.5f, 0.5f, 0.5f), 0.f, 0));
    for (int i = 0; i < n; i++) {
        for (int j = 0; j < n; j++) {
            float m = R.xi();
            int mat;
            if (m < 0.5f) {
                mat = AddMaterial(B, CRT_DIELECTRIC, -1, 1.5f, 0);
            }
            else {
                float cr = 0.5f * (1.0f + R.xi()), cg = 0.5f * (1.0f + R.xi()), cb = 0.5f * (1.0f + R.xi());
                float fuzz = 0.3f * R.xi();
                mat = AddMaterial(B, CRT_METAL, AddConstant(B, cr, cg, cb), fuzz, 0);
            }
            AddHittable(B, CRT_SPHERE, ((float)i - half) + 0.5f, 0.4f, ((float)j - half) + 0.5f, 0.4f, 0.f, mat);
        }
    }
    float s = (float)n / 100.0f;
    LookAt(B, 0.f, 12.f * s, 40.f * s, 0.f, 0.f, 0.f);
    B->cam_fov_deg = 45.f;
}

// The reference's start-up scene, CudaLayer.cpp:103-256 (RND replaced by splitmix64, same draw order).
void GenReferenceDefault(crt_scene_builder* B, uint64_t seed)
{
    SplitMix R(seed);
    const float odd[3] = {0.2f, 0.3f, 0.1f}, even[3] = {0.9f, 0.9f, 0.9f};
    AddHittable(B, CRT_XZRECT, 0.f, -0.5f, 0.f, 1000.f, 1000.f, AddMaterial(B, CRT_LAMBERTIAN, AddChecker(B, odd, even), 0.f, 0));
    for (int a = -2; a < 2; a++) {
        for (int b = -2; b < 2; b++) {
            float choose = R.xi();
            float cx = (float)a + R.xi(), cz = (float)b + R.xi();
            if (choose < 0.5f) {
                float r = R.xi() * R.xi(), g = R.xi() * R.xi(), bl = R.xi() * R.xi();
                AddHittable(B, CRT_SPHERE, cx, 0.2f, cz, 0.2f, 0.f, AddMaterial(B, CRT_LAMBERTIAN, AddConstant(B, r, g, bl), 0.f, 0));
            }
            else if (choose < 0.8f) {
                float r = 0.5f * (1.0f + R.xi()), g = 0.5f * (1.0f + R.xi()), bl = 0.5f * (1.0f + R.xi());
                float fuzz = 0.5f * R.xi();
                AddHittable(B, CRT_SPHERE, cx, 0.2f, cz, 0.2f, 0.f, AddMaterial(B, CRT_METAL, AddConstant(B, r, g, bl), fuzz, 0));
            }
            else if (choose < 0.9f) {
                AddHittable(B, CRT_SPHERE, cx, 0.2f, cz, 0.3f, 0.f, AddMaterial(B, CRT_DIELECTRIC, -1, 1.5f, 0));
            }
            else {
                AddHittable(B, CRT_SPHERE, cx, 0.2f, cz, 0.5f, 0.f, AddMaterial(B, CRT_DIFFUSELIGHT, AddConstant(B, 1.f, 1.f, 1.f), 0.f, 3));
            }
        }
    }
    // Camera(W, H, (0,2,12)), orientation (0,0,-1), fov 45 (CudaLayer.cpp:43, Camera.h:39-45): the builder defaults.
}

} // namespace

extern "C" {

crt_status crt_scene_generate(int32_t kind, uint64_t seed, uint64_t size_hint, crt_scene_builder** out)
{
    if (!out)
        return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_scene_generate: out is NULL");
    crt_scene_builder* B = new (std::nothrow) crt_scene_builder();
    if (!B)
        return crt_set_error(nullptr, CRT_ERR_NOMEM, "crt_scene_generate: out of memory");
    switch (kind) {
    case CRT_SCENE_RTIOW: GenRtiow(B, seed); break;
    case CRT_SCENE_TEXTURED: GenTextured(B, seed, size_hint); break;
    case CRT_SCENE_GRID_1M: GenGrid(B, seed, size_hint); break;
    case CRT_SCENE_STRESS_10K: GenStress(B, seed, size_hint); break;
    case CRT_SCENE_REFERENCE_DEFAULT: GenReferenceDefault(B, seed); break;
    default:
        delete B;
        return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_scene_generate: unknown scene kind");
    }
    B->desc.hittables = B->hittables.data(); B->desc.n_hittables = B->hittables.size();
    B->desc.materials = B->materials.data(); B->desc.n_materials = B->materials.size();
    B->desc.textures = B->textures.data();   B->desc.n_textures = B->textures.size();
    B->desc.images = B->images.data();       B->desc.n_images = B->images.size();
    *out = B;
    return CRT_OK;
}

const crt_scene_desc* crt_scene_builder_desc(const crt_scene_builder* b)
{
    return b ? &b->desc : nullptr;
}

void crt_scene_builder_free(crt_scene_builder* b)
{
    delete b;
}

// CudaLayer.cpp:45-62: rightV = normalize(cross(orientation, up)); upV = normalize(cross(orientation, rightV));
// fov = radians(fov_deg).
crt_status crt_camera_look(const float position[3], const float orientation[3], const float world_up[3], float fov_degrees,
                           float near_plane, float far_plane, const float background_start[3],
                           const float background_end[3], crt_input_struct* out)
{
    if (!position || !orientation || !world_up || !out)
        return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_camera_look: NULL argument");
    auto cross = [](const float* a, const float* b, float* c) {
        c[0] = a[1] * b[2] - a[2] * b[1];
        c[1] = a[2] * b[0] - a[0] * b[2];
        c[2] = a[0] * b[1] - a[1] * b[0];
    };
    auto normalize = [](float* v) {
        float inv = 1.0f / sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        v[0] *= inv; v[1] *= inv; v[2] *= inv;
    };
    float right[3], up[3];
    cross(orientation, world_up, right);
    normalize(right);
    cross(orientation, right, up);
    normalize(up);
    out->origin_x = position[0]; out->origin_y = position[1]; out->origin_z = position[2];
    out->orientation_x = orientation[0]; out->orientation_y = orientation[1]; out->orientation_z = orientation[2];
    out->up_x = up[0]; out->up_y = up[1]; out->up_z = up[2];
    out->far_plane = far_plane;
    out->near_plane = near_plane;
    out->fov = fov_degrees * 0.01745329251994329576923690768489f;
    for (int i = 0; i < 3; i++) {
        out->background_start[i] = background_start ? background_start[i] : 1.0f;
        out->background_end[i] = background_end ? background_end[i] : (i == 0 ? 0.5f : i == 1 ? 0.7f : 1.0f);
    }
    return CRT_OK;
}

crt_status crt_scene_builder_camera(const crt_scene_builder* b, uint32_t width, uint32_t height, crt_input_struct* out)
{
    (void)width; (void)height; // the reference camera divides both u and v by the width (Kernel.cu:141-142)
    if (!b || !out)
        return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_scene_builder_camera: NULL argument");
    const float up[3] = {0.f, 1.f, 0.f};
    return crt_camera_look(b->cam_pos, b->cam_orient, up, b->cam_fov_deg, b->cam_near, b->cam_far, b->bg0, b->bg1, out);
}

} // extern "C"

// ------------------------------------------------------------------------------------------
// Scripted fly camera (Renderer/Camera.cpp:37-118) and PPM reader.
// ------------------------------------------------------------------------------------------
extern "C" {

crt_status crt_flycam_init(crt_flycam* cam, const float position[3])
{
    if (!cam || !position) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_flycam_init: NULL argument");
    for (int i = 0; i < 3; i++) cam->position[i] = position[i];
    cam->orientation[0] = 0.f; cam->orientation[1] = 0.f; cam->orientation[2] = -1.f; // Camera.h:41
    cam->up[0] = 0.f; cam->up[1] = 1.f; cam->up[2] = 0.f;
    cam->far_plane = 10.0f; cam->near_plane = 0.1f; cam->fov_degrees = 45.0f;         // Camera.h:44-46
    cam->speed = 0.05f; cam->sensitivity = 0.1f;                                      // Camera.h:6-7
    cam->yaw = 270.0f; cam->pitch = 0.0f;                                             // Camera.h:59-60
    return CRT_OK;
}

crt_status crt_flycam_step(crt_flycam* cam, uint32_t keys, float mouse_dx, float mouse_dy, float scroll, int* moved)
{
    if (!cam) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_flycam_step: NULL argument");
    const crt_flycam before = *cam;
    float* P = cam->position;
    const float* O = cam->orientation;
    const float* U = cam->up;
    float right[3] = {O[1] * U[2] - O[2] * U[1], O[2] * U[0] - O[0] * U[2], O[0] * U[1] - O[1] * U[0]}; // cross(orientation, up)
    const float inv = 1.0f / sqrtf(right[0] * right[0] + right[1] * right[1] + right[2] * right[2]);
    for (int i = 0; i < 3; i++) right[i] *= inv;
    const float sp = cam->speed; // the speed set by the PREVIOUS call applies to this call's moves (Camera.cpp:40-64)
    if (keys & CRT_KEY_W) for (int i = 0; i < 3; i++) P[i] += sp * O[i];
    if (keys & CRT_KEY_A) for (int i = 0; i < 3; i++) P[i] += sp * -right[i];
    if (keys & CRT_KEY_S) for (int i = 0; i < 3; i++) P[i] += sp * -O[i];
    if (keys & CRT_KEY_D) for (int i = 0; i < 3; i++) P[i] += sp * right[i];
    if (keys & CRT_KEY_SPACE) for (int i = 0; i < 3; i++) P[i] += sp * U[i];
    if (keys & CRT_KEY_LCTRL) for (int i = 0; i < 3; i++) P[i] += sp * -U[i];
    cam->speed = (keys & CRT_KEY_LSHIFT) ? 0.05f * 2.0f : 0.05f;
    if (keys & CRT_KEY_C) { P[0] = 0.0f; P[1] = 2.0f; P[2] = 12.0f; }
    if (mouse_dx != 0.f || mouse_dy != 0.f) { // right button held: yaw/pitch (Camera.cpp:89-109); dy > 0 = cursor moved UP
        cam->yaw += mouse_dx * cam->sensitivity;
        cam->pitch += mouse_dy * cam->sensitivity;
        if (cam->pitch > 89.0f) cam->pitch = 89.0f;
        if (cam->pitch < -89.0f) cam->pitch = -89.0f;
        const float k = 0.01745329251994329576923690768489f;
        float d[3] = {cosf(k * cam->yaw) * cosf(k * cam->pitch), sinf(k * cam->pitch), sinf(k * cam->yaw) * cosf(k * cam->pitch)};
        const float il = 1.0f / sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        for (int i = 0; i < 3; i++) cam->orientation[i] = d[i] * il;
    }
    if (scroll != 0.f) { // ProcessMouseScroll, Camera.cpp:28-35
        cam->fov_degrees -= scroll;
        if (cam->fov_degrees < 1.0f) cam->fov_degrees = 1.0f;
        if (cam->fov_degrees > 120.0f) cam->fov_degrees = 120.0f;
    }
    if (moved) {
        *moved = 0;
        for (int i = 0; i < 3; i++)
            if (before.position[i] != cam->position[i] || before.orientation[i] != cam->orientation[i]) *moved = 1;
        if (before.fov_degrees != cam->fov_degrees) *moved = 1;
    }
    return CRT_OK;
}

crt_status crt_flycam_inputs(const crt_flycam* cam, const float background_start[3], const float background_end[3], crt_input_struct* out)
{
    if (!cam || !out) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_flycam_inputs: NULL argument");
    return crt_camera_look(cam->position, cam->orientation, cam->up, cam->fov_degrees, cam->near_plane, cam->far_plane, background_start,
                           background_end, out);
}

crt_status crt_load_ppm(const char* path, uint8_t** out_pixels, int32_t* out_width, int32_t* out_height)
{
    if (!path || !out_pixels || !out_width || !out_height) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_load_ppm: NULL argument");
    *out_pixels = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_load_ppm: cannot open the file");
    int vals[3], got = 0, c;
    char magic[3] = {0, 0, 0};
    bool ok = fread(magic, 1, 2, f) == 2 && magic[0] == 'P' && magic[1] == '6';
    while (ok && got < 3) { // header tokens, '#' comments allowed
        c = fgetc(f);
        if (c == '#') { while ((c = fgetc(f)) != '\n' && c != EOF) { } }
        else if (c >= '0' && c <= '9') {
            long v = 0;
            do { v = v * 10 + (c - '0'); c = fgetc(f); } while (c >= '0' && c <= '9' && v < (1 << 28));
            vals[got++] = (int)v; // the character after the number is consumed (for maxval it is the single whitespace before the data)
        }
        else if (c == EOF) ok = false;
    }
    ok = ok && vals[0] > 0 && vals[1] > 0 && vals[2] == 255 && (uint64_t)vals[0] * vals[1] <= (1ull << 30);
    if (!ok) { fclose(f); return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_load_ppm: not a binary PPM (P6) with maxval 255"); }
    const size_t n = (size_t)vals[0] * vals[1] * 3;
    uint8_t* px = (uint8_t*)malloc(n);
    if (!px) { fclose(f); return crt_set_error(nullptr, CRT_ERR_NOMEM, "crt_load_ppm: out of memory"); }
    if (fread(px, 1, n, f) != n) { free(px); fclose(f); return crt_set_error(nullptr, CRT_ERR_INVALID, "crt_load_ppm: truncated pixel data"); }
    fclose(f);
    *out_pixels = px; *out_width = vals[0]; *out_height = vals[1];
    return CRT_OK;
}

void crt_free(void* p) { free(p); }

} // extern "C"
