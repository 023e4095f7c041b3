// crt_internal.h — shared between the host-side translation units of libcrt.so (not installed).
#pragma once

#include "crt.h"

#include <cstdint>
#include <string>
#include <vector>

// Records `msg` as the last error of `ctx` (and process-wide when ctx is NULL) and returns `st`.
crt_status crt_set_error(crt_context* ctx, crt_status st, const char* msg);

// ---------------------------------------------------------------------------------------------
// Packed device scene (what the kernels read). Built on the host by scene_pack.cpp.
//
// BVH2, one primitive per leaf. Node = 64 bytes = 4 x 128-bit words:
//   child boxes as CENTRE c and HALF-EXTENT h (rounded outwards): with t_mid = c * (1/d) - o * (1/d) per axis, the slab
//   interval of an axis is t_mid -/+ h * |1/d| whatever the sign of d, so the test needs no per-plane min/max
//   w0 = {c0.c.x, c0.c.y, c0.c.z, c0.h.x}
//   w1 = {c0.h.y, c0.h.z, c1.c.x, c1.c.y}
//   w2 = {c1.c.z, c1.h.x, c1.h.y, c1.h.z}
//   w3 = {child0, child1, 0, 0} (int bits)
// child >= 0: index of an inner node. child < 0: leaf, ~child = (type << 28) | (material type << 26) | prim index.
// Nodes are stored breadth-first so the first K nodes are the hottest (they are staged in shared memory).
//
// Primitive i (leaf order): prim[i] = sphere {cx, cy, cz, r}
//                                    | rect {a0, a1, b0, b1} (bounds on the two in-plane axes) with prim_k[i] = plane offset
// Material record i (leaf order, one per primitive, 32 bytes = 2 words):
//   m0 = {r, g, b, param}      constant colour | checker odd colour; param = fuzz | ir | (float)light_intensity
//   m1 = {er, eg, eb, bits}    checker even colour; bits = mat_type | tex_type << 4 | (image + 1) << 8
// prim_id[i] = index of the hittable in the list the caller uploaded.
// ---------------------------------------------------------------------------------------------
struct CrtPackedNode
{
    float w[12];
    int32_t child[2];
    int32_t pad[2];
};
static_assert(sizeof(CrtPackedNode) == 64, "node is 64 bytes");

#define CRT_LEAF_TYPE_SHIFT 28
#define CRT_LEAF_MAT_SHIFT 26 /* 2 bits: the material type of the primitive, so that a hit can be binned without a load */
#define CRT_LEAF_INDEX_MASK 0x03FFFFFF

struct CrtPackedScene
{
    std::vector<CrtPackedNode> nodes;
    std::vector<float> prims;    // 4 floats per primitive
    std::vector<float> prim_k;   // 1 float per primitive (rect plane offset; 0 for spheres)
    std::vector<float> mats;     // 8 floats per primitive
    std::vector<int32_t> prim_id;
    // for the GPU refit (scene edits that keep the topology): where every node / leaf slot hangs, as parent_node * 2 + child
    std::vector<int32_t> parent;   // per node; -1 for the root
    std::vector<int32_t> leaf_ref; // per primitive slot
    float abs_pad = 0.f;           // absolute box padding used by this build (scene_pack.cpp)
    crt_bvh_stats stats;
};

// Peels the chain of huge leaves off the top of a packed tree (DevScene::entry / top_code, device_common.cuh): starting at
// the root, while a node has exactly one leaf child whose box is at least a quarter of the node's own, that leaf is recorded
// (its code, and the slot in it) and the walk goes on in the other child. `nodes` are the first `n_nodes` packed nodes of the
// tree (the chain is at the top, so a short prefix is enough). Returns the number of peeled leaves (<= max_top).
int crt_peel_top(const CrtPackedNode* nodes, size_t n_nodes, uint64_t n_prims, int max_top, int32_t* entry, int32_t* codes);

// Builds the BVH over the ACTIVE hittables (Hittable.cuh:311-320 filters inactive ones) and packs everything.
// Returns false and fills err on invalid input.
bool crt_pack_scene(const crt_scene_desc* scene, CrtPackedScene* out, std::string* err);
// One primitive / one material record in the packed layout (also used by the edit path).
bool crt_prepare_gpu_build(const crt_scene_desc* scene, std::vector<uint32_t>* active, float cmin[3], float cmax[3], float blo[3],
                           float bhi[3], float* abs_pad, std::string* err);
bool crt_check_material(const crt_scene_desc* scene, int64_t index, std::string* err);
void crt_pack_prim(const crt_hittable& h, float prim[4], float* k);
void crt_pack_material(const crt_scene_desc* scene, const crt_material& m, float record[8]);

// Host-side layout mirror of the reference scene graph (ref_abi.h) -> flat description.
struct CrtFlatScene
{
    std::vector<crt_hittable> hittables;
    std::vector<crt_material> materials;
    std::vector<crt_texture> textures;
    std::vector<crt_image> images;
    crt_scene_desc desc() const
    {
        crt_scene_desc d;
        d.hittables = hittables.data(); d.n_hittables = hittables.size();
        d.materials = materials.data(); d.n_materials = materials.size();
        d.textures = textures.data();   d.n_textures = textures.size();
        d.images = images.data();       d.n_images = images.size();
        return d;
    }
};
bool crt_flatten_reference_world(const void* world, CrtFlatScene* out, std::string* err);
