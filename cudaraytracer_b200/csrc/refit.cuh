// refit.cuh — scene edits without a rebuild: scatter the edited records, then refit every box bottom-up on the GPU.
//
// Replaces what the reference does after ANY GUI edit: `bvh_node->Destroy()` + placement-new `BVHNode(m_List, 0, n)`
// (CudaLayer.cpp:493-494,967-968,1375-1377 -> Hittable.cuh:303-385,447-465), a full host rebuild with one managed allocation
// per split. Moving / resizing primitives or switching materials keeps the tree topology, so here only the boxes change:
//   k_apply_edits  writes the edited primitive, material record and leaf code (it carries the material type) in place;
//   k_bvh_refit    one thread per primitive recomputes its (padded) leaf box and climbs: at every inner node the SECOND
//                  thread to arrive (atomic counter) owns both finished child boxes, stores their union in the parent and goes on.
// Closest hits do not depend on the tree, only on the boxes being conservative, so parity after a refit is the same bar as
// after a build (tests/test_gpu_edit.py: primary-hit IDs bit-exact vs the oracle on the edited scene).
#pragma once

#include "device_common.cuh"

struct EditRec
{
    float4 prim, m0, m1;
    float k;
    int slot, leaf_ref, child_code;
};
static_assert(sizeof(EditRec) == 64, "edit record is 64 bytes");

__global__ void k_apply_edits(const EditRec* __restrict__ e, int n, float4* prims, float* prim_k, float4* mats, float4* nodes,
                              const int* __restrict__ leaf_ref)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    EditRec r = e[i];
    r.leaf_ref = leaf_ref[r.slot];
    prims[r.slot] = r.prim;
    prim_k[r.slot] = r.k;
    mats[2 * (size_t)r.slot] = r.m0;
    mats[2 * (size_t)r.slot + 1] = r.m1;
    reinterpret_cast<int*>(nodes + 4 * (size_t)(r.leaf_ref >> 1) + 3)[r.leaf_ref & 1] = r.child_code;
}

__device__ __forceinline__ float refit_pad_lo(float v, float abs_pad) { return v - (fabsf(v) * 4e-6f + abs_pad); }
__device__ __forceinline__ float refit_pad_hi(float v, float abs_pad) { return v + (fabsf(v) * 4e-6f + abs_pad); }

// Boxes are stored as centre / half-extent (crt_internal.h); h is rounded up so that [c - h, c + h] contains [lo, hi].
__device__ __forceinline__ void refit_store(float* box, int axis, float lo, float hi)
{
    const float c = __fadd_rn(__fmul_rn(0.5f, lo), __fmul_rn(0.5f, hi));
    box[axis] = c;
    box[3 + axis] = fmaxf(__fsub_ru(hi, c), __fsub_ru(c, lo));
}

__global__ void k_bvh_refit(float4* nodes, const float4* __restrict__ prims, const float* __restrict__ prim_k,
                            const int* __restrict__ parent, const int* __restrict__ leaf_ref, unsigned int* visits, int n_prims,
                            float abs_pad)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_prims) return;
    int ref = leaf_ref[i];
    float* nd = reinterpret_cast<float*>(nodes + 4 * (size_t)(ref >> 1));
    const int code = ~reinterpret_cast<const int*>(nd + 12)[ref & 1];
    const int type = (code >> 28) & 3;
    const float4 p = prims[i];
    float lo[3], hi[3];
    if (type == 0) { // sphere: center -/+ radius (Hittable.cuh:112-116); a negative radius still gives an ordered box
        const float r = fabsf(p.w);
        lo[0] = p.x - r; lo[1] = p.y - r; lo[2] = p.z - r;
        hi[0] = p.x + r; hi[1] = p.y + r; hi[2] = p.z + r;
    }
    else {           // rects: in-plane bounds {a0, a1, b0, b1}, plane offset k -/+ 0.0001 (Hittable.cuh:171-181)
        const float k = prim_k[i];
        const int na = type == 1 ? 2 : type == 2 ? 1 : 0;         // normal axis: XY -> z, XZ -> y, YZ -> x
        const int aa = type == 3 ? 1 : 0, ba = type == 1 ? 1 : 2; // in-plane axes
        lo[aa] = p.x; hi[aa] = p.y; lo[ba] = p.z; hi[ba] = p.w;
        lo[na] = k - 0.0001f; hi[na] = k + 0.0001f;
    }
    float* slot = nd + 6 * (ref & 1);
    for (int a = 0; a < 3; a++)
        refit_store(slot, a, refit_pad_lo(lo[a], abs_pad), refit_pad_hi(hi[a], abs_pad));
    // climb: the second thread to reach a node owns both of its child boxes
    while (true) {
        const int node = ref >> 1;
        __threadfence();
        if (atomicAdd(&visits[node], 1u) == 0u) return;
        __threadfence();
        const int up = parent[node];
        if (up < 0) return; // the root's own box is not stored anywhere
        const volatile float* c = reinterpret_cast<const volatile float*>(nodes + 4 * (size_t)node);
        float* dst = reinterpret_cast<float*>(nodes + 4 * (size_t)(up >> 1)) + 6 * (up & 1);
        for (int a = 0; a < 3; a++) // union of the two child boxes, each end rounded outwards
            refit_store(dst, a, fminf(__fsub_rd(c[a], c[3 + a]), __fsub_rd(c[6 + a], c[9 + a])),
                        fmaxf(__fadd_ru(c[a], c[3 + a]), __fadd_ru(c[6 + a], c[9 + a])));
        ref = up;
    }
}
