// ref_flatten.cpp — walks the reference's pointer graph (host-readable: cudaMallocManaged or plain host
// memory) starting at the BVH root that LaunchKernel receives, and produces the flat crt_scene_desc.
// The reference's tree only tells us WHICH hittables are in the scene; our own BVH is rebuilt from them.

#include "crt_internal.h"
#include "ref_abi.h"

#include <cstring>
#include <unordered_set>

namespace {

using namespace ref_abi;

struct Walker
{
    CrtFlatScene* out;
    std::string* err;
    std::unordered_set<const Hittable*> seen;
    int depth_guard = 0;

    int add_texture(const Texture* t)
    {
        crt_texture ct;
        memset(&ct, 0, sizeof(ct));
        ct.image = -1;
        if (!t || !t->Object) {
            *err = "reference scene: material without texture";
            return -1;
        }
        void* obj = *t->Object;
        switch (t->type) {
        case CRT_CONSTANT: {
            const Constant* c = (const Constant*)obj;
            ct.type = CRT_CONSTANT;
            for (int i = 0; i < 3; i++) ct.color[i] = c->color.e[i];
            break;
        }
        case CRT_CHECKER: {
            const Checker* c = (const Checker*)obj;
            ct.type = CRT_CHECKER;
            for (int i = 0; i < 3; i++) { ct.odd[i] = c->odd->color.e[i]; ct.even[i] = c->even->color.e[i]; }
            break;
        }
        case CRT_IMAGE: {
            const Image* im = (const Image*)obj;
            ct.type = CRT_IMAGE;
            crt_image ci;
            ci.pixels = im->data;
            ci.width = im->width;
            ci.height = im->height;
            // share one crt_image per distinct texel buffer
            int found = -1;
            for (size_t k = 0; k < out->images.size(); k++)
                if (out->images[k].pixels == ci.pixels && out->images[k].width == ci.width && out->images[k].height == ci.height) found = (int)k;
            if (found < 0) { out->images.push_back(ci); found = (int)out->images.size() - 1; }
            ct.image = found;
            break;
        }
        default:
            *err = "reference scene: unknown texture type";
            return -1;
        }
        out->textures.push_back(ct);
        return (int)out->textures.size() - 1;
    }

    int add_material(const Material* m)
    {
        if (!m || !m->Object) {
            *err = "reference scene: hittable without material";
            return -1;
        }
        void* obj = *m->Object;
        crt_material cm;
        cm.type = m->type; cm.texture = -1; cm.param = 0.f; cm.light_intensity = 0;
        switch (m->type) {
        case CRT_LAMBERTIAN: cm.texture = add_texture(((const Lambertian*)obj)->albedo); break;
        case CRT_METAL:
            cm.texture = add_texture(((const Metal*)obj)->albedo);
            cm.param = ((const Metal*)obj)->fuzz;
            break;
        case CRT_DIELECTRIC: cm.param = ((const Dielectric*)obj)->ir; break;
        case CRT_DIFFUSELIGHT:
            cm.texture = add_texture(((const DiffuseLight*)obj)->albedo);
            cm.light_intensity = ((const DiffuseLight*)obj)->light_intensity;
            break;
        default:
            *err = "reference scene: unknown material type";
            return -1;
        }
        if (m->type != CRT_DIELECTRIC && cm.texture < 0) return -1;
        out->materials.push_back(cm);
        return (int)out->materials.size() - 1;
    }

    bool add_leaf(const Hittable* h)
    {
        if (!seen.insert(h).second) return true; // single-object leaves have left == right (Hittable.cuh:326-328)
        if (!h->Object) {
            *err = "reference scene: hittable without object";
            return false;
        }
        void* obj = *h->Object;
        crt_hittable ch;
        memset(&ch, 0, sizeof(ch));
        ch.type = h->type;
        ch.active = h->isActive ? 1 : 0;
        const Material* mat;
        if (h->type == HT_SPHERE) {
            const Sphere* s = (const Sphere*)obj;
            ch.cx = s->center.e[0]; ch.cy = s->center.e[1]; ch.cz = s->center.e[2];
            ch.a = s->radius;
            mat = s->mat_ptr;
        }
        else {
            const Rect* r = (const Rect*)obj;
            ch.cx = r->center.e[0]; ch.cy = r->center.e[1]; ch.cz = r->center.e[2];
            ch.a = r->width; ch.b = r->height;
            mat = r->mat_ptr;
        }
        ch.material = add_material(mat);
        if (ch.material < 0) return false;
        out->hittables.push_back(ch);
        return true;
    }

    bool walk(const Hittable* h)
    {
        if (!h) return true;
        if (h->type == HT_BVHNODE) {
            if (++depth_guard > 4096) {
                *err = "reference scene: BVH too deep or cyclic";
                return false;
            }
            const BVHNode* n = h->Object ? (const BVHNode*)*h->Object : nullptr;
            bool ok = !n || (walk(n->left) && (n->right == n->left || walk(n->right)));
            depth_guard--;
            return ok;
        }
        if (h->type >= HT_SPHERE && h->type <= HT_YZRECT) return add_leaf(h);
        *err = "reference scene: unsupported hittable type in BVH (only Sphere / XYRect / XZRect / YZRect leaves)";
        return false;
    }
};

} // namespace

bool crt_flatten_reference_world(const void* world, CrtFlatScene* out, std::string* err)
{
    const ref_abi::Hittable* w = (const ref_abi::Hittable*)world;
    Walker W;
    W.out = out;
    W.err = err;
    if (w->type == ref_abi::HT_HITTABLELIST) {
        // HittableList::Hit (Hittable.cuh:545-560) takes the closest hit over its elements and treats every one of them as a
        // BVHNode hittable: the scene is the union of the hittables under those trees
        const ref_abi::HittableList* l = w->Object ? (const ref_abi::HittableList*)*w->Object : nullptr;
        if (!l || l->list_size < 0 || (l->list_size > 0 && !l->list)) {
            *err = "reference scene: HittableList without a list";
            return false;
        }
        for (int i = 0; i < l->list_size; i++) {
            const ref_abi::Hittable* e = l->list[i];
            if (!e || e->type != ref_abi::HT_BVHNODE) {
                *err = "reference scene: the elements of a HittableList must be BVHNode hittables (Hittable.cuh:552)";
                return false;
            }
            if (!W.walk(e)) return false;
        }
        return true;
    }
    if (w->type != ref_abi::HT_BVHNODE) {
        *err = "reference scene: world must be a Hittable of type BVHNODE (Kernel.cu:40) or HITTABLELIST (Hittable.cuh:532-581)";
        return false;
    }
    return W.walk(w);
}
