// b2_scene_handle.hpp -- the opaque B2Scene of include/b2pt.h: host copy of the flattened scene, its device tables,
// the wide BVH, a private stream and the wavefront workspace.  Shared by b2_api.cu and b2_multi.cu.
#pragma once
#include <cstring>
#include <mutex>
#include "b2_device_scene.hpp"
#include "b2_host_scene.hpp"
#include "b2_render.hpp"
#include "b2_queries.hpp"

struct B2Scene {
    int device = 0;
    int n_sms = 148;                      // multiProcessorCount of `device`, read once at scene creation: every persistent grid is sized from it
    HostScene hs;
    DevBuf<BuildTri> d_in;
    DevBuf<float4> sph, lights, mats, groups;
    DeviceBvh bvh;
    bool built = false;
    int max_leaf = 3;
    cudaStream_t stream = nullptr;
    DevBuf<unsigned long long> isect_counters;
    RaySortScratch ray_sort;              // b2_intersect*: reordering of large ray batches (b2_queries.cu)
    RenderWorkspace *ws_ptr = nullptr;    // taken from the per-device cache at first use, returned to it by b2_scene_destroy
    RenderWorkspace &ws_ref();            // (allocating and freeing ~6 GB of wavefront state per scene handle is what made
                                          //  a create -> render -> destroy cycle 5x slower than the render itself)
    std::mutex mtx;                       // one call in flight per handle
    // L2-resident top of the tree (the lever BASELINE.json's north_star names): the wide nodes are laid out level by level, so
    // the first l2_window_bytes of the node array are the top levels (all of it when the array is small).  Every traversal
    // launch runs with a persisting access-policy window over that range (L2Window below); sized by b2_scene_build_bvh.
    size_t l2_window_bytes = 0; float l2_hit_ratio = 1.0f;
    DevScene view() const {
        DevScene v; memset(&v, 0, sizeof(v));
        v.tri = bvh.tri.p; v.tri_obj = bvh.tri_obj.p; v.nodes = bvh.nodes.p;
        v.sph = sph.p; v.lights = lights.p; v.mats = mats.p; v.groups = groups.p;
        v.n_tris = bvh.n_tris; v.n_nodes = bvh.n_nodes; v.n_sph = hs.n_sph; v.n_lights = hs.n_lights; v.n_mats = hs.n_mats; v.n_groups = hs.n_groups;
        v.scene_min = hs.scene_min; v.scene_max = hs.scene_max;
        v.cvt_magic = 0x4B000000u; v.n_sms = n_sms;
        return v;
    }
};


void b2_workspace_release(B2Scene *s);
// RAII: persisting-L2 access-policy window over the top of s's node array on stream `st` for the launches in scope.
struct L2Window {
    cudaStream_t st = nullptr; bool on = false;
    L2Window(const B2Scene *s, cudaStream_t stream) : st(stream) {
        if (!s || s->l2_window_bytes == 0 || !s->bvh.nodes.p) return;
        cudaStreamAttrValue v; memset(&v, 0, sizeof(v));
        v.accessPolicyWindow.base_ptr = (void *) s->bvh.nodes.p; v.accessPolicyWindow.num_bytes = s->l2_window_bytes;
        v.accessPolicyWindow.hitRatio = s->l2_hit_ratio; v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting; v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        on = cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v) == cudaSuccess;
        if (!on) cudaGetLastError();
    }
    ~L2Window() {
        if (!on) return;
        cudaStreamAttrValue v; memset(&v, 0, sizeof(v));       // num_bytes = 0 switches the window off for whatever the caller launches next
        cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v);
    }
    L2Window(const L2Window &) = delete; L2Window &operator=(const L2Window &) = delete;
};
extern "C" int b2_check_built(B2Scene *s);   // internal helper (not part of include/b2pt.h)
