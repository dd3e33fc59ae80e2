// b2_scene_handle.hpp -- the opaque B2Scene of include/b2pt.h: host copy of the flattened scene, its device tables,
// the wide BVH, a private stream and the wavefront workspace.  Shared by b2_api.cu and b2_multi.cu.
#pragma once
#include <cstring>
#include <mutex>
#include "b2_device_scene.hpp"
#include "b2_host_scene.hpp"
#include "b2_render.hpp"

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
    RenderWorkspace *ws_ptr = nullptr;    // taken from the per-device cache at first use, returned to it by b2_scene_destroy
    RenderWorkspace &ws_ref();            // (allocating and freeing ~6 GB of wavefront state per scene handle is what made
                                          //  a create -> render -> destroy cycle 5x slower than the render itself)
    std::mutex mtx;                       // one call in flight per handle
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
extern "C" int b2_check_built(B2Scene *s);   // internal helper (not part of include/b2pt.h)
