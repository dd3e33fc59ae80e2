// b2_device_scene.hpp -- host-side owners of the device-resident scene (B2Scene behind the C ABI).
#pragma once
#include "b2_cuda_util.hpp"
#include "b2_bvh_build.cuh"
#include "b2_bvh.cuh"

struct DeviceBvh {
    DevBuf<uint4> nodes;
    DevBuf<float4> tri;
    DevBuf<int32_t> tri_obj;
    int n_tris = 0, n_nodes = 0, depth = 0;
    float build_ms = 0.f;
    int build_passes = 0;          // PLOC passes of the last build
    float sah_cost = 0.f;          // cost of the wide tree / root area, one wide-node test = 1, one triangle test = B2_SAH_C_PRIM
    void release() { nodes.release(); tri.release(); tri_obj.release(); n_tris = n_nodes = depth = 0; build_ms = 0.f; build_passes = 0; sah_cost = 0.f; }
};

int b2_build_bvh_device(const BuildTri *d_in, int n, float pad, int max_leaf, int builder, int ploc_radius, cudaStream_t st, DeviceBvh &out);
