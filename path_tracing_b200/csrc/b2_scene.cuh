// b2_scene.cuh -- device-resident scene layout (what the kernels read from HBM / L2).
//
//   triangles : BVH-leaf order, 48 B = 3 x float4  {v0 | order}, {e1 = v1-v0 | material + transparent flag},
//               {e2 = v2-v0 | group index}; one 128-bit load each, coalesced within a leaf.
//   nodes     : 8-wide BVH, 80 B = 5 x uint4 with 8-bit quantised child boxes (layout below).
//   spheres, light balls, materials, group boxes: short tables (dozens of entries), float4-packed.
// "order" is the reference's object visiting order (std::map group key, then insertion order,
// cpu_bdpt.cpp:43-46): on exact t ties the LATER object wins (object.cpp:92 accepts t == tMax).
#pragma once
#include "b2_common.cuh"
#include "b2_bsdf.cuh"

#define B2_PRIM_MISS (-1)
#define B2_KIND_TRI 0
#define B2_KIND_SPHERE 1
#define B2_KIND_LIGHT 2
B2_HD int prim_make(int kind, int index) { return (kind << 28) | index; }   // up to 2^28 primitives per kind
B2_HD int prim_kind(int prim) { return prim >> 28; }
B2_HD int prim_index(int prim) { return prim & 0x0fffffff; }

#define B2_MAT_TRANSPARENT 0x80000000u   // material eta > 0: never blocks shadow rays (cpu_bdpt.cpp:102)

// 8-wide node, 80 bytes.  w0 = {origin.x, origin.y, origin.z, ex | ey<<8 | ez<<16 | imask<<24}
//   w1 = {child_base, tri_base, vmask, 0}
//   w2 = {qlo_x[0..3], qlo_x[4..7], qlo_y[0..3], qlo_y[4..7]}
//   w3 = {qlo_z[0..3], qlo_z[4..7], qhi_x[0..3], qhi_x[4..7]}
//   w4 = {qhi_y[0..3], qhi_y[4..7], qhi_z[0..3], qhi_z[4..7]}
// child box of slot s = origin + q[s] * 2^(e-127) per axis (e = biased IEEE exponent byte).  An empty slot has the
// inverted box qlo = 255, qhi = 0 (never hit).  imask bit s: slot s is an internal child, its node index is
// child_base + popc(imask & ((1<<s)-1)).  vmask bits 3s .. 3s+2: leaf slot s holds 1..3 triangles (the low cnt bits
// of its field are set); the node's triangles sit compactly from tri_base in slot order, so the triangle of vmask
// bit b is tri_base + popc(vmask & ((1<<b)-1)).  The traversal gets one hit bit per slot from eight branch-free slab
// tests and derives both child lists from it with a handful of logic operations (b2_bvh.cuh: node_test).
// Slot s sits towards the (s&1 ? +x : -x, s&2 ? +y : -y, s&4 ? +z : -z) corner of the node, so a ray with
// octant bits `oct` (bit set where dir < 0) meets slots in roughly increasing (s ^ oct).
struct WideNode { uint4 w[5]; };
#define B2_NODE_UINT4 5

struct DevScene {
    const float4 *tri;      // 3 per triangle
    const int32_t *tri_obj; // obj_id per triangle slot
    const uint4 *nodes;     // 5 per node
    const float4 *sph;      // 2 per sphere: {c, r}, {order, material|flag, group, obj_id} (int bits)
    const float4 *lights;   // 4 per light: {pos, ball_r}, {dir, cutoff}, {illum, is_parallel}, {normalize(dir), cos(cutoff)}
    const float4 *mats;     // 2 per material: {base, roughness}, {metallic, eta, 0, 0}
    const float4 *groups;   // 2 per group: {min, 0}, {max, 0}  (degenerate extents pre-widened, object.cpp:108-111)
    int n_tris, n_nodes, n_sph, n_lights, n_mats, n_groups;
    float3 scene_min, scene_max;   // union of the group boxes (cpu_bdpt.cpp:181-187)
    int ggx_fix;                   // != 0: materials are marked for the normalised GGX D (mat_mark_fix); 0 in parity mode
    int n_sms;                     // SMs of the device the scene lives on (host side: sizes the persistent grids)
    uint32_t cvt_magic;            // 0x4B000000 = 2^23 as float bits, for the PRMT byte -> float conversion of b2_bvh.cuh (kept a runtime value on purpose)
    int gpu_sem;                   // != 0: intersection semantics of the reference's GPU integrators (pt_cu / ppm_cu through
                                   // geometric.cuh:240-390) instead of cpu_bdpt.cpp's: two-root spheres, opaque shadow rays
};

B2_HD Mat load_mat(const DevScene &sc, int m) {
    float4 a = sc.mats[2 * m], b = sc.mats[2 * m + 1];
    Mat r; r.base = mk3(a); r.roughness = a.w; r.metallic = b.x; r.eta = b.y;
    if (sc.ggx_fix) mat_mark_fix(r);
    return r;
}
