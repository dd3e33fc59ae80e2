// b2_bvh.cuh -- traversal of the 8-wide quantised BVH (layout in b2_scene.cuh).
//
// One ray per thread.  The traversal stack holds "node groups": {child_base, hits<<8 | imask}, i.e. the
// still-unvisited hit children of one wide node, 8 bytes per level.  Children are popped in increasing
// (slot ^ octant), which is roughly front-to-back thanks to the builder's slot assignment.
// Leaf triangles run through exactly the same tri_test() as the exhaustive kernels, so the result
// (min t, ties -> highest reference order) is independent of visiting order and bit-identical to them
// as long as no box containing an accepted hit is culled: boxes are padded at build time and the slab
// comparison carries a (1 + 4 ulp) slack.
#pragma once
#include "b2_isect.cuh"

#define B2_STACK_CAP 64

struct RayPre { float3 o, d, idir; uint32_t oct; };
// 1 / d for the slab tests of the wide nodes ONLY (the boxes are padded and the comparison carries a 4-ulp slack; hit distances
// never come from it): the approximate reciprocal on the device in every translation unit - b2_render.cu and b2_integrators.cu
// got it from -prec-div=false already, b2_queries.cu paid three IEEE divisions per ray
B2_HD float safe_rcp(float d) { return fabsf(d) > 1e-30f ? fast_rcp(d) : copysignf(1e30f, d); }
B2_HD RayPre ray_pre(float3 o, float3 d, float3 idir) {          // idir = safe_rcp of d, computed elsewhere (k_connect's prepare phase)
    RayPre r; r.o = o; r.d = d; r.idir = idir;
    r.oct = (d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u);
    return r;
}
B2_HD RayPre ray_pre(float3 o, float3 d) { return ray_pre(o, d, mk3(safe_rcp(d.x), safe_rcp(d.y), safe_rcp(d.z))); }
// AABB::intersectAABB (object.cpp:104-121) with the ray's precomputed approximate reciprocals instead of three IEEE
// divisions: only for the render's shadow rays, next to tri_test_fast (statistical parity; closest hits, b2_intersect
// and every bit-exactness test keep group_pass).  Same accept / reject except for rays that graze a group box within rounding.
B2_HD bool group_pass_fast(const DevScene &sc, int g, const RayPre &r, float tmin, float tmax) {
    const float4 mn = sc.groups[2 * g], mx = sc.groups[2 * g + 1];
    const float x0 = (mn.x - r.o.x) * r.idir.x, x1 = (mx.x - r.o.x) * r.idir.x;
    const float y0 = (mn.y - r.o.y) * r.idir.y, y1 = (mx.y - r.o.y) * r.idir.y;
    const float z0 = (mn.z - r.o.z) * r.idir.z, z1 = (mx.z - r.o.z) * r.idir.z;
    tmin = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin));
    tmax = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), tmax));
    return !(tmax <= tmin);
}
// occluded_spheres (b2_isect.cuh) with the fast sphere and group-box tests; cpu_bdpt.cpp semantics only (not sc.gpu_sem)
B2_HD bool occluded_spheres_fast(const DevScene &sc, const RayPre &r, float tMin, float tMax) {
    for (int i = 0; i < sc.n_sph; ++i) {
        const float4 a = sc.sph[2 * i], b = sc.sph[2 * i + 1];
        if (f2u(b.y) & B2_MAT_TRANSPARENT) continue;
        if (sphere_test_fast(r.o, r.d, mk3(a), a.w, tMin, tMax) && group_pass_fast(sc, (int) f2u(b.z), r, tMin, tMax)) return true;
    }
    return false;
}
B2_HD uint32_t byte_of(uint32_t lo_word, uint32_t hi_word, int i) { return ((i < 4 ? lo_word : hi_word) >> (8 * (i & 3))) & 0xffu; }
// (float) byte i of {lo_word, hi_word}.  Two exact forms on the device: I2F.U8 with a byte selector (one instruction on the
// quarter-rate conversion pipe, 8 issue cycles per warp) or PRMT + FADD (one PRMT builds the float 2^23 + byte, one FADD
// removes the 2^23: ALU + FMA pipes, 2 cycles each).  A node test converts 48 bytes; ncu on the C3 connection stage
// (profiles/r02_node_test.md) had the conversion pipe 40 % busy with I2F alone, so the six plane arrays are split between
// the two forms: bit p of B2_CVT_PRMT_PLANES set = plane array p (x near, x far, y near, y far, z near, z far) uses PRMT.
#ifndef B2_CVT_PRMT_PLANES
#define B2_CVT_PRMT_PLANES 0x2A
#endif
// `magic` is 0x4B000000 read from the kernel's parameters (DevScene::cvt_magic), not a literal: PRMT takes ONE immediate,
// and with a literal there ptxas moves every byte selector into a register first (18 MOVs per node test in the SASS).
template <int PLANE>
B2_HD float byte_f32(uint32_t lo_word, uint32_t hi_word, int i, uint32_t magic) {
#if defined(__CUDA_ARCH__)
    if ((B2_CVT_PRMT_PLANES >> PLANE) & 1)
        return __uint_as_float(__byte_perm(i < 4 ? lo_word : hi_word, magic, 0x7650u | (uint32_t) (i & 3))) - 8388608.0f;
#endif
    return (float) byte_of(lo_word, hi_word, i);
}

// triangle slot of bit `bit` of a node's vmask-relative leaf mask (b2_scene.cuh)
B2_HD int tri_of_bit(uint32_t tri_base, uint32_t vmask, int bit) { return (int) tri_base + popc32(vmask & ((1u << bit) - 1u)); }

// One 16-byte row of a wide node.  B2_NODE_L1_HINT = 1: loaded with the L1 eviction priority "evict_last" (LDG.E.EL), meant to
// let the node lines outlive the triangle rows, vertex records and local-memory lines that share the 60 KB of L1 with them.
// Measured: C3 1003.4 -> 1006.0 ms, C5 unchanged - off (profiles/r02_connect_summary.md).
#ifndef B2_NODE_L1_HINT
#define B2_NODE_L1_HINT 0
#endif
B2_HD uint4 ld_node(const uint4 *p) {
#if defined(__CUDA_ARCH__) && B2_NODE_L1_HINT
    uint4 v;
    asm volatile("ld.global.L1::evict_last.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
#else
    return *p;
#endif
}
// Tests the 8 children of `node`: one hit bit per slot from eight branch-free slab tests (an empty slot carries an
// inverted box).  Returns the hit internal children as bits (slot ^ oct) in hits_p and the hit leaf triangles as a
// sparse 24-bit mask (bit 3*slot + j; triangle slot = tri_of_bit(tri_base, vmask, bit)).
B2_HD void node_test(const DevScene &sc, uint32_t node, const RayPre &r, float tmin, float tmax,
                     uint32_t &hits_p, uint32_t &tribits, uint32_t &child_base, uint32_t &tri_base, uint32_t &imask, uint32_t &vmask) {
    const uint4 *__restrict__ np = sc.nodes + (size_t) node * B2_NODE_UINT4;
    const uint4 w0 = ld_node(np), w1 = ld_node(np + 1), w2 = ld_node(np + 2), w3 = ld_node(np + 3), w4 = ld_node(np + 4);
    const float sx = u2f((w0.w & 0xffu) << 23), sy = u2f(((w0.w >> 8) & 0xffu) << 23), sz = u2f(((w0.w >> 16) & 0xffu) << 23);
    imask = w0.w >> 24;
    child_base = w1.x; tri_base = w1.y; vmask = w1.z;
    const float ax = sx * r.idir.x, ay = sy * r.idir.y, az = sz * r.idir.z;
    const float bx = (u2f(w0.x) - r.o.x) * r.idir.x, by = (u2f(w0.y) - r.o.y) * r.idir.y, bz = (u2f(w0.z) - r.o.z) * r.idir.z;
    // near/far byte planes per axis, chosen by the ray's direction sign
    const bool nx = r.oct & 1u, ny = r.oct & 2u, nz = r.oct & 4u;
    const uint32_t xl0 = nx ? w3.z : w2.x, xl1 = nx ? w3.w : w2.y, xh0 = nx ? w2.x : w3.z, xh1 = nx ? w2.y : w3.w;
    const uint32_t yl0 = ny ? w4.x : w2.z, yl1 = ny ? w4.y : w2.w, yh0 = ny ? w2.z : w4.x, yh1 = ny ? w2.w : w4.y;
    const uint32_t zl0 = nz ? w4.z : w3.x, zl1 = nz ? w4.w : w3.y, zh0 = nz ? w3.x : w4.z, zh1 = nz ? w3.y : w4.w;
    uint32_t hit = 0u;
    const uint32_t mg = sc.cvt_magic;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float tlx = byte_f32<0>(xl0, xl1, i, mg) * ax + bx, thx = byte_f32<1>(xh0, xh1, i, mg) * ax + bx;
        const float tly = byte_f32<2>(yl0, yl1, i, mg) * ay + by, thy = byte_f32<3>(yh0, yh1, i, mg) * ay + by;
        const float tlz = byte_f32<4>(zl0, zl1, i, mg) * az + bz, thz = byte_f32<5>(zh0, zh1, i, mg) * az + bz;
        const float tn = fmaxf(fmaxf(tlx, tly), fmaxf(tlz, tmin));
        const float tf = fminf(fminf(thx, thy), fminf(thz, tmax));
        hit |= (tn <= tf * 1.0000005f + 1e-30f) ? (1u << i) : 0u;
    }
    // internal children: slot space -> visiting order p = slot ^ oct, three conditional swaps whose shift IS the octant bit
    uint32_t h = hit & imask;
    // (a zero shift leaves the mask as it is: (h & hi) | (h & lo) = h)
    { const uint32_t s = r.oct & 4u; h = ((h << s) & 0xf0u) | ((h >> s) & 0x0fu); }
    { const uint32_t s = r.oct & 2u; h = ((h << s) & 0xccu) | ((h >> s) & 0x33u); }
    { const uint32_t s = r.oct & 1u; h = ((h << s) & 0xaau) | ((h >> s) & 0x55u); }
    hits_p = h;
    // leaf children: every hit bit tripled (bit s -> bits 3s .. 3s+2), masked with the triangles that exist
    uint32_t x = hit & ~imask;
    x = (x | (x << 8)) & 0x00f00fu;
    x = (x | (x << 4)) & 0x0c30c3u;
    x = (x | (x << 2)) & 0x249249u;
    tribits = (x * 7u) & vmask;
}

// The still-unvisited part of the tree: the current group of hit children plus the shared-memory stack of older groups.
struct TravPos { int sp; uint2 grp; uint32_t node; bool have; };
#ifdef B2_TRACK_STACK_DEPTH
static int b2_max_stack_depth = 0;
#endif
B2_HD void trav_reset(TravPos &p, bool has_root) { p.sp = 0; p.grp = make_uint2(0u, 0u); p.node = 0u; p.have = has_root; }
// Makes p.node the next wide node to test (the root at first, then the nearest unvisited hit child); false = exhausted.
B2_HD bool trav_next(TravPos &p, uint32_t oct, const uint2 *stack, int sstride) {
    if (!p.have) {
        if ((p.grp.y >> 8) == 0u) { if (p.sp == 0) return false; p.grp = stack[(--p.sp) * sstride]; }
        const int k = ffs32(p.grp.y >> 8) - 1;
        p.grp.y &= ~(1u << (8 + k));
        const uint32_t slot = (uint32_t) k ^ oct;
        p.node = p.grp.x + popc32(p.grp.y & 0xffu & ((1u << slot) - 1u));
    }
    p.have = false;
    return true;
}
// The hit internal children of the node just tested become the current group; the rest of the old one is stacked.
B2_HD void trav_push(TravPos &p, uint32_t hits_p, uint32_t child_base, uint32_t imask, uint2 *stack, int sstride) {
    if (hits_p) {
        if ((p.grp.y >> 8) != 0u && p.sp < B2_STACK_CAP) stack[(p.sp++) * sstride] = p.grp;
#ifdef B2_TRACK_STACK_DEPTH         // host compilation of the tests only: the deepest stack any walk reached (tests/hostsim)
        if (p.sp > b2_max_stack_depth) b2_max_stack_depth = p.sp;
#endif
        p.grp = make_uint2(child_base, (hits_p << 8) | imask);
    }
}

template <bool COUNT>
B2_HD HitRec closest_hit_bvh(const DevScene &sc, float3 ro, float3 rd, uint2 *stack, int sstride, TravCounters *cnt) {
    HitRec best; best.prim = B2_PRIM_MISS; best.t = B2_T_INF;
    uint32_t best_order = 0;
    if (sc.n_nodes > 0) {
        const RayPre r = ray_pre(ro, rd);
        TravPos p; trav_reset(p, true);
        while (trav_next(p, r.oct, stack, sstride)) {
            uint32_t hits_p, tribits, child_base, tri_base, imask, vmask;
            node_test(sc, p.node, r, 1e-4f, best.t, hits_p, tribits, child_base, tri_base, imask, vmask);
            if (COUNT) cnt->nodes++;
            while (tribits) {
                const int bit = ffs32(tribits) - 1;
                tribits &= tribits - 1u;
                const int ti = tri_of_bit(tri_base, vmask, bit);
                const float4 a = sc.tri[3 * ti], b = sc.tri[3 * ti + 1], c = sc.tri[3 * ti + 2];
                if (COUNT) cnt->tris++;
                float t;
                if (tri_test(ro, rd, mk3(a), mk3(b), mk3(c), 1e-4f, best.t, t)) {
                    const uint32_t order = f2u(a.w);
                    if (accept_hit(sc, ro, rd, t, order, (int) f2u(c.w), best, best_order)) { best.t = t; best.prim = prim_make(B2_KIND_TRI, ti); best_order = order; }
                }
            }
            trav_push(p, hits_p, child_base, imask, stack, sstride);
        }
    }
    closest_spheres(sc, ro, rd, best, best_order);
    closest_lights(sc, ro, rd, best);
    return best;
}

// Closest-hit traversal as a resumable state machine, for kernels that hand a finished lane its next ray while the other
// lanes of the warp keep walking (k_intersect): same node / triangle arithmetic and acceptance rule as closest_hit_bvh.
struct ClosestTrav { RayPre r; HitRec best; uint32_t best_order; TravPos p; };
B2_HD void closest_init(ClosestTrav &t, const DevScene &sc, float3 ro, float3 rd) {
    t.r = ray_pre(ro, rd); t.best.prim = B2_PRIM_MISS; t.best.t = B2_T_INF; t.best_order = 0u;
    trav_reset(t.p, sc.n_nodes > 0);
}
// The node half of a step: pops the next wide node, tests its 8 children, stacks the hit internal ones and returns the hit
// leaf triangles as (tri_base, vmask, sparse 24-bit mask).  Returns false when the tree is exhausted (spheres / light
// balls: closest_finish).
template <bool COUNT>
B2_HD bool closest_node_step(const DevScene &sc, ClosestTrav &t, uint2 *stack, int sstride, TravCounters *cnt, uint32_t &tribits, uint32_t &tri_base, uint32_t &vmask) {
    tribits = 0u; tri_base = 0u; vmask = 0u;
    if (!trav_next(t.p, t.r.oct, stack, sstride)) return false;
    uint32_t hits_p, child_base, imask;
    node_test(sc, t.p.node, t.r, 1e-4f, t.best.t, hits_p, tribits, child_base, tri_base, imask, vmask);
    if (COUNT) cnt->nodes++;
    trav_push(t.p, hits_p, child_base, imask, stack, sstride);
    return true;
}
// The hit leaf triangles of one node test, by the ray's own lane, with the full acceptance rule (accept_hit).
template <bool COUNT>
B2_HD void closest_test_tris(const DevScene &sc, ClosestTrav &t, uint32_t tribits, uint32_t tri_base, uint32_t vmask, TravCounters *cnt) {
    while (tribits) {
        const int bit = ffs32(tribits) - 1;
        tribits &= tribits - 1u;
        const int ti = tri_of_bit(tri_base, vmask, bit);
        const float4 a = sc.tri[3 * ti], b = sc.tri[3 * ti + 1], c = sc.tri[3 * ti + 2];
        if (COUNT) cnt->tris++;
        float tt;
        if (tri_test(t.r.o, t.r.d, mk3(a), mk3(b), mk3(c), 1e-4f, t.best.t, tt)) {
            const uint32_t order = f2u(a.w);
            if (accept_hit(sc, t.r.o, t.r.d, tt, order, (int) f2u(c.w), t.best, t.best_order)) { t.best.t = tt; t.best.prim = prim_make(B2_KIND_TRI, ti); t.best_order = order; }
        }
    }
}
// One wide node and its leaf triangles, all by the ray's own lane.
template <bool COUNT>
B2_HD bool closest_step(const DevScene &sc, ClosestTrav &t, uint2 *stack, int sstride, TravCounters *cnt) {
    uint32_t tribits, tri_base, vmask;
    if (!closest_node_step<COUNT>(sc, t, stack, sstride, cnt, tribits, tri_base, vmask)) return false;
    closest_test_tris<COUNT>(sc, t, tribits, tri_base, vmask, cnt);
    return true;
}
B2_HD HitRec closest_finish(const DevScene &sc, ClosestTrav &t) {
    closest_spheres(sc, t.r.o, t.r.d, t.best, t.best_order);
    closest_lights(sc, t.r.o, t.r.d, t.best);
    return t.best;
}

// Any-hit traversal as a resumable state machine: the connection stage keeps one of these per lane and hands a finished
// lane its next shadow ray while the other lanes of the warp are still walking the tree (k_connect, trace phase).
struct AnyTrav { RayPre r; float tmax; TravPos p; };
enum { B2_TRAV_CONTINUE = 0, B2_TRAV_HIT = 1, B2_TRAV_EXHAUSTED = 2 };
B2_HD void any_init(AnyTrav &t, const DevScene &sc, const RayPre &r, float tmax) { t.r = r; t.tmax = tmax; trav_reset(t.p, sc.n_nodes > 0); }
B2_HD void any_init(AnyTrav &t, const DevScene &sc, float3 ro, float3 rd, float tmax) { any_init(t, sc, ray_pre(ro, rd), tmax); }
// One wide node (+ its leaf triangles).  B2_TRAV_EXHAUSTED: no triangle blocks the segment (spheres are the caller's).
template <bool COUNT, bool FAST_TRI = false, bool OPAQUE = false>     // OPAQUE: every triangle blocks, no group cull (sc.gpu_sem callers)
B2_HD int any_step(const DevScene &sc, AnyTrav &t, uint2 *stack, int sstride, TravCounters *cnt) {
    const float tMin = 1e-3f;
    if (!trav_next(t.p, t.r.oct, stack, sstride)) return B2_TRAV_EXHAUSTED;
    uint32_t hits_p, tribits, child_base, tri_base, imask, vmask;
    node_test(sc, t.p.node, t.r, tMin, t.tmax, hits_p, tribits, child_base, tri_base, imask, vmask);
    if (COUNT) cnt->nodes++;
    while (tribits) {
        const int bit = ffs32(tribits) - 1;
        tribits &= tribits - 1u;
        const int ti = tri_of_bit(tri_base, vmask, bit);
        const float4 a = sc.tri[3 * ti], b = sc.tri[3 * ti + 1], c = sc.tri[3 * ti + 2];   // one round trip: all rows before the flag test
        if (!OPAQUE && (f2u(b.w) & B2_MAT_TRANSPARENT)) continue;
        if (COUNT) cnt->tris++;
        float tt;
        const bool hit = FAST_TRI ? tri_test_fast(t.r.o, t.r.d, mk3(a), mk3(b), mk3(c), tMin, t.tmax) : tri_test(t.r.o, t.r.d, mk3(a), mk3(b), mk3(c), tMin, t.tmax, tt);
        if (hit && (OPAQUE || (FAST_TRI ? group_pass_fast(sc, (int) f2u(c.w), t.r, tMin, t.tmax) : group_pass(sc, (int) f2u(c.w), t.r.o, t.r.d, tMin, t.tmax)))) return B2_TRAV_HIT;
    }
    trav_push(t.p, hits_p, child_base, imask, stack, sstride);
    return B2_TRAV_CONTINUE;
}

// The node half of any_step: the hit leaf triangles come back as (tri_base, vmask, sparse 24-bit mask) for a caller that
// tests the triangles of a whole warp together (k_connect's trace phase).  false = the tree is exhausted.
template <bool COUNT>
B2_HD bool any_node_step(const DevScene &sc, AnyTrav &t, uint2 *stack, int sstride, TravCounters *cnt, uint32_t &tribits, uint32_t &tri_base, uint32_t &vmask) {
    tribits = 0u; tri_base = 0u; vmask = 0u;
    if (!trav_next(t.p, t.r.oct, stack, sstride)) return false;
    uint32_t hits_p, child_base, imask;
    node_test(sc, t.p.node, t.r, 1e-3f, t.tmax, hits_p, tribits, child_base, tri_base, imask, vmask);
    if (COUNT) cnt->nodes++;
    trav_push(t.p, hits_p, child_base, imask, stack, sstride);
    return true;
}

// true = blocked by an opaque object within [1e-3, tMax] (cpu_check_visibility, cpu_bdpt.cpp:82-107).  The plain loop
// (state in registers, no resumable struct): b2_intersect's any-hit kernel measured 629 vs 485 Mrays/s against a loop
// over any_step on the 10 M-triangle soup.
template <bool COUNT, bool OPAQUE = false>       // OPAQUE: transparent materials block too (sc.gpu_sem callers, b2_isect.cuh)
B2_HD bool occluded_bvh(const DevScene &sc, float3 ro, float3 rd, float tMax, uint2 *stack, int sstride, TravCounters *cnt) {
    const float tMin = 1e-3f;
    if (sc.n_nodes > 0) {
        const RayPre r = ray_pre(ro, rd);
        TravPos p; trav_reset(p, true);
        while (trav_next(p, r.oct, stack, sstride)) {
            uint32_t hits_p, tribits, child_base, tri_base, imask, vmask;
            node_test(sc, p.node, r, tMin, tMax, hits_p, tribits, child_base, tri_base, imask, vmask);
            if (COUNT) cnt->nodes++;
            while (tribits) {
                const int bit = ffs32(tribits) - 1;
                tribits &= tribits - 1u;
                const int ti = tri_of_bit(tri_base, vmask, bit);
                const float4 a = sc.tri[3 * ti], b = sc.tri[3 * ti + 1], c = sc.tri[3 * ti + 2];
                if (!OPAQUE && (f2u(b.w) & B2_MAT_TRANSPARENT)) continue;
                if (COUNT) cnt->tris++;
                float t;
                if (tri_test(ro, rd, mk3(a), mk3(b), mk3(c), tMin, tMax, t) && (OPAQUE || group_pass(sc, (int) f2u(c.w), ro, rd, tMin, tMax))) return true;
            }
            trav_push(p, hits_p, child_base, imask, stack, sstride);
        }
    }
    return occluded_spheres(sc, ro, rd, tMin, tMax);
}
