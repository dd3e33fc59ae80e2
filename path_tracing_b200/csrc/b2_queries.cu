// b2_queries.cu -- batch kernels behind b2_intersect* (ray queries) and the BSDF / MIS / Philox table entry
// points of include/b2pt.h.  One item per thread; the closest-hit / any-hit kernels are the same traversal
// code the renderer uses (b2_bvh.cuh), with shared-memory stacks.
#include <cub/cub.cuh>
#include "b2_queries.hpp"
#include "b2_bdpt.cuh"

namespace {

constexpr int Q_THREADS = 128;

__device__ __forceinline__ int public_id(const DevScene &sc, int prim) {
    if (prim == B2_PRIM_MISS) return -1;
    const int k = prim_kind(prim), i = prim_index(prim);
    if (k == B2_KIND_TRI) return sc.tri_obj[i];
    if (k == B2_KIND_SPHERE) return (int) __float_as_uint(sc.sph[2 * i + 1].w);
    return -2 - i;
}
__device__ __forceinline__ void warp_add64(unsigned long long *p, uint32_t v) {
    v = __reduce_add_sync(0xffffffffu, v);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(p, (unsigned long long) v);
}
__device__ __forceinline__ float3 ld3(const float *p, size_t i) { return make_float3(p[3 * i], p[3 * i + 1], p[3 * i + 2]); }
__device__ __forceinline__ void st3(float *p, size_t i, float3 v) { p[3 * i] = v.x; p[3 * i + 1] = v.y; p[3 * i + 2] = v.z; }

// ---- ray reordering ----
// A batch of incoherent rays is walked in the order of (cell of the origin on a grid over the scene box with ~256 rays per
// cell, direction octant, coarse direction inside the octant): <= 30-bit keys, one radix sort of (key, ray index) pairs, and the traversal kernels read ray perm[i] and write its
// result to slot perm[i].  Neighbouring lanes then start in the same part of the tree and fetch the same nodes (one L1 / L2
// line instead of 32 DRAM sectors); what every ray computes is untouched, so ids, t and visibility are bit-identical.
// Cost on C5 (100 M rays): keys + sort + the gathered ray loads, a few ms against ~100 ms of traversal.
#ifndef B2_SORT_MIN_RAYS
#define B2_SORT_MIN_RAYS (1u << 18)
#endif
#ifndef B2_SORT_MIN_TRIS
#define B2_SORT_MIN_TRIS 100000       // below: the tree sits in L1 / L2 whatever the order
#endif
__device__ __forceinline__ uint32_t spread7(uint32_t x) {           // bit k of a (<= 10-bit) value -> bit 3k
    x &= 0x3ffu;
    x = (x ^ (x << 16)) & 0xff0000ffu; x = (x ^ (x << 8)) & 0x0300f00fu; x = (x ^ (x << 4)) & 0x030c30c3u; x = (x ^ (x << 2)) & 0x09249249u;
    return x;
}
template <bool ANY>
__global__ void __launch_bounds__(256) k_ray_keys(DevScene sc, const float *__restrict__ org, const float *__restrict__ dir, uint32_t n, int cell_bits,
                                                  uint32_t *__restrict__ keys, uint32_t *__restrict__ idx) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float3 o = ld3(org, i), b = ld3(dir, i);
    const float3 d = ANY ? make_float3(b.x - o.x, b.y - o.y, b.z - o.z) : b;      // any hit: `dir` holds the far end of the segment
    const float3 ext = make_float3(sc.scene_max.x - sc.scene_min.x, sc.scene_max.y - sc.scene_min.y, sc.scene_max.z - sc.scene_min.z);
    const float qx = (o.x - sc.scene_min.x) / fmaxf(ext.x, 1e-30f), qy = (o.y - sc.scene_min.y) / fmaxf(ext.y, 1e-30f), qz = (o.z - sc.scene_min.z) / fmaxf(ext.z, 1e-30f);
    const float cells = (float) (1 << cell_bits);
    const uint32_t cx = (uint32_t) fminf(fmaxf(qx * cells, 0.0f), cells - 1.0f), cy = (uint32_t) fminf(fmaxf(qy * cells, 0.0f), cells - 1.0f), cz = (uint32_t) fminf(fmaxf(qz * cells, 0.0f), cells - 1.0f);
    const uint32_t oct = (d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u);
    // inside an octant: the direction's components relative to its largest one, 2 bits each (rays that leave a cell the same way are neighbours)
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z), am = fmaxf(fmaxf(ax, ay), fmaxf(az, 1e-30f));
    const uint32_t dsub = ((uint32_t) fminf(ax / am * 4.0f, 3.0f) << 4) | ((uint32_t) fminf(ay / am * 4.0f, 3.0f) << 2) | (uint32_t) fminf(az / am * 4.0f, 3.0f);
    keys[i] = ((spread7(cx) | (spread7(cy) << 1) | (spread7(cz) << 2)) << 9) | (oct << 6) | dsub;    // NaN coordinates land in cell 0: any order is a valid order
    idx[i] = i;
}

// MODE bit 0: any-hit; BRUTE: exhaustive kernels; COUNT: visit counters.
// BVH variants: ONE loop per thread in which a lane whose ray is decided stores its result and loads its next ray at once,
// while the other lanes of the warp keep walking.  With a ray loop around a traversal loop a warp waited for its longest ray
// every round: ncu on C5 showed 9.9 of 32 lanes active (incoherent rays, 40 wide nodes per ray on average, several
// times that for the longest).  B2_QUERY_REFILL=0 restores the nested loops for A/B runs.
#ifndef B2_QUERY_REFILL
#define B2_QUERY_REFILL 1
#endif
template <bool ANY, bool BRUTE, bool COUNT, bool FAST = false>     // FAST: the render's shadow-ray arithmetic (tri_test_fast + group_pass_fast)
__global__ void __launch_bounds__(Q_THREADS) k_intersect(DevScene sc, const float *__restrict__ org, const float *__restrict__ dir, size_t n,
                                                         int32_t *__restrict__ prim_id, float *__restrict__ t_out, unsigned long long *counters) {
    extern __shared__ uint2 s_stack[];
    TravCounters tc{ 0u, 0u };
    uint32_t rays = 0;
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (!BRUTE && B2_QUERY_REFILL) {
        uint2 *stack = s_stack + threadIdx.x;
        bool busy = false;
        ClosestTrav ct; AnyTrav at;
        for (;;) {
            if (!busy) {
                if (i >= n) break;
                const float3 a = ld3(org, i), b = ld3(dir, i);
                rays++;
                if (!ANY) closest_init(ct, sc, a, b);
                else { const ShadowRay sr = make_shadow_ray(a, b); any_init(at, sc, sr.o, sr.d, sr.tmax); }
                busy = true;
            }
            if (!ANY) {
                if (!closest_step<COUNT>(sc, ct, stack, blockDim.x, &tc)) {
                    const HitRec h = closest_finish(sc, ct);
                    prim_id[i] = public_id(sc, h.prim);
                    if (t_out) t_out[i] = h.t;
                    i += stride; busy = false;
                }
            } else {
                int r = any_step<COUNT, FAST>(sc, at, stack, blockDim.x, &tc);
                if (r != B2_TRAV_CONTINUE) {
                    if (r == B2_TRAV_EXHAUSTED && ((FAST && !sc.gpu_sem) ? occluded_spheres_fast(sc, at.r, 1e-3f, at.tmax) : occluded_spheres(sc, at.r.o, at.r.d, 1e-3f, at.tmax))) r = B2_TRAV_HIT;
                    prim_id[i] = r == B2_TRAV_HIT ? 0 : 1;
                    i += stride; busy = false;
                }
            }
        }
    } else {
        for (; i < n; i += stride) {
            const float3 a = ld3(org, i), b = ld3(dir, i);
            rays++;
            if (!ANY) {
                const HitRec h = BRUTE ? closest_hit_brute(sc, a, b) : closest_hit_bvh<COUNT>(sc, a, b, s_stack + threadIdx.x, blockDim.x, &tc);
                prim_id[i] = public_id(sc, h.prim);
                if (t_out) t_out[i] = h.t;
            } else {
                const ShadowRay sr = make_shadow_ray(a, b);
                const bool occ = BRUTE ? occluded_brute(sc, sr.o, sr.d, sr.tmax) : occluded_bvh<COUNT>(sc, sr.o, sr.d, sr.tmax, s_stack + threadIdx.x, blockDim.x, &tc);
                prim_id[i] = occ ? 0 : 1;
            }
        }
    }
    if (COUNT) {
        // every thread of a warp reaches here (the loops above have no early return)
        warp_add64(&counters[0], rays); warp_add64(&counters[1], tc.nodes); warp_add64(&counters[2], tc.tris);
    }
}

// Closest hit with the leaf triangles of a whole warp POOLED: every lane walks its own ray through the wide nodes, but the
// (ray, triangle) pairs the 32 node tests produce are numbered with a warp prefix sum and tested 32 at a time, whichever
// lane the ray belongs to.  ncu on C5 before (per-lane triangle loop): the triangle fetch + exact Moeller-Trumbore were 51 %
// of the executed warp instructions and 60 % of the stall samples at 2.6 - 4.9 active lanes, most of them waiting on three
// dependent 128-bit loads of one lane.  A candidate goes back to its ray through a 64-bit shared-memory atomicMin on
// {t bits, ~order}: min t, ties -> the later object, the acceptance rule of the per-lane loop (closer()), so ids and t stay
// bit-identical (the tests are the same tri_test calls against a tMax that is at least the final one).
// C5: 511 -> 557 Mrays/s, 9.9 -> 17.1 active lanes per instruction; what remains is load latency (long_scoreboard 4.7 per
// issue on the node and triangle fetches, DRAM at 1.9 TB/s of random 32-byte sectors, L2 hit rate 41 %).  A prefetch of the
// next node into L1 (CCTL.E.PF1 right after the node test) changed nothing (554) and was removed.
#ifndef B2_QUERY_POOL
#define B2_QUERY_POOL 1
#endif
__device__ __forceinline__ int nth_bit(uint32_t bits, int j) {          // position of the (j+1)-th set bit (j is small)
    for (; j > 0; --j) bits &= bits - 1u;
    return __ffs((int) bits) - 1;
}
#ifndef B2_CLOSEST_MIN_BLOCKS
#define B2_CLOSEST_MIN_BLOCKS 8      // 64 registers (16 B spilled): 8 CTAs = 32 warps per SM hide more of the DRAM latency than 6 at 80 registers (C5: 599 -> 685 Mrays/s)
#endif
template <bool COUNT>
__global__ void __launch_bounds__(Q_THREADS, B2_CLOSEST_MIN_BLOCKS) k_closest_pooled(DevScene sc, const float *__restrict__ org, const float *__restrict__ dir, size_t n,
                                                              int32_t *__restrict__ prim_id, float *__restrict__ t_out, unsigned long long *counters,
                                                              const uint32_t *__restrict__ perm) {        // perm: the i-th ray to walk (nullptr: ray i)
    extern __shared__ uint2 s_stack[];
    __shared__ unsigned long long s_key[Q_THREADS];
    __shared__ int s_ti[Q_THREADS];
    __shared__ int s_tie[Q_THREADS];
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, wbase = threadIdx.x & ~31;
    TravCounters tc{ 0u, 0u };
    uint32_t rays = 0;
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    uint2 *stack = s_stack + threadIdx.x;
    bool busy = false;
    ClosestTrav ct;
    ct.r.o = ct.r.d = make_float3(0.f, 0.f, 0.f); ct.best.t = 0.0f;
    for (;;) {
        if (!busy && i < n) {
            const size_t j = perm ? (size_t) perm[i] : i;
            closest_init(ct, sc, ld3(org, j), ld3(dir, j));
            rays++;
            busy = true;
        }
        if (!__any_sync(FULL, busy)) break;
        uint32_t tribits = 0u, tri_base = 0u, vmask = 0u;
        bool exhausted = false;
        if (busy) exhausted = !closest_node_step<COUNT>(sc, ct, stack, blockDim.x, &tc, tribits, tri_base, vmask);
        const int cnt = __popc(tribits);
        int incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += v; }
        const int total = __shfl_sync(FULL, incl, 31);
        if (total > 0) {
            s_key[threadIdx.x] = ~0ull; s_tie[threadIdx.x] = 0;
            __syncwarp();
            for (int base = 0; base < total; base += 32) {
                const int k = base + lane;
                int owner = 0;                                  // first lane whose inclusive count exceeds k
#pragma unroll
                for (int step = 16; step >= 1; step >>= 1) { const int v = __shfl_sync(FULL, incl, owner + step - 1); if (v <= k) owner += step; }
                owner = min(owner, 31);
                const int o_excl = __shfl_sync(FULL, incl - cnt, owner);
                const uint32_t o_bits = __shfl_sync(FULL, tribits, owner), o_base = __shfl_sync(FULL, tri_base, owner), o_vmask = __shfl_sync(FULL, vmask, owner);
                const float3 ro = make_float3(__shfl_sync(FULL, ct.r.o.x, owner), __shfl_sync(FULL, ct.r.o.y, owner), __shfl_sync(FULL, ct.r.o.z, owner));
                const float3 rd = make_float3(__shfl_sync(FULL, ct.r.d.x, owner), __shfl_sync(FULL, ct.r.d.y, owner), __shfl_sync(FULL, ct.r.d.z, owner));
                const float o_tmax = __shfl_sync(FULL, ct.best.t, owner);
                unsigned long long key = ~0ull;
                int ti = 0;
                if (k < total) {
                    ti = tri_of_bit(o_base, o_vmask, nth_bit(o_bits, k - o_excl));
                    const float4 a = sc.tri[3 * ti], b = sc.tri[3 * ti + 1], c = sc.tri[3 * ti + 2];
                    if (COUNT) tc.tris++;
                    float tt;
                    if (tri_test(ro, rd, mk3(a), mk3(b), mk3(c), 1e-4f, o_tmax, tt) && group_pass(sc, (int) f2u(c.w), ro, rd, 1e-4f, B2_T_INF)) {
                        key = ((unsigned long long) __float_as_uint(tt) << 32) | (unsigned long long) (~f2u(a.w));
                        const unsigned long long old = atomicMin(&s_key[wbase + owner], key);
                        // an exact tie -- with the ray's current hit, or between two candidates of this step (whichever arrives second
                        // sees the first in `old`) -- is resolved by the ray's own lane with the full rule (tie classes, accept_hit)
                        if (tt == o_tmax || (old != ~0ull && (uint32_t) (old >> 32) == __float_as_uint(tt))) s_tie[wbase + owner] = 1;
                    }
                }
                __syncwarp();
                if (key != ~0ull && s_key[wbase + owner] == key) s_ti[wbase + owner] = ti;      // one object = one order: the winner is unique
                __syncwarp();
            }
            if (cnt > 0) {
                if (s_tie[threadIdx.x]) closest_test_tris<false>(sc, ct, tribits, tri_base, vmask, &tc);
                else {
                    const unsigned long long key = s_key[threadIdx.x];
                    if (key != ~0ull) {                          // strictly closer than the current hit, and alone at its t
                        const int wi = s_ti[threadIdx.x];
                        ct.best.t = __uint_as_float((uint32_t) (key >> 32)); ct.best.prim = prim_make(B2_KIND_TRI, wi); ct.best_order = ~(uint32_t) key;
                    }
                }
            }
            __syncwarp();
        }
        if (busy && exhausted) {
            const HitRec h = closest_finish(sc, ct);
            const size_t j = perm ? (size_t) perm[i] : i;
            prim_id[j] = public_id(sc, h.prim);
            if (t_out) t_out[j] = h.t;
            i += stride; busy = false;
        }
    }
    if (COUNT) { warp_add64(&counters[0], rays); warp_add64(&counters[1], tc.nodes); warp_add64(&counters[2], tc.tris); }
}

// Any hit with the leaf triangles of a whole warp pooled: k_connect's trace phase (b2_render.cu) as a ray-query kernel.  Every
// lane walks its own segment through the wide nodes and takes its next segment the moment this one is decided; the candidate
// triangles of the 32 node tests are numbered with a warp prefix sum and tested 32 at a time by whichever lanes are free.  The
// answer does not depend on the order in which candidates are tested (any opaque hit inside the segment decides), so it is
// the answer of the per-lane loop, bit for bit.  ncu on C5 with the per-lane loop (k_intersect<1,...>): 9.7 of 32 lanes active.
#ifndef B2_QUERY_ANY_POOL
#define B2_QUERY_ANY_POOL 1
#endif
#ifndef B2_QUERY_ANY_POOL_MIN_TRIS
#define B2_QUERY_ANY_POOL_MIN_TRIS 2048     // below: a segment meets 1-2 triangles, the per-lane loop wins (as in k_connect)
#endif
#ifndef B2_ANY_MIN_BLOCKS
#define B2_ANY_MIN_BLOCKS 8          // 64 registers, as ptxas chooses without a bound; pinned (9 / 10 / 12 CTAs per SM measured slower: 1186 / 922 Mrays/s)
#endif
template <bool COUNT, bool FAST>
__global__ void __launch_bounds__(Q_THREADS, B2_ANY_MIN_BLOCKS) k_any_pooled(DevScene sc, const float *__restrict__ p1, const float *__restrict__ p2, size_t n,
                                                          int32_t *__restrict__ vis_out, unsigned long long *counters, const uint32_t *__restrict__ perm) {
    extern __shared__ uint2 s_stack[];
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    TravCounters tc{ 0u, 0u };
    uint32_t rays = 0;
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    uint2 *stack = s_stack + threadIdx.x;
    bool busy = false;
    AnyTrav at;
    at.r.o = at.r.d = at.r.idir = make_float3(0.f, 0.f, 0.f); at.r.oct = 0u; at.tmax = 0.0f;
    for (;;) {
        if (!busy && i < n) {
            const size_t j = perm ? (size_t) perm[i] : i;
            const ShadowRay sr = make_shadow_ray(ld3(p1, j), ld3(p2, j));
            any_init(at, sc, sr.o, sr.d, sr.tmax);
            rays++;
            busy = true;
        }
        if (!__any_sync(FULL, busy)) break;
        uint32_t tribits = 0u, tri_base = 0u, vmask = 0u;
        bool exhausted = false;
        if (busy) exhausted = !any_node_step<COUNT>(sc, at, stack, blockDim.x, &tc, tribits, tri_base, vmask);
        const int cnt = __popc(tribits);
        int incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += v; }
        const int total = __shfl_sync(FULL, incl, 31);
        uint32_t blocked = 0u;                                  // bit j: the segment of lane j met an opaque triangle
        for (int base = 0; base < total; base += 32) {
            const int k = base + lane;
            int owner = 0;
#pragma unroll
            for (int step = 16; step >= 1; step >>= 1) { const int v = __shfl_sync(FULL, incl, owner + step - 1); if (v <= k) owner += step; }
            owner = min(owner, 31);
            const int o_excl = __shfl_sync(FULL, incl - cnt, owner);
            const uint32_t o_bits = __shfl_sync(FULL, tribits, owner), o_base = __shfl_sync(FULL, tri_base, owner), o_vmask = __shfl_sync(FULL, vmask, owner);
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a, c = a;
            if (k < total) {
                const int ti = tri_of_bit(o_base, o_vmask, nth_bit(o_bits, k - o_excl));
                a = sc.tri[3 * ti]; b = sc.tri[3 * ti + 1]; c = sc.tri[3 * ti + 2];
            }
            RayPre r;
            r.o = make_float3(__shfl_sync(FULL, at.r.o.x, owner), __shfl_sync(FULL, at.r.o.y, owner), __shfl_sync(FULL, at.r.o.z, owner));
            r.d = make_float3(__shfl_sync(FULL, at.r.d.x, owner), __shfl_sync(FULL, at.r.d.y, owner), __shfl_sync(FULL, at.r.d.z, owner));
            const float o_tmax = __shfl_sync(FULL, at.tmax, owner);
            if (FAST) r.idir = make_float3(__shfl_sync(FULL, at.r.idir.x, owner), __shfl_sync(FULL, at.r.idir.y, owner), __shfl_sync(FULL, at.r.idir.z, owner));
            bool hit = false;
            if (k < total && !(f2u(b.w) & B2_MAT_TRANSPARENT)) {
                if (COUNT) tc.tris++;
                float tt;
                if (FAST) hit = tri_test_fast(r.o, r.d, mk3(a), mk3(b), mk3(c), 1e-3f, o_tmax) && group_pass_fast(sc, (int) f2u(c.w), r, 1e-3f, o_tmax);
                else hit = tri_test(r.o, r.d, mk3(a), mk3(b), mk3(c), 1e-3f, o_tmax, tt) && group_pass(sc, (int) f2u(c.w), r.o, r.d, 1e-3f, o_tmax);
            }
            blocked |= __reduce_or_sync(FULL, hit ? 1u << owner : 0u);
        }
        if (busy) {
            const bool occ = (blocked >> lane) & 1u;
            if (occ || exhausted) {
                bool o2 = occ;
                if (!o2) o2 = (FAST && !sc.gpu_sem) ? occluded_spheres_fast(sc, at.r, 1e-3f, at.tmax) : occluded_spheres(sc, at.r.o, at.r.d, 1e-3f, at.tmax);
                vis_out[perm ? (size_t) perm[i] : i] = o2 ? 0 : 1;
                i += stride; busy = false;
            }
        }
    }
    if (COUNT) { warp_add64(&counters[0], rays); warp_add64(&counters[1], tc.nodes); warp_add64(&counters[2], tc.tris); }
}

__device__ __forceinline__ Mat ld_mat(const float *m, size_t i) {
    Mat r; r.base = make_float3(m[6 * i], m[6 * i + 1], m[6 * i + 2]); r.roughness = m[6 * i + 3]; r.metallic = m[6 * i + 4]; r.eta = m[6 * i + 5];
    return r;
}
__global__ void k_bsdf_evaluate(size_t n, const float *mtl, const float *wo, const float *wi, const float *N, float *out) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) st3(out, i, bsdf_evaluate(ld_mat(mtl, i), ld3(wo, i), ld3(wi, i), ld3(N, i)));
}
__global__ void k_bsdf_pdf(size_t n, const float *mtl, const float *wo, const float *wi, const float *N, float *out) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = bsdf_pdf(ld_mat(mtl, i), ld3(wo, i), ld3(wi, i), ld3(N, i));
}
__global__ void k_bsdf_sample(size_t n, const float *mtl, const float *wo, const float *N, const float *u, const float *cur_eta,
                              float *wi_out, float *f_out, float *pdf_out, int32_t *delta_out, float *eta_out) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    BsdfSample s; s.wi = make_float3(0, 0, 0); s.f = make_float3(0, 0, 0); s.pdf = 0.0f; s.new_eta = 0.0f; s.is_delta = false;
    bsdf_sample(ld_mat(mtl, i), ld3(wo, i), ld3(N, i), u[3 * i], u[3 * i + 1], u[3 * i + 2], cur_eta[i], s);
    st3(wi_out, i, s.wi); st3(f_out, i, s.f); pdf_out[i] = s.pdf; delta_out[i] = s.is_delta ? 1 : 0; eta_out[i] = s.new_eta;
}

// flat vertex accessors: eye (17) pos3 normal3 thr3 mtl6 pdf_fwd pdf_rev; light (20) ... is_src is_parallel cutoff pdf_fwd pdf_rev
struct FlatEyeAcc {
    const float *v; int s;
    __device__ float3 pos(int i) const { return make_float3(v[17 * i], v[17 * i + 1], v[17 * i + 2]); }
    __device__ float eta(int i) const { return v[17 * i + 14]; }
    __device__ float pdf_fwd(int i) const { return v[17 * i + 15]; }
    __device__ float pdf_rev(int i) const { return v[17 * i + 16]; }
    __device__ float3 cur_n() const { return make_float3(v[17 * s + 3], v[17 * s + 4], v[17 * s + 5]); }
    __device__ Mat cur_mtl() const { return ld_mat(v + 17 * (size_t) s + 9, 0); }
};
struct FlatLightAcc {
    const float *v; int t;
    __device__ float3 pos(int i) const { return make_float3(v[20 * i], v[20 * i + 1], v[20 * i + 2]); }
    __device__ float eta(int i) const { return v[20 * i + 14]; }
    __device__ bool is_src(int i) const { return v[20 * i + 15] != 0.0f; }
    __device__ float pdf_fwd(int i) const { return v[20 * i + 18]; }
    __device__ float pdf_rev(int i) const { return v[20 * i + 19]; }
    __device__ float3 cur_n() const { return make_float3(v[20 * t + 3], v[20 * t + 4], v[20 * t + 5]); }
    __device__ Mat cur_mtl() const { return ld_mat(v + 20 * (size_t) t + 9, 0); }
};
__global__ void k_mis_weight(size_t n, const float *eye17, const int32_t *n_eye, const int32_t *s_idx, const float *lit20, const int32_t *n_lit,
                             const int32_t *t_idx, const float *d_vec, const float *dist2, const float *cam_pos, int max_verts, float *out) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int s = s_idx[i], t = t_idx[i];
    if (s < 0 || t < 0 || s >= n_eye[i] || t >= n_lit[i] || s >= max_verts || t >= max_verts) { out[i] = 0.0f; return; }
    const FlatEyeAcc E{ eye17 + (size_t) 17 * max_verts * i, s };
    const FlatLightAcc L{ lit20 + (size_t) 20 * max_verts * i, t };
    out[i] = mis_weight_core(E, s, L, t, ld3(d_vec, i), dist2[i], ld3(cam_pos, i));
}
__global__ void k_philox(size_t n, const uint32_t *ctr4, const uint32_t *key2, uint32_t *out4, float *unit4) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 o = philox4x32_10(make_uint4(ctr4[4 * i], ctr4[4 * i + 1], ctr4[4 * i + 2], ctr4[4 * i + 3]), make_uint2(key2[2 * i], key2[2 * i + 1]));
    out4[4 * i] = o.x; out4[4 * i + 1] = o.y; out4[4 * i + 2] = o.z; out4[4 * i + 3] = o.w;
    if (unit4) { unit4[4 * i] = u32_to_unit(o.x); unit4[4 * i + 1] = u32_to_unit(o.y); unit4[4 * i + 2] = u32_to_unit(o.z); unit4[4 * i + 3] = u32_to_unit(o.w); }
}

template <bool ANY, bool BRUTE, bool COUNT, bool FAST = false>
int launch_isect(const DevScene &sc, int bvh_depth, const float *o, const float *d, size_t n, int32_t *id, float *t, unsigned long long *cnt, cudaStream_t st,
                 RaySortScratch *ss, bool keep_order) {
    // large batches against large scenes are walked in the order of their (origin cell, direction octant) keys
    const uint32_t *perm = nullptr;
    // (closest hits only: C5 100 M rays 695 -> 775 Mrays/s, keys and sort included; any-hit segments end after a few nodes and
    //  do not pay the sort back: 1569 -> 1544 Mrays/s.  B2_SORT_ANY=1 builds the any-hit variant for A/B runs.)
#ifndef B2_SORT_ANY
#define B2_SORT_ANY 0
#endif
    if (!BRUTE && (!ANY || B2_SORT_ANY) && ss && !keep_order && n >= B2_SORT_MIN_RAYS && n < ((size_t) 1 << 31) && sc.n_tris >= B2_SORT_MIN_TRIS) {
        const int n32 = (int) n;
        B2_TRY(ss->keys.ensure(n)); B2_TRY(ss->keys_sorted.ensure(n)); B2_TRY(ss->idx.ensure(n)); B2_TRY(ss->idx_sorted.ensure(n));
        // cells per axis: ~256 rays per cell (32 per octant), between 8 and 128
        int cell_bits = 3;
        while (cell_bits < 7 && ((size_t) 1 << (3 * (cell_bits + 1))) * 256 <= n) cell_bits++;
        const int end_bit = 3 * cell_bits + 9;
        size_t tmp_bytes = 0;
        B2_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, ss->keys.p, ss->keys_sorted.p, ss->idx.p, ss->idx_sorted.p, n32, 0, end_bit, st));
        B2_TRY(ss->tmp.ensure(tmp_bytes));
        k_ray_keys<ANY><<<(unsigned) ((n + 255) / 256), 256, 0, st>>>(sc, o, d, (uint32_t) n, cell_bits, ss->keys.p, ss->idx.p);
        B2_CUDA(cub::DeviceRadixSort::SortPairs(ss->tmp.p, tmp_bytes, ss->keys.p, ss->keys_sorted.p, ss->idx.p, ss->idx_sorted.p, n32, 0, end_bit, st));
        perm = ss->idx_sorted.p;
    }
    const size_t sb = stack_bytes(Q_THREADS, bvh_depth);
    static size_t configured = 0;
    if (sb > 48 * 1024 && sb > configured) {
        B2_CUDA(cudaFuncSetAttribute(k_intersect<ANY, BRUTE, COUNT, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb));
    }
    // grid: enough blocks to fill every SM's resident CTAs, grid-stride beyond that
    const size_t want = (n + Q_THREADS - 1) / Q_THREADS;
    const unsigned grid = (unsigned) std::min<size_t>(want, (size_t) (sc.n_sms > 0 ? sc.n_sms : 148) * 64);
    if (!ANY && !BRUTE && B2_QUERY_POOL) {
        if (sb + 2048 > 48 * 1024) B2_CUDA(cudaFuncSetAttribute(k_closest_pooled<COUNT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb));
        k_closest_pooled<COUNT><<<grid, Q_THREADS, sb, st>>>(sc, o, d, n, id, t, cnt, perm);
    } else if (ANY && !BRUTE && B2_QUERY_ANY_POOL && sc.n_tris >= B2_QUERY_ANY_POOL_MIN_TRIS) {
        if (sb > 48 * 1024) B2_CUDA(cudaFuncSetAttribute(k_any_pooled<COUNT, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb));
        k_any_pooled<COUNT, FAST><<<grid, Q_THREADS, sb, st>>>(sc, o, d, n, id, cnt, perm);
    } else
        k_intersect<ANY, BRUTE, COUNT, FAST><<<grid, Q_THREADS, sb, st>>>(sc, o, d, n, id, t, cnt);
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}

}  // namespace

int b2_sort_reserve(RaySortScratch &ss, size_t n) {
    if (n >= ((size_t) 1 << 31)) { b2_set_error("more than 2^31 sort keys"); return B2_ERR_INVALID; }
    B2_TRY(ss.keys.ensure(n)); B2_TRY(ss.keys_sorted.ensure(n)); B2_TRY(ss.idx.ensure(n)); B2_TRY(ss.idx_sorted.ensure(n));
    size_t tmp_bytes = 0;
    B2_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, ss.keys.p, ss.keys_sorted.p, ss.idx.p, ss.idx_sorted.p, (int) n, 0, 32, 0));
    return ss.tmp.ensure(tmp_bytes);
}
int b2_sort_pairs(RaySortScratch &ss, size_t n, int end_bit, cudaStream_t st) {
    size_t tmp_bytes = ss.tmp.n;
    B2_CUDA(cub::DeviceRadixSort::SortPairs(ss.tmp.p, tmp_bytes, ss.keys.p, ss.keys_sorted.p, ss.idx.p, ss.idx_sorted.p, (int) n, 0, end_bit, st));
    return B2_OK;
}

int b2_intersect_launch(const DevScene &sc, int bvh_depth, const float *d_org, const float *d_dir, size_t n, int mode,
                        int32_t *d_prim_id, float *d_t, unsigned long long *d_counters, cudaStream_t st, RaySortScratch *ss) {
    if (n == 0) return B2_OK;
    const bool any = mode & B2_ANY, brute = mode & B2_BRUTE_FORCE, count = (mode & B2_COUNT_VISITS) && !brute, keep = (mode & B2_KEEP_RAY_ORDER) != 0;
    if (!any) {
        if (brute) return launch_isect<false, true, false>(sc, bvh_depth, d_org, d_dir, n, d_prim_id, d_t, d_counters, st, ss, keep);
        if (count) return launch_isect<false, false, true>(sc, bvh_depth, d_org, d_dir, n, d_prim_id, d_t, d_counters, st, ss, keep);
        return launch_isect<false, false, false>(sc, bvh_depth, d_org, d_dir, n, d_prim_id, d_t, d_counters, st, ss, keep);
    }
    if (brute) return launch_isect<true, true, false>(sc, bvh_depth, d_org, d_dir, n, d_prim_id, d_t, d_counters, st, ss, keep);
    if (mode & B2_FAST_SHADOW) return launch_isect<true, false, false, true>(sc, bvh_depth, d_org, d_dir, n, d_prim_id, d_t, d_counters, st, ss, keep);
    if (count) return launch_isect<true, false, true>(sc, bvh_depth, d_org, d_dir, n, d_prim_id, d_t, d_counters, st, ss, keep);
    return launch_isect<true, false, false>(sc, bvh_depth, d_org, d_dir, n, d_prim_id, d_t, d_counters, st, ss, keep);
}

// ---- table entry points: host arrays in, host arrays out (device kernels in between) ----
namespace {
struct Up {          // upload helper
    DevBuf<unsigned char> b;
    template <class T> int put(const T *h, size_t count, T **d) {
        B2_TRY(b.alloc(std::max<size_t>(count, 1) * sizeof(T)));
        if (count) B2_CUDA(cudaMemcpy(b.p, h, count * sizeof(T), cudaMemcpyHostToDevice));
        *d = (T *) b.p;
        return B2_OK;
    }
    template <class T> int make(size_t count, T **d) { B2_TRY(b.alloc(std::max<size_t>(count, 1) * sizeof(T))); *d = (T *) b.p; return B2_OK; }
};
template <class T> int down(T *h, const T *d, size_t count) { if (count) B2_CUDA(cudaMemcpy(h, d, count * sizeof(T), cudaMemcpyDeviceToHost)); return B2_OK; }
}  // namespace

int b2_tables_bsdf_evaluate(size_t n, const float *mtl6, const float *wo, const float *wi, const float *N, float *out3) {
    Up a, b, c, d, o; float *dm, *dwo, *dwi, *dn, *dout;
    B2_TRY(a.put(mtl6, 6 * n, &dm)); B2_TRY(b.put(wo, 3 * n, &dwo)); B2_TRY(c.put(wi, 3 * n, &dwi)); B2_TRY(d.put(N, 3 * n, &dn)); B2_TRY(o.make(3 * n, &dout));
    if (n) k_bsdf_evaluate<<<blocks_for(n, 128), 128>>>(n, dm, dwo, dwi, dn, dout);
    B2_CUDA(cudaGetLastError()); B2_CUDA(cudaDeviceSynchronize());
    return down(out3, dout, 3 * n);
}
int b2_tables_bsdf_pdf(size_t n, const float *mtl6, const float *wo, const float *wi, const float *N, float *out) {
    Up a, b, c, d, o; float *dm, *dwo, *dwi, *dn, *dout;
    B2_TRY(a.put(mtl6, 6 * n, &dm)); B2_TRY(b.put(wo, 3 * n, &dwo)); B2_TRY(c.put(wi, 3 * n, &dwi)); B2_TRY(d.put(N, 3 * n, &dn)); B2_TRY(o.make(n, &dout));
    if (n) k_bsdf_pdf<<<blocks_for(n, 128), 128>>>(n, dm, dwo, dwi, dn, dout);
    B2_CUDA(cudaGetLastError()); B2_CUDA(cudaDeviceSynchronize());
    return down(out, dout, n);
}
int b2_tables_bsdf_sample(size_t n, const float *mtl6, const float *wo, const float *N, const float *u3, const float *cur_eta,
                          float *wi_out, float *f_out, float *pdf_out, int32_t *delta_out, float *eta_out) {
    Up a, b, c, d, e, o1, o2, o3, o4, o5; float *dm, *dwo, *dn, *du, *de, *dwi, *df, *dp, *dne; int32_t *dd;
    B2_TRY(a.put(mtl6, 6 * n, &dm)); B2_TRY(b.put(wo, 3 * n, &dwo)); B2_TRY(c.put(N, 3 * n, &dn)); B2_TRY(d.put(u3, 3 * n, &du)); B2_TRY(e.put(cur_eta, n, &de));
    B2_TRY(o1.make(3 * n, &dwi)); B2_TRY(o2.make(3 * n, &df)); B2_TRY(o3.make(n, &dp)); B2_TRY(o4.make(n, &dd)); B2_TRY(o5.make(n, &dne));
    if (n) k_bsdf_sample<<<blocks_for(n, 128), 128>>>(n, dm, dwo, dn, du, de, dwi, df, dp, dd, dne);
    B2_CUDA(cudaGetLastError()); B2_CUDA(cudaDeviceSynchronize());
    B2_TRY(down(wi_out, dwi, 3 * n)); B2_TRY(down(f_out, df, 3 * n)); B2_TRY(down(pdf_out, dp, n)); B2_TRY(down(delta_out, dd, n));
    return down(eta_out, dne, n);
}
int b2_tables_mis_weight(size_t n, const float *eye17, const int32_t *n_eye, const int32_t *s_idx, const float *lit20, const int32_t *n_lit,
                         const int32_t *t_idx, const float *d_vec, const float *dist2, const float *cam_pos, int32_t max_verts, float *out) {
    if (max_verts <= 0) { b2_set_error("max_verts must be > 0"); return B2_ERR_INVALID; }
    Up a, b, c, d, e, f, g, h, i, o; float *de, *dl, *dd, *d2, *dc, *dout; int32_t *dne, *ds, *dnl, *dt;
    B2_TRY(a.put(eye17, (size_t) 17 * max_verts * n, &de)); B2_TRY(b.put(n_eye, n, &dne)); B2_TRY(c.put(s_idx, n, &ds));
    B2_TRY(d.put(lit20, (size_t) 20 * max_verts * n, &dl)); B2_TRY(e.put(n_lit, n, &dnl)); B2_TRY(f.put(t_idx, n, &dt));
    B2_TRY(g.put(d_vec, 3 * n, &dd)); B2_TRY(h.put(dist2, n, &d2)); B2_TRY(i.put(cam_pos, 3 * n, &dc)); B2_TRY(o.make(n, &dout));
    if (n) k_mis_weight<<<blocks_for(n, 128), 128>>>(n, de, dne, ds, dl, dnl, dt, dd, d2, dc, max_verts, dout);
    B2_CUDA(cudaGetLastError()); B2_CUDA(cudaDeviceSynchronize());
    return down(out, dout, n);
}
int b2_tables_philox(size_t n, const uint32_t *ctr4, const uint32_t *key2, uint32_t *out4, float *unit_out4) {
    Up a, b, o, u; uint32_t *dc, *dk, *dout; float *du;
    B2_TRY(a.put(ctr4, 4 * n, &dc)); B2_TRY(b.put(key2, 2 * n, &dk)); B2_TRY(o.make(4 * n, &dout)); B2_TRY(u.make(4 * n, &du));
    if (n) k_philox<<<blocks_for(n, 128), 128>>>(n, dc, dk, dout, du);
    B2_CUDA(cudaGetLastError()); B2_CUDA(cudaDeviceSynchronize());
    B2_TRY(down(out4, dout, 4 * n));
    if (unit_out4) B2_TRY(down(unit_out4, du, 4 * n));
    return B2_OK;
}
