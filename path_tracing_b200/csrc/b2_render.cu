// b2_render.cu -- wavefront BDPT on the device: __global__ wrappers around the stage bodies of b2_bdpt.cuh and
// the host wave loop behind b2_render_bdpt / b2_render_bdpt_device / b2_light_pool (include/b2pt.h).
//
// Replaces run_cpu_bdpt (src/cpu_bdpt.cpp:173-488) / run_cuda_bdpt (include/bdpt_cu_helper.h:6).  One render:
//
//   k_light_paths  one thread per light subpath -> pool (5 x float4 per vertex)              cpu_bdpt.cpp:218-325
//   k_pool_active  ordered compaction of the pool vertices that can contribute (:391) into 48-byte pre-test records
//   per batch of <= max_paths eye samples, per wave (one bounce of every live path):
//     k_extend     wave 0: primary ray generated in registers (:349-361); later waves: path state from HBM.
//                  Closest-hit traversal of the 8-wide BVH with shared-memory stacks (:369), then the bounce
//                  (:372-385, :442-472): eye-vertex record appended to a COMPACT list, BSDF sample, next ray.
//                  Path slots, eye-vertex slots and the next wave's queue are handed out with warp ballot +
//                  prefix popcount and one atomic per warp.  A sample whose primary ray escapes never touches HBM.
//     k_connect    eye vertex x pool (:389-439), one warp per 32 eye vertices, warp-autonomous: every lane
//                  pre-tests its vertex against the pool records (uniform broadcast loads); surviving pairs are
//                  compacted into the warp's shared-memory ring and finished 32 at a time at full occupancy
//                  (BSDFs, shadow any-hit, MIS); per-vertex sums go to the film with one atomicAdd per channel.
//   k_resolve      film / spp                                                                 :479
#include "b2_render.hpp"
#include "b2_strategies.cuh"
#include "b2_warp.cuh"
#include <algorithm>
#include <cstdlib>
#include <vector>

namespace {

constexpr int EXT_THREADS = 128;      // extend kernel block
constexpr int CONN_THREADS = 128;     // connect kernel block: 4 independent warps
constexpr int CONN_WARPS = CONN_THREADS / 32;
#ifndef B2_CONN_BLOCKS
#define B2_CONN_BLOCKS 6
#endif
constexpr int CONN_BLOCKS_PER_SM = B2_CONN_BLOCKS; // persistent grid = 148 SMs x this (also the register budget: launch bounds)
#ifndef B2_CONN_CHUNK
#define B2_CONN_CHUNK 4
#endif
#ifndef B2_MIS_WALKS_HOISTED
#define B2_MIS_WALKS_HOISTED 1        // MIS weight: the light-side walk once per pool vertex (k_pool_active), the eye-side walk once per
#endif                                // eye vertex and work item (k_connect), instead of both once per visible pair (0: round-1 form)
constexpr int CONN_CHUNK = B2_CONN_CHUNK;         // pool vertices pre-tested per compaction round; the pre-test table is padded to a
                                      // multiple of this with records that never pass, and pool slices are whole chunks

template <bool COUNT>
__global__ void __launch_bounds__(EXT_THREADS) k_light_paths(RenderCtx c, int n_paths) {
    extern __shared__ uint2 s_stack[];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    TravCounters tc{ 0u, 0u };
    uint32_t ncl = 0, ncap = 0;
    if (i < n_paths) light_path_body<COUNT>(c, i, s_stack + threadIdx.x, blockDim.x, &tc, ncl, ncap);
    warp_count_add(&c.counters[B2_CNT_CLOSEST], ncl);
    warp_count_add(&c.counters[B2_CNT_DELTA_CAP], ncap);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}

// 30-bit Morton code of a position inside the scene box (NaN / outside: clamped, any order is a valid order)
__device__ __forceinline__ uint32_t spread10(uint32_t x) {
    x &= 0x3ffu;
    x = (x | (x << 16)) & 0x030000ffu; x = (x | (x << 8)) & 0x0300f00fu; x = (x | (x << 4)) & 0x030c30c3u; x = (x | (x << 2)) & 0x09249249u;
    return x;
}
__device__ __forceinline__ uint32_t morton30(const DevScene &sc, float x, float y, float z) {
    const float3 lo = sc.scene_min, hi = sc.scene_max;
    const float qx = (x - lo.x) / fmaxf(hi.x - lo.x, 1e-30f), qy = (y - lo.y) / fmaxf(hi.y - lo.y, 1e-30f), qz = (z - lo.z) / fmaxf(hi.z - lo.z, 1e-30f);
    const uint32_t cx = (uint32_t) fminf(fmaxf(qx * 1024.0f, 0.0f), 1023.0f), cy = (uint32_t) fminf(fmaxf(qy * 1024.0f, 0.0f), 1023.0f), cz = (uint32_t) fminf(fmaxf(qz * 1024.0f, 0.0f), 1023.0f);
    return spread10(cx) | (spread10(cy) << 1) | (spread10(cz) << 2);
}
#ifndef B2_POOL_SORT
#define B2_POOL_SORT 0                // 1: the active pool vertices (<= 2048 of them) in position order instead of pool order
#endif
#ifndef B2_CV_KEY_NORMAL
#define B2_CV_KEY_NORMAL 0            // 1: eye-vertex sort key = 27-bit position code + octant of the normal
#endif
// Single block: pre-test records of the pool vertices with |throughput| >= 1e-6, in ascending pool order (deterministic).
__global__ void __launch_bounds__(1024) k_pool_active(RenderCtx c) {
    __shared__ int s_warp[32];
    __shared__ int s_base;
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    for (int i0 = 0; i0 < c.n_pool; i0 += blockDim.x) {
        const int i = i0 + threadIdx.x;
        const bool act = i < c.n_pool && pool_vertex_active(c.pool, i);
        const uint32_t m = __ballot_sync(0xffffffffu, act);
        if (lane == 0) s_warp[warp] = __popc(m);
        __syncthreads();
        int off = 0, tot = 0;
        for (int w = 0; w < nwarps; ++w) { const int v = s_warp[w]; if (w < warp) off += v; tot += v; }
        if (act) {
            make_pretest_rec(c, i, c.pre + (size_t) B2_PRE_F4 * (s_base + off + __popc(m & lanemask_lt())));
#if B2_MIS_WALKS_HOISTED
            // the light-side walk of the MIS weight, once per pool vertex instead of once per visible pair (b2_bdpt.cuh: mis_light_sum)
            const LightVtx lv = load_lv(c.pool, i);
            const int t_idx = i % c.light_depth;
            const float ls = mis_light_sum(RenderLightAcc{ c, lv, i - t_idx }, t_idx);
            reinterpret_cast<float *>(c.pool + B2_LV_F4 * i + 4)[3] = isfinite(ls) ? ls : 3.0e38f;
#endif
        }
        __syncthreads();
        if (threadIdx.x == 0) s_base += tot;
        __syncthreads();
    }
    // pad to a whole chunk with records no eye vertex can pass (zero normal: dot(n, d) < 0 is false)
    const int n = s_base, n_pad = (n + CONN_CHUNK - 1) / CONN_CHUNK * CONN_CHUNK;
    if (threadIdx.x < n_pad - n) {
        float4 *rec = c.pre + (size_t) B2_PRE_F4 * (n + threadIdx.x);
        rec[0] = make_float4(0.0f, 0.0f, 0.0f, B2_NO_GATE); rec[1] = make_float4(0.0f, 0.0f, 0.0f, 0.0f); rec[2] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
    if (threadIdx.x == 0) *c.n_pool_active = n;
#if B2_POOL_SORT
    // position order: the pool vertices a warp's 32 (neighbouring) eye vertices meet in one pre-test window are neighbours too
    __shared__ uint32_t s_key[2048];
    if (n > 1 && n <= 2048) {
        float4 r[2][B2_PRE_F4]; int rank[2] = { 0, 0 };
        for (int k = 0; k < 2; ++k) {
            const int t = threadIdx.x + k * 1024;
            if (t < n) {
                for (int j = 0; j < B2_PRE_F4; ++j) r[k][j] = c.pre[(size_t) B2_PRE_F4 * t + j];
                s_key[t] = morton30(c.sc, r[k][0].x, r[k][0].y, r[k][0].z);
            }
        }
        __syncthreads();
        for (int k = 0; k < 2; ++k) {
            const int t = threadIdx.x + k * 1024;
            if (t < n) { const uint32_t kt = s_key[t]; int cnt = 0; for (int u = 0; u < n; ++u) { const uint32_t ku = s_key[u]; cnt += (ku < kt || (ku == kt && u < t)) ? 1 : 0; } rank[k] = cnt; }
        }
        __syncthreads();
        for (int k = 0; k < 2; ++k) {
            const int t = threadIdx.x + k * 1024;
            if (t < n) for (int j = 0; j < B2_PRE_F4; ++j) c.pre[(size_t) B2_PRE_F4 * rank[k] + j] = r[k][j];
        }
    }
#endif
}

// Connection stage on position-sorted eye vertices (B2_CONN_SORT_CV; A/B in profiles/r02_connect_summary.md, section 5): a work
// item of k_connect is 32 CONSECUTIVE records, whose shadow rays to one pool vertex share BVH nodes only if the 32 vertices are
// close together -- true for primary hits of neighbouring pixels, not after a bounce.  Key = 30-bit Morton code of the position
// inside the scene box; records past n_cv (the launch is sized by the host's upper bound n_live) sort last.
__global__ void __launch_bounds__(256) k_cv_keys(RenderCtx c, uint32_t n_upper, uint32_t *__restrict__ keys, uint32_t *__restrict__ idx) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_upper) return;
    uint32_t key = 1u << 30;
    if ((int) i < *c.n_cv) {
        const float4 f0 = c.cv[B2_CV_F4 * (size_t) i];
        key = morton30(c.sc, f0.x, f0.y, f0.z);
#if B2_CV_KEY_NORMAL
        const float4 f1 = c.cv[B2_CV_F4 * (size_t) i + 1];
        key = (key & ~7u) | (f1.x < 0.0f ? 1u : 0u) | (f1.y < 0.0f ? 2u : 0u) | (f1.z < 0.0f ? 4u : 0u);   // lowest position bits -> normal octant
#endif
    }
    keys[i] = key; idx[i] = i;
}
__global__ void __launch_bounds__(256) k_cv_gather(RenderCtx c, uint32_t n_upper, const uint32_t *__restrict__ order, float4 *__restrict__ out) {
    const size_t t = (size_t) blockIdx.x * blockDim.x + threadIdx.x;          // one 16-byte row per thread
    const uint32_t rec = (uint32_t) (t / B2_CV_F4), row = (uint32_t) (t % B2_CV_F4);
    if (rec >= n_upper || (int) rec >= *c.n_cv) return;
    out[t] = c.cv[(size_t) B2_CV_F4 * order[rec] + row];
}

// One bounce of every live path.  PRIMARY: item = sample w0 + i of the batch (no state in HBM yet).
template <bool PRIMARY, bool COUNT>
__global__ void __launch_bounds__(EXT_THREADS) k_extend(RenderCtx c, int n_items, uint32_t pix0, int s0) {
    extern __shared__ uint2 s_stack[];
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = q < n_items;
    TravCounters tc{ 0u, 0u };
    PathState ps;
    ShadeOut so; so.has_cv = so.alive = so.store_hist = so.emitter_hit = so.delta_cap = false;
    int path = -1;
    if (active) {
        if (PRIMARY) ps = primary_path(c, pix0, s0, (uint32_t) q);
        else { path = c.live[q]; ps = load_path(c, path); }
        const HitRec h = closest_hit_bvh<COUNT>(c.sc, ps.ro, ps.rd, s_stack + threadIdx.x, blockDim.x, &tc);
        shade_step(c, ps, h, so);
    }
    if (PRIMARY) path = warp_alloc(c.n_paths, so.has_cv);          // a path gets its slot with its first eye vertex
    const int cvi = warp_alloc(c.n_cv, so.has_cv);
    if (so.has_cv) {
        so.ev.path = path;
        store_cv(c.cv, cvi, so.ev);
        if (so.store_hist) {
            c.hist_pe[(size_t) so.ev.depth * c.max_paths + path] = mk4(so.ev.pos, so.ev.mtl.eta);
            c.hist_pdf[(size_t) so.ev.depth * c.max_paths + path] = make_float2(so.pdf_fwd, so.pdf_rev);
        }
        if (so.alive) store_path(c, path, ps);
    }
    const int li = warp_alloc(c.n_live_next, so.alive);
    if (so.alive) c.live_next[li] = path;
    if (so.emitter_hit && valid_color(so.emitted)) {                // depth-0 emitter hit, cpu_bdpt.cpp:372-375
        atomicAdd(&c.film[3 * ps.pixel], so.emitted.x); atomicAdd(&c.film[3 * ps.pixel + 1], so.emitted.y); atomicAdd(&c.film[3 * ps.pixel + 2], so.emitted.z);
    }
    warp_count_add(&c.counters[B2_CNT_PATH_SAMPLES], so.emitter_hit ? 1u : 0u);
    warp_count_add(&c.counters[B2_CNT_DELTA_CAP], so.delta_cap ? 1u : 0u);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}

// Connection stage.  Work item = (group of 32 eye-vertex records) x (slice of the active pool, whole chunks);
// persistent warps claim items from a global cursor.  The pool is sliced only when there are too few eye vertices to
// fill the machine (late waves), so that the serial latency of a nearly empty wave stays short.  No block-level
// synchronisation.  Per item a warp runs ONE loop that picks, warp-uniformly, the next of four phases; three per-warp
// queues in shared memory decouple them, so that each phase runs at (nearly) full lane occupancy:
//   weight   : >= 32 visible pairs queued: MIS weight, clamp, per-owner sums in shared memory (connect_weight)
//   prepare  : fewer than 32 READY rays and >= 32 pairs stacked (or the pool slice is exhausted): 32 pairs at a time run
//              connect_prepare (both BSDFs, black skip, shadow ray, contribution without the MIS weight); survivors are
//              appended to the ready queue (ray + contribution, 44 B each)
//   pre-test : fewer than 32 pairs stacked and pool chunks remain: every lane tests ITS eye vertex against CONN_CHUNK pool
//              records (warp-uniform addresses, fully unrolled: 2 broadcast loads + ~16 FP instructions per pair), one
//              bit per record; survivors go on the PAIR stack (owner lane, pool index), pool vertex by pool vertex
//   trace    : every idle lane first takes a ready ray (7 shared-memory loads), then one any-hit step (wide node + its
//              leaf triangles) for every lane holding a ray; an occluded ray frees its lane, a visible one is queued
//              for the weight phase
// so a lane whose ray is decided gets the next ray in the very next step while the others keep walking (occluded rays
// end after a few steps, visible ones walk the whole segment).  Round 1 refilled lanes by running connect_prepare on the
// idle lanes only once >= 20 of them were idle: ncu on the 1 M-triangle scene showed 21 of 32 lanes active in the node
// tests and 20 in the BSDFs (profiles/r02_connect_summary.md).  One film atomicAdd per eye vertex and channel per item.
constexpr int CONN_LOW = 32;          // pre-test another chunk while fewer pairs than this are stacked
constexpr int CONN_MQ = 64;           // visible-pair queue: < 32 left + 32 new per trace step
// ready-ray queue: < low left + 32 new per prepare round.  POOL (large scenes): a trace step frees ~4 lanes, 16 ready rays are
// plenty and the shared memory goes to the traversal stacks; small scenes: every lane finishes its ray within the step (below),
// so a step wants 32 ready rays
#ifndef B2_CONN_RQ_POOL
#define B2_CONN_RQ_POOL 40
#endif
template <bool POOL> struct ConnCfg { static constexpr int RQ = POOL ? B2_CONN_RQ_POOL : 64, RQ_LOW = POOL ? B2_CONN_RQ_POOL - 32 : 32; };
#ifndef B2_CONN_SORT_CV
#define B2_CONN_SORT_CV 2             // 1: eye vertices of the waves after the primary one sorted by position before k_connect; 2: every wave
#endif
#ifndef B2_CONN_SORT_MIN_RAYS
#define B2_CONN_SORT_MIN_RAYS 16      // ... when the first connection launch sent at least this many shadow rays per eye vertex
#endif
#ifndef B2_CONN_SORT_MIN
#define B2_CONN_SORT_MIN (1 << 16)    // ... for waves of at least this many live paths
#endif
#ifndef B2_FAST_SHADOW_TRI
#define B2_FAST_SHADOW_TRI 1          // contracted Moeller-Trumbore for the shadow rays of the connection stage (b2_isect.cuh)
#endif
#ifndef B2_DEFAULT_BATCH_LOG2
#define B2_DEFAULT_BATCH_LOG2 26      // most eye samples in flight per batch when the caller leaves max_batch_paths at 0
#endif
#ifndef B2_CONN_POOL_MIN_TRIS
#define B2_CONN_POOL_MIN_TRIS 2048    // scenes with at least this many triangles run the trace phase with the leaf triangles of the
#endif                                // whole warp pooled (C3 spl 8: 882 -> 820 ms); below it the per-lane loop of any_step wins
                                      // (C2, 48 triangles, 1.5 triangle tests per ray: connect stage 2.94 ms vs 3.51 ms pooled)
template <bool POOL>
struct ConnSmem {
    static constexpr int CONN_RQ = ConnCfg<POOL>::RQ;
    uint32_t pair[CONN_LOW + 32 * CONN_CHUNK];
    float rq_ox[CONN_RQ], rq_oy[CONN_RQ], rq_oz[CONN_RQ], rq_dx[CONN_RQ], rq_dy[CONN_RQ], rq_dz[CONN_RQ], rq_tmax[CONN_RQ];
    float rq_ix[CONN_RQ], rq_iy[CONN_RQ], rq_iz[CONN_RQ];      // 1 / direction, computed by the prepare phase at full lane occupancy
    float cur_ix[32], cur_iy[32], cur_iz[32];                  // ... of the ray each lane is tracing (read by the lane that finds a pooled triangle hit)
    float rq_px[CONN_RQ], rq_py[CONN_RQ], rq_pz[CONN_RQ]; uint32_t rq_e[CONN_RQ];
    float mq_px[CONN_MQ], mq_py[CONN_MQ], mq_pz[CONN_MQ]; uint32_t mq_e[CONN_MQ];
    float acc[3][32];
    float eye_sum[32];                                         // mis_eye_sum of the item's 32 eye vertices (B2_MIS_WALKS_HOISTED)
};
// position of the (j+1)-th set bit of a non-empty mask with more than j bits (j is small: a lane has 0.5 candidate
// triangles per step on average; the generic __fns costs ~25 instructions whatever j is)
__device__ __forceinline__ int nth_bit(uint32_t bits, int j) {
    for (; j > 0; --j) bits &= bits - 1u;
    return __ffs((int) bits) - 1;
}
template <bool COUNT, bool POOL>
__global__ void __launch_bounds__(CONN_THREADS, CONN_BLOCKS_PER_SM) k_connect(RenderCtx c) {
    extern __shared__ uint2 s_stack[];                    // CONN_THREADS x stack depth
    __shared__ ConnSmem<POOL> s_q[CONN_WARPS];
    constexpr int CONN_RQ_LOW = ConnCfg<POOL>::RQ_LOW;
    const int n_cv = *c.n_cv, n_act = *c.n_pool_active;
    if (n_cv <= 0 || n_act <= 0) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t lt = lanemask_lt();
    const int n_groups = (n_cv + 31) >> 5;
    const int n_chunks = (n_act + CONN_CHUNK - 1) / CONN_CHUNK;
    const int total_warps = gridDim.x * CONN_WARPS;
    const int n_slices = max(1, min(total_warps / n_groups, n_chunks));
    const long long n_items = (long long) n_groups * n_slices;
    ConnSmem<POOL> &Q = s_q[warp];
    uint2 *stack = s_stack + threadIdx.x;
    const float4 *__restrict__ pre = c.pre;
    uint32_t n_shadow = 0, n_path = 0;
    unsigned long long n_pairs = 0;
    TravCounters tc{ 0u, 0u };

    // Items are claimed from a global cursor, not grid-strided: the work of an item varies with the number of surviving
    // pairs, and a fixed stride aliases with the image width (2368 warps = 74 rows of 32 groups at 1024 px: every warp
    // would only ever see one image column; ncu showed max/avg = 1.33 instructions per scheduler).
    for (;;) {
        unsigned int item = 0u;
        if (lane == 0) item = atomicAdd(c.conn_cursor, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if ((long long) item >= n_items) break;
        int group, slice;
        if (n_slices == 1) { group = (int) item; slice = 0; }
        else { group = (int) (item / n_slices); slice = (int) (item % n_slices); }
        const int ch_begin = (int) ((long long) slice * n_chunks / n_slices), ch_end = (int) ((long long) (slice + 1) * n_chunks / n_slices);
        const int q0 = group << 5, q = q0 + lane;
        const bool valid = q < n_cv;
        float3 my_pos = mk3(0, 0, 0), my_n = mk3(0, 0, 0);            // a zero normal never passes: idle lanes need no branch
        uint32_t my_pixel = 0u;
        if (valid) { const float4 f0 = c.cv[B2_CV_F4 * q]; my_pos = mk3(f0); my_pixel = __float_as_uint(f0.w) >> 8; my_n = mk3(c.cv[B2_CV_F4 * q + 1]); }
        Q.acc[0][lane] = 0.0f; Q.acc[1][lane] = 0.0f; Q.acc[2][lane] = 0.0f;
#if B2_MIS_WALKS_HOISTED
        {   // the eye-side walk of the MIS weight, once per eye vertex: the weight phase below multiplies it by the pair's pdf
            float es = 0.0f;
            if (valid) { const EyeVtx ev = load_cv(c.cv, q); es = mis_eye_sum(RenderEyeAcc{ c, ev }, ev.depth); }
            Q.eye_sum[lane] = isfinite(es) ? es : 3.0e38f;     // an overflowing walk: the weight is zero either way (sum = inf)
        }
#endif
        __syncwarp();
        int n_buf = 0, n_rq = 0, n_mq = 0, ch = ch_begin;  // warp-uniform
        bool busy = false;                                 // this lane holds a shadow ray: tv, pc, e
        AnyTrav tv; float3 pc = mk3(0, 0, 0); uint32_t e = 0u;
        tv.r.o = tv.r.d = mk3(0, 0, 0); tv.tmax = 0.0f;

        for (;;) {
            const uint32_t busy_m = __ballot_sync(0xffffffffu, busy);
            const bool pairs_done = ch >= ch_end;
            if (n_mq >= 32 || (n_mq > 0 && busy_m == 0u && n_rq == 0 && n_buf == 0 && pairs_done)) {
                // ---- weight: top min(32, n_mq) visible pairs ----
                const int count = min(32, n_mq);
                if (lane < count) {
                    const int k = n_mq - count + lane;
                    const uint32_t ee = Q.mq_e[k];
                    const int owner = (int) (ee >> 27);
                    const EyeVtx ev = load_cv(c.cv, q0 + owner);
#if B2_MIS_WALKS_HOISTED
                    const float3 r = connect_weight(c, ev, (int) (ee & 0x7ffffffu), mk3(Q.mq_px[k], Q.mq_py[k], Q.mq_pz[k]), Q.eye_sum[owner], true);
#else
                    const float3 r = connect_weight(c, ev, (int) (ee & 0x7ffffffu), mk3(Q.mq_px[k], Q.mq_py[k], Q.mq_pz[k]));
#endif
                    n_path++;
                    if (r.x != 0.0f) atomicAdd(&Q.acc[0][owner], r.x);
                    if (r.y != 0.0f) atomicAdd(&Q.acc[1][owner], r.y);
                    if (r.z != 0.0f) atomicAdd(&Q.acc[2][owner], r.z);
                }
                n_mq -= count;
                __syncwarp();
            } else if (n_rq < CONN_RQ_LOW && (n_buf >= 32 || (n_buf > 0 && pairs_done))) {
                // ---- prepare: the top min(32, n_buf) pairs become ready shadow rays ----
                const int count = min(32, n_buf);
                bool ok = false;
                ShadowRay sr; float3 ppc = mk3(0, 0, 0); uint32_t pe = 0u;
                sr.o = sr.d = mk3(0, 0, 0); sr.tmax = 0.0f;
                if (lane < count) {
                    pe = Q.pair[n_buf - 1 - lane];
                    const EyeVtx ev = load_cv(c.cv, q0 + (int) (pe >> 27));
                    ok = connect_prepare(c, ev, (int) (pe & 0x7ffffffu), sr, ppc);
                }
                n_buf -= count;
                const uint32_t okm = __ballot_sync(0xffffffffu, ok);
                if (ok) {
                    const int k = n_rq + __popc(okm & lt);
                    Q.rq_ox[k] = sr.o.x; Q.rq_oy[k] = sr.o.y; Q.rq_oz[k] = sr.o.z; Q.rq_dx[k] = sr.d.x; Q.rq_dy[k] = sr.d.y; Q.rq_dz[k] = sr.d.z;
                    Q.rq_ix[k] = safe_rcp(sr.d.x); Q.rq_iy[k] = safe_rcp(sr.d.y); Q.rq_iz[k] = safe_rcp(sr.d.z);
                    Q.rq_tmax[k] = sr.tmax; Q.rq_px[k] = ppc.x; Q.rq_py[k] = ppc.y; Q.rq_pz[k] = ppc.z; Q.rq_e[k] = pe;
                    n_shadow++;
                }
                n_rq += __popc(okm);
                __syncwarp();
            } else if (n_rq < CONN_RQ_LOW && n_buf < CONN_LOW && !pairs_done) {
                // ---- pre-test chunks of pool records until 32 pairs are stacked or the slice ends (one pass through the phase
                //      selector per CHUNK cost ~30 of ~190 instructions on scenes where almost every pair is rejected: C2) ----
              do {
                const float4 *__restrict__ rec = pre + (size_t) B2_PRE_F4 * CONN_CHUNK * ch;
                ++ch;
                uint32_t mask = 0u;
#pragma unroll
                for (int k = 0; k < CONN_CHUNK; ++k) {
                    const float4 r0 = __ldg(&rec[B2_PRE_F4 * k]), r1 = __ldg(&rec[B2_PRE_F4 * k + 1]);
                    const float3 d = mk3(r0) - my_pos;
                    const float dist2 = dot3(d, d);
                    bool pass = !(dist2 < 1e-6f) && dot3(my_n, d) > 0.0f && dot3(mk3(r1), d) < 0.0f;
                    if (r0.w > -1.5f && pass) {            // spot-cone gate (warp-uniform first operand)
                        const float4 r2 = __ldg(&rec[B2_PRE_F4 * k + 2]);
                        pass = !(-dot3(mk3(r2), d) < r0.w * fast_sqrt(dist2));
                    }
                    mask |= (pass ? 1u : 0u) << k;
                }
                if (__ballot_sync(0xffffffffu, mask != 0u)) {
                    // pool vertex by pool vertex, owners in lane order: consecutive entries are shadow rays from
                    // neighbouring eye vertices to the same pool vertex, which keeps the traversal of a warp coherent
#pragma unroll
                    for (int k = 0; k < CONN_CHUNK; ++k) {
                        const bool mine = (mask >> k) & 1u;
                        const uint32_t m = __ballot_sync(0xffffffffu, mine);
                        if (mine) Q.pair[n_buf + __popc(m & lt)] = ((uint32_t) lane << 27) | __float_as_uint(__ldg(&rec[B2_PRE_F4 * k + 1].w));
                        n_buf += __popc(m);
                    }
                    __syncwarp();
                }
              } while (n_buf < CONN_LOW && ch < ch_end);
            } else if (busy_m != 0u || n_rq > 0) {
                // ---- trace: idle lanes take a ready ray, then one step per lane that holds a ray ----
                if (n_rq > 0 && busy_m != 0xffffffffu) {
                    const int rank = __popc(~busy_m & lt);
                    if (!busy && rank < n_rq) {
                        const int k = n_rq - 1 - rank;
                        e = Q.rq_e[k]; pc = mk3(Q.rq_px[k], Q.rq_py[k], Q.rq_pz[k]);
                        const float3 idir = mk3(Q.rq_ix[k], Q.rq_iy[k], Q.rq_iz[k]);
                        any_init(tv, c.sc, ray_pre(mk3(Q.rq_ox[k], Q.rq_oy[k], Q.rq_oz[k]), mk3(Q.rq_dx[k], Q.rq_dy[k], Q.rq_dz[k]), idir), Q.rq_tmax[k]);
                        if (POOL) { Q.cur_ix[lane] = idir.x; Q.cur_iy[lane] = idir.y; Q.cur_iz[lane] = idir.z; }
                        busy = true;
                    }
                    n_rq -= min(32 - __popc(busy_m), n_rq);
                    __syncwarp();
                }
                bool visible = false;
              if (POOL) {
                // node tests per lane; the leaf triangles of the whole warp are numbered with a prefix sum and tested 32 at a
                // time by whichever lanes are free (per-lane triangle loops ran at ~4 active lanes, profiles/r01_v11_summary.md)
                uint32_t tribits = 0u, tri_base = 0u, vmask = 0u;
                bool exhausted = false;
                if (busy) exhausted = !any_node_step<COUNT>(c.sc, tv, stack, CONN_THREADS, &tc, tribits, tri_base, vmask);
                // (an L1 prefetch of the node the NEXT step will test - trav_peek + prefetch.global.L1 here, before the pooled triangle
                //  tests - was measured: C3 1110 -> 1756 ms; 33 KB of queues and stacks per block leave ~50 KB of L1 per SM, the
                //  prefetched lines evict the triangle rows and each other.  profiles/r02_connect_summary.md)
                const int cnt = __popc(tribits);
                int incl = cnt;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += v; }
                const int total = __shfl_sync(0xffffffffu, incl, 31);
                uint32_t blocked = 0u;                          // bit j: the ray of lane j hit an opaque triangle
                for (int base = 0; base < total; base += 32) {
                    const int k = base + lane;
                    int owner = 0;
#pragma unroll
                    for (int step = 16; step >= 1; step >>= 1) { const int v = __shfl_sync(0xffffffffu, incl, owner + step - 1); if (v <= k) owner += step; }
                    owner = min(owner, 31);
                    const int o_excl = __shfl_sync(0xffffffffu, incl - cnt, owner);
                    const uint32_t o_bits = __shfl_sync(0xffffffffu, tribits, owner), o_base = __shfl_sync(0xffffffffu, tri_base, owner);
                    const uint32_t o_vmask = __shfl_sync(0xffffffffu, vmask, owner);
                    // the three rows of the triangle are requested together, BEFORE the owner's ray is fetched: round 2's first
                    // version read row b, tested its transparency flag and only then asked for rows a and c - two dependent
                    // L2 round trips, 15 % of the kernel's stall samples (profiles/r02_connect_summary.md)
                    float4 a = mk4(mk3(0, 0, 0), 0.0f), b = a, cc = a;
                    if (k < total) {
                        const int ti = tri_of_bit(o_base, o_vmask, nth_bit(o_bits, k - o_excl));
                        a = c.sc.tri[3 * ti]; b = c.sc.tri[3 * ti + 1]; cc = c.sc.tri[3 * ti + 2];   // (plain loads: L2-only (__ldcg) 1042 -> 1078 ms, L1::evict_first 1003 -> 1030 ms)
                    }
                    const float3 ro = mk3(__shfl_sync(0xffffffffu, tv.r.o.x, owner), __shfl_sync(0xffffffffu, tv.r.o.y, owner), __shfl_sync(0xffffffffu, tv.r.o.z, owner));
                    const float3 rd = mk3(__shfl_sync(0xffffffffu, tv.r.d.x, owner), __shfl_sync(0xffffffffu, tv.r.d.y, owner), __shfl_sync(0xffffffffu, tv.r.d.z, owner));
                    const float o_tmax = __shfl_sync(0xffffffffu, tv.tmax, owner);
                    bool hit = false;
                    if (k < total && !(f2u(b.w) & B2_MAT_TRANSPARENT)) {
                        if (COUNT) tc.tris++;
                        float tt;
                        if (B2_FAST_SHADOW_TRI != 0) {
                            if (tri_test_fast(ro, rd, mk3(a), mk3(b), mk3(cc), 1e-3f, o_tmax)) {
                                const RayPre rp = ray_pre(ro, rd, mk3(Q.cur_ix[owner], Q.cur_iy[owner], Q.cur_iz[owner]));
                                hit = group_pass_fast(c.sc, (int) f2u(cc.w), rp, 1e-3f, o_tmax);
                            }
                        } else hit = tri_test(ro, rd, mk3(a), mk3(b), mk3(cc), 1e-3f, o_tmax, tt) && group_pass(c.sc, (int) f2u(cc.w), ro, rd, 1e-3f, o_tmax);
                    }
                    // every lane learns which owners were hit: OR of (1 << owner) over the lanes that hit
                    blocked |= __reduce_or_sync(0xffffffffu, hit ? 1u << owner : 0u);
                }
                if (busy) {
                    if ((blocked >> lane) & 1u) busy = false;
                    else if (exhausted) {
                        if (!(B2_FAST_SHADOW_TRI != 0 ? occluded_spheres_fast(c.sc, tv.r, 1e-3f, tv.tmax) : occluded_spheres(c.sc, tv.r.o, tv.r.d, 1e-3f, tv.tmax))) visible = true;
                        busy = false;
                    }
                }
              } else if (busy) {
                    // small scenes (a ray visits 1-3 wide nodes): every lane walks its ray to the end, so the next step starts with
                    // 32 free lanes and 32 ready rays.  One step per pass through the phase selector ran the node tests at 14 of 32
                    // lanes and the sphere tests at 12 on C1 (profiles/r02_small_scenes.md)
                    int r;
                    do r = any_step<COUNT, B2_FAST_SHADOW_TRI != 0>(c.sc, tv, stack, CONN_THREADS, &tc); while (r == B2_TRAV_CONTINUE);
                    if (r == B2_TRAV_EXHAUSTED) visible = !(B2_FAST_SHADOW_TRI != 0 ? occluded_spheres_fast(c.sc, tv.r, 1e-3f, tv.tmax) : occluded_spheres(c.sc, tv.r.o, tv.r.d, 1e-3f, tv.tmax));
                    busy = false;
              }
                const uint32_t vm = __ballot_sync(0xffffffffu, visible);
                if (vm) {
                    if (visible) {
                        const int k = n_mq + __popc(vm & lt);
                        Q.mq_px[k] = pc.x; Q.mq_py[k] = pc.y; Q.mq_pz[k] = pc.z; Q.mq_e[k] = e;
                    }
                    n_mq += __popc(vm);
                    __syncwarp();
                }
            } else break;                                  // nothing stacked, nothing ready, nothing in flight, no chunk left, queue empty
        }
        if (valid) {
            const float3 r = mk3(Q.acc[0][lane], Q.acc[1][lane], Q.acc[2][lane]);
            if (r.x != 0.0f) atomicAdd(&c.film[3 * my_pixel], r.x);
            if (r.y != 0.0f) atomicAdd(&c.film[3 * my_pixel + 1], r.y);
            if (r.z != 0.0f) atomicAdd(&c.film[3 * my_pixel + 2], r.z);
        }
        n_pairs += (unsigned long long) (min(ch_end * CONN_CHUNK, n_act) - ch_begin * CONN_CHUNK) * (unsigned long long) min(32, n_cv - q0);
        __syncwarp();
    }
    warp_count_add(&c.counters[B2_CNT_SHADOW], n_shadow);
    warp_count_add(&c.counters[B2_CNT_PATH_SAMPLES], n_path);
    if (lane == 0 && n_pairs) atomicAdd(&c.counters[B2_CNT_PAIRS], n_pairs);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}

// ---- strategies beyond the reference set (b2_strategies.cuh): t = 1 light tracing with film splats, s = 1 next-event ----
template <bool COUNT>
__global__ void __launch_bounds__(EXT_THREADS) k_light_trace(RenderCtx c, CamFrame cf, int n_items, uint32_t pix0, int s0, float alpha0_scale) {
    extern __shared__ uint2 s_stack[];
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    TravCounters tc{ 0u, 0u };
    uint32_t ncl = 0, nsh = 0, nsp = 0;
    if (q < n_items) {
        uint32_t slot;                                      // same (sample, slot) split as the eye samples, so shards partition the subpaths
        const uint32_t ds = udiv_magic(pix0 + (uint32_t) q, (uint32_t) (c.W * c.H), c.npix_magic, slot);
        const unsigned long long sample = (unsigned long long) (c.sample_offset + (s0 + (int) ds) * c.sample_stride);
        light_trace_body<COUNT>(c, cf, sample * (unsigned long long) (c.W * c.H) + slot, alpha0_scale, s_stack + threadIdx.x, blockDim.x, &tc, ncl, nsh, nsp);
    }
    warp_count_add(&c.counters[B2_CNT_CLOSEST], ncl);
    warp_count_add(&c.counters[B2_CNT_SHADOW], nsh);
    warp_count_add(&c.counters[B2_CNT_PATH_SAMPLES], nsp);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}
template <bool COUNT>
__global__ void __launch_bounds__(EXT_THREADS) k_nee(RenderCtx c, float w) {
    extern __shared__ uint2 s_stack[];
    const int n_cv = *c.n_cv;
    TravCounters tc{ 0u, 0u };
    uint32_t nsh = 0, npa = 0;
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < n_cv; q += gridDim.x * blockDim.x) {
        const EyeVtx ev = load_cv(c.cv, q);
        const float3 r = nee_body<COUNT>(c, ev, w, s_stack + threadIdx.x, blockDim.x, &tc, nsh, npa);
        if (r.x != 0.0f) atomicAdd(&c.film[3 * ev.pixel], r.x);
        if (r.y != 0.0f) atomicAdd(&c.film[3 * ev.pixel + 1], r.y);
        if (r.z != 0.0f) atomicAdd(&c.film[3 * ev.pixel + 2], r.z);
    }
    warp_count_add(&c.counters[B2_CNT_SHADOW], nsh);
    warp_count_add(&c.counters[B2_CNT_PATH_SAMPLES], npa);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}

__global__ void __launch_bounds__(256) k_resolve(const float *__restrict__ sum, float *__restrict__ out, size_t n, float inv) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = sum[i] * inv;
}
__global__ void __launch_bounds__(256) k_resolve_acc(const float *__restrict__ sum, float *__restrict__ out, size_t n, float spp) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] += sum[i] / spp;                  // one progressive pass into the running sum (main.cpp:443-446)
}
__global__ void __launch_bounds__(256) k_resolve_div(const float *__restrict__ sum, float *__restrict__ out, size_t n, float spp) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = sum[i] / spp;                   // accum / spp, cpu_bdpt.cpp:479
}

__global__ void k_pool_export(RenderCtx c, float *out20) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= c.n_pool) return;
    const LightVtx v = load_lv(c.pool, i);
    float *o = out20 + 20 * (size_t) i;
    o[0] = v.pos.x; o[1] = v.pos.y; o[2] = v.pos.z; o[3] = v.n.x; o[4] = v.n.y; o[5] = v.n.z; o[6] = v.thr.x; o[7] = v.thr.y; o[8] = v.thr.z;
    o[9] = v.mtl.base.x; o[10] = v.mtl.base.y; o[11] = v.mtl.base.z; o[12] = v.mtl.roughness; o[13] = v.mtl.metallic; o[14] = v.mtl.eta;
    o[15] = (v.flags & B2_LV_SRC) ? 1.f : 0.f; o[16] = (v.flags & B2_LV_PARALLEL) ? 1.f : 0.f; o[17] = v.cutoff; o[18] = v.pdf_fwd; o[19] = v.pdf_rev;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------
// RenderWorkspace: device buffers that persist across calls on a scene (re-allocated only when they grow)
// ---------------------------------------------------------------------------------------------------------
int RenderWorkspace::ensure(int max_paths, int eye_depth, int n_pool, size_t film_floats) {
    if (max_paths > cap_paths || eye_depth > cap_depth) {
        const int mp = std::max(max_paths, cap_paths), ed = std::max(eye_depth, cap_depth);
        B2_TRY(ray_o.alloc(mp)); B2_TRY(ray_d.alloc(mp)); B2_TRY(thr.alloc(mp)); B2_TRY(last_pos.alloc(mp)); B2_TRY(last_n.alloc(mp));
        B2_TRY(hist_pe.alloc((size_t) ed * mp)); B2_TRY(hist_pdf.alloc((size_t) ed * mp));
        B2_TRY(live_a.alloc(mp)); B2_TRY(live_b.alloc(mp)); B2_TRY(cv.alloc((size_t) B2_CV_F4 * mp));
        cap_paths = mp; cap_depth = ed;
    }
    if (n_pool > cap_pool) { B2_TRY(pool.alloc((size_t) B2_LV_F4 * n_pool)); B2_TRY(pre.alloc((size_t) B2_PRE_F4 * (n_pool + CONN_CHUNK))); cap_pool = n_pool; }
    if (film_floats > film.n) B2_TRY(film.alloc(film_floats));
    if (!ints.p) B2_TRY(ints.alloc(8));
    if (!counters.p) B2_TRY(counters.alloc(B2_CNT_N));
    if (!h_ints) B2_CUDA(cudaMallocHost((void **) &h_ints, 8 * sizeof(int)));
    return B2_OK;
}
int RenderWorkspace::ensure_cv_sort(int max_paths) {
    if ((size_t) B2_CV_F4 * max_paths > cv_sorted.n) B2_TRY(cv_sorted.alloc((size_t) B2_CV_F4 * max_paths));
    return b2_sort_reserve(cv_sort, (size_t) max_paths);
}
int RenderWorkspace::ensure_out(size_t film_floats) {
    if (film_floats > film_out.n) B2_TRY(film_out.alloc(film_floats));
    if (film_floats > h_film_n) {
        if (h_film) cudaFreeHost(h_film);
        h_film = nullptr; h_film_n = 0;
        B2_CUDA(cudaMallocHost((void **) &h_film, film_floats * sizeof(float)));
        h_film_n = film_floats;
    }
    return B2_OK;
}
void RenderWorkspace::release() {
    film_out.release();
    if (h_film) { cudaFreeHost(h_film); h_film = nullptr; h_film_n = 0; }
    ray_o.release(); ray_d.release(); thr.release(); last_pos.release(); last_n.release();
    hist_pe.release(); hist_pdf.release(); live_a.release(); live_b.release(); cv.release();
    pool.release(); pre.release(); film.release(); ints.release(); counters.release();
    cv_sorted.release(); cv_sort.keys.release(); cv_sort.keys_sorted.release(); cv_sort.idx.release(); cv_sort.idx_sorted.release(); cv_sort.tmp.release();
    hp_in.release(); hp.release(); flux.release(); hp_keys.release(); hp_keys_sorted.release(); hp_order.release(); hp_order_sorted.release();
    cell_table.release(); sort_tmp.release();
    mw_f4.release(); mw_i32.release(); mw_pairs.release();
    if (h_ints) { cudaFreeHost(h_ints); h_ints = nullptr; }
    cap_paths = cap_depth = cap_pool = 0;
}

static int validate_params(const B2RenderParams &p) {
    if (p.width <= 0 || p.height <= 0 || p.spp <= 0 || p.spl < 0 || p.eye_depth < 0 || p.light_depth <= 0) {
        b2_set_error("invalid render parameters (width/height/spp must be > 0, light_depth > 0, spl/eye_depth >= 0)"); return B2_ERR_INVALID;
    }
    if (p.eye_depth > 255) { b2_set_error("eye_depth > 255 is not supported"); return B2_ERR_INVALID; }
    if ((long long) p.width * p.height > (1ll << 24)) { b2_set_error("film larger than 2^24 pixels is not supported"); return B2_ERR_INVALID; }
    if (p.sample_offset < 0 || p.sample_stride < 0) { b2_set_error("invalid sample_offset / sample_stride"); return B2_ERR_INVALID; }
    if (p.strategies < B2_STRATEGIES_REFERENCE || p.strategies > B2_STRATEGIES_MIS) { b2_set_error("unknown strategies value %d", p.strategies); return B2_ERR_INVALID; }
    if (p.integrator < B2_INTEGRATOR_BDPT || p.integrator > B2_INTEGRATOR_PPM) { b2_set_error("unknown integrator value %d", p.integrator); return B2_ERR_INVALID; }
    if (p.integrator != B2_INTEGRATOR_BDPT && p.strategies != B2_STRATEGIES_REFERENCE) { b2_set_error("strategies apply to the bdpt integrator only"); return B2_ERR_INVALID; }
    if (!(p.ppm_radius >= 0.0f)) { b2_set_error("invalid ppm_radius"); return B2_ERR_INVALID; }
    if ((long long) p.spl * p.light_depth > (1ll << 26)) { b2_set_error("spl * light_depth too large"); return B2_ERR_INVALID; }
    return B2_OK;
}

void b2_setup_ctx(RenderCtx &c, const DevScene &sc, const B2Camera &cam, const B2RenderParams &p, RenderWorkspace &ws, int max_paths) {
    memset(&c, 0, sizeof(c));
    c.sc = sc;
    c.sc.ggx_fix = p.strategies != B2_STRATEGIES_REFERENCE;      // normalised GGX D only outside parity mode (b2_bsdf.cuh)
    camera_basis(cam.eye, cam.look_at, cam.view_up, cam.fov, p.width, p.height, c.cam_eye, c.cUL, c.cdx, c.cdy);
    c.W = p.width; c.H = p.height; c.spl = p.spl; c.eye_depth = p.eye_depth; c.light_depth = p.light_depth;
    c.w_magic = udiv_magic_of((uint32_t) p.width); c.npix_magic = udiv_magic_of((uint32_t) p.width * (uint32_t) p.height);
    c.max_delta = p.max_delta > 0 ? p.max_delta : 256;
    c.sample_offset = p.sample_offset; c.sample_stride = p.sample_stride > 0 ? p.sample_stride : 1;
    c.key = make_uint2((uint32_t) p.seed, (uint32_t) (p.seed >> 32));
    c.pool = ws.pool.p; c.n_pool = sc.n_lights * p.spl * p.light_depth;
    c.pre = ws.pre.p; c.n_pool_active = ws.ints.p + 0;
    c.max_paths = max_paths;
    c.ray_o = ws.ray_o.p; c.ray_d = ws.ray_d.p; c.thr = ws.thr.p; c.last_pos = ws.last_pos.p; c.last_n = ws.last_n.p;
    c.hist_pe = ws.hist_pe.p; c.hist_pdf = ws.hist_pdf.p;
    c.live = ws.live_a.p; c.live_next = ws.live_b.p; c.n_live_next = ws.ints.p + 1;
    c.n_cv = ws.ints.p + 2; c.conn_cursor = (unsigned int *) (ws.ints.p + 3); c.n_paths = ws.ints.p + 4;
    c.cv = ws.cv.p; c.film = ws.film.p; c.counters = ws.counters.p;
}

// Optional per-stage device timing: one CUDA event pair per launch, summed per stage after each wave's sync.
struct StageTimer {
    enum { PRIMARY, EXTEND, CONNECT, N };
    bool on = false;
    cudaStream_t st = nullptr;
    std::vector<cudaEvent_t> pool;
    struct Rec { int stage; cudaEvent_t a, b; };
    std::vector<Rec> pending;
    float ms[N] = { 0, 0, 0 };
    cudaEvent_t get() { if (pool.empty()) { cudaEvent_t e; cudaEventCreate(&e); return e; } cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
    void begin(int stage) { if (!on) return; Rec r{ stage, get(), get() }; cudaEventRecord(r.a, st); pending.push_back(r); }
    void end() { if (!on) return; cudaEventRecord(pending.back().b, st); }
    void collect() {             // call after a stream synchronize
        for (Rec &r : pending) { float t = 0.f; if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess) ms[r.stage] += t; pool.push_back(r.a); pool.push_back(r.b); }
        pending.clear();
    }
    ~StageTimer() { for (cudaEvent_t e : pool) cudaEventDestroy(e); for (Rec &r : pending) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); } }
};


// Light pool generation (shared by render and b2_light_pool).  Leaves *n_pool_active in ws.ints[0].
static int build_pool(RenderCtx &c, int bvh_depth, bool count, cudaStream_t st, uint64_t &n_launches) {
    if (c.n_pool <= 0) { B2_CUDA(cudaMemsetAsync(c.n_pool_active, 0, sizeof(int), st)); return B2_OK; }
    B2_CUDA(cudaMemsetAsync(c.pool, 0, sizeof(float4) * B2_LV_F4 * (size_t) c.n_pool, st));    // value-initialised pool, cpu_bdpt.cpp:203
    const int n_paths = c.sc.n_lights * c.spl;
    const size_t sb = stack_bytes(EXT_THREADS, bvh_depth);
    if (count) k_light_paths<true><<<blocks_for(n_paths, EXT_THREADS), EXT_THREADS, sb, st>>>(c, n_paths);
    else k_light_paths<false><<<blocks_for(n_paths, EXT_THREADS), EXT_THREADS, sb, st>>>(c, n_paths);
    k_pool_active<<<1, 1024, 0, st>>>(c);
    n_launches += 2;
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}

static int set_smem_limits(int bvh_depth) {
    static int configured[64];                 // per device: bvh_depth + 1 already configured (0 = never)
    int dev = 0;
    B2_CUDA(cudaGetDevice(&dev));
    if (dev < 64 && bvh_depth + 1 <= configured[dev]) return B2_OK;
    const size_t sb_t = stack_bytes(EXT_THREADS, bvh_depth), sb_c = stack_bytes(CONN_THREADS, bvh_depth);
    if (sb_c + std::max(sizeof(ConnSmem<true>), sizeof(ConnSmem<false>)) * CONN_WARPS + 64 > 227 * 1024) {
        b2_set_error("BVH depth %d needs more shared memory for traversal stacks than a CTA has", bvh_depth); return B2_ERR_STATE;
    }
    B2_CUDA(cudaFuncSetAttribute(k_light_paths<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_t));
    B2_CUDA(cudaFuncSetAttribute(k_light_paths<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_t));
    B2_CUDA(cudaFuncSetAttribute(k_extend<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_t));
    B2_CUDA(cudaFuncSetAttribute(k_extend<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_t));
    B2_CUDA(cudaFuncSetAttribute(k_extend<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_t));
    B2_CUDA(cudaFuncSetAttribute(k_extend<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_t));
    B2_CUDA(cudaFuncSetAttribute(k_connect<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_c));
    B2_CUDA(cudaFuncSetAttribute(k_connect<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_c));
    B2_CUDA(cudaFuncSetAttribute(k_connect<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_c));
    B2_CUDA(cudaFuncSetAttribute(k_connect<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_c));
    B2_CUDA(cudaFuncSetAttribute(k_light_trace<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_t));
    B2_CUDA(cudaFuncSetAttribute(k_light_trace<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_t));
    B2_CUDA(cudaFuncSetAttribute(k_nee<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_t));
    B2_CUDA(cudaFuncSetAttribute(k_nee<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_t));
    if (dev < 64) configured[dev] = bvh_depth + 1;
    return B2_OK;
}

int b2_render_device_impl(const DevScene &sc, int bvh_depth, RenderWorkspace &ws, const B2Camera &cam, const B2RenderParams &p,
                          float *d_film_out, cudaStream_t st, B2Stats *stats) {
    B2_TRY(validate_params(p));
    if (p.integrator == B2_INTEGRATOR_PT) return b2_render_pt_impl(sc, bvh_depth, ws, cam, p, d_film_out, st, stats);
    if (p.integrator == B2_INTEGRATOR_PPM) return b2_render_ppm_impl(sc, bvh_depth, ws, cam, p, d_film_out, st, stats);
    if (p.strategies == B2_STRATEGIES_MIS) return b2_render_mis_impl(sc, bvh_depth, ws, cam, p, d_film_out, st, stats);
    const int npix = p.width * p.height;
    const size_t film_floats = (size_t) 3 * npix;
    const int stride = p.sample_stride > 0 ? p.sample_stride : 1;
    int n_my = 0;
    for (long long s = p.sample_offset; s < p.spp; s += stride) n_my++;
    const long long total = (long long) npix * n_my;
    long long mp = p.max_batch_paths;
    if (mp <= 0) {
        // default: up to 2^26 samples in flight (C2 in one batch: the per-batch tails of the extend/connect waves cost ~0.35 ms
        // each, 8.2 -> 10.0 Gsamples/s from 2^24 to 2^26), halved until the wavefront state (104 + 24 B per depth and sample,
        // 24 GB at depth 8) fits in half of the free HBM
        const size_t per_path = 5 * sizeof(float4) + (size_t) std::max(p.eye_depth, 1) * (sizeof(float4) + sizeof(float2)) + 2 * sizeof(int) + B2_CV_F4 * sizeof(float4)
                                + (B2_CONN_SORT_CV ? B2_CV_F4 * sizeof(float4) + 24 : 0);      // sorted records, keys, order, radix scratch
        mp = std::min<long long>(1ll << B2_DEFAULT_BATCH_LOG2, std::max<long long>(total, 1));
        if (mp > ws.cap_paths || p.eye_depth > ws.cap_depth) {            // only when the workspace has to grow (cudaMemGetInfo costs > 1 ms)
            size_t free_b = 0, total_b = 0;
            B2_CUDA(cudaMemGetInfo(&free_b, &total_b));
            const size_t have = (size_t) std::max(ws.cap_paths, 0) * (5 * sizeof(float4) + (size_t) std::max(ws.cap_depth, 0) * (sizeof(float4) + sizeof(float2)) + 2 * sizeof(int) + B2_CV_F4 * sizeof(float4))
                                + ws.cv_sorted.n * sizeof(float4) + ws.cv_sort.keys.n * 16 + ws.cv_sort.tmp.n;
            while (mp > (1ll << 20) && (size_t) mp * per_path > (free_b + have) / 2) mp >>= 1;
            if (mp < ws.cap_paths && p.eye_depth <= ws.cap_depth) mp = ws.cap_paths;
        }
    }
    mp = std::max<long long>(std::min<long long>(mp, std::max<long long>(total, 1)), 1);
    const int max_paths = (int) mp;
    if ((long long) sc.n_lights * p.spl * p.light_depth >= (1ll << 27)) { b2_set_error("light pool larger than 2^27 vertices is not supported"); return B2_ERR_INVALID; }
    const int n_pool = sc.n_lights * p.spl * p.light_depth;
    B2_TRY(ws.ensure(max_paths, std::max(p.eye_depth, 1), std::max(n_pool, 1), film_floats));
    B2_TRY(set_smem_limits(bvh_depth));
    RenderCtx c; b2_setup_ctx(c, sc, cam, p, ws, max_paths);
    const bool count = p.count_traversal != 0;
    uint64_t n_launches = 0, n_batches = 0, n_waves = 0;
    const bool conn_pool = sc.n_tris >= B2_CONN_POOL_MIN_TRIS;
    // Eye vertices sorted by position ahead of k_connect: -1 = not decided yet, 0 = no, 1 = yes.  It pays when an eye vertex sends
    // enough shadow rays (C3: 52 per vertex, 936 -> 837 ms; C1: 74, 187 -> 148 ms; C2: ~2, 5.99 -> 9.7 ms), which only the first
    // connection launch can tell: that one runs unsorted and its shadow-ray counter decides for the rest of the render; the
    // workspace remembers the answer for the next call with the same sampling parameters.
    const unsigned long long sort_cfg = ((unsigned long long) p.spl << 48) ^ ((unsigned long long) p.light_depth << 40) ^ ((unsigned long long) p.eye_depth << 32)
                                        ^ ((unsigned long long) p.width << 16) ^ (unsigned long long) p.height ^ 0x8000000000000000ull;
    int sort_cv = (B2_CONN_SORT_CV != 0 && p.strategies == B2_STRATEGIES_REFERENCE && max_paths >= B2_CONN_SORT_MIN) ? -1 : 0;
    if (sort_cv != 0) {                                       // B2PT_CONN_SORT=0 / 1 in the environment: never / always (tests, A/B runs)
        const char *force = getenv("B2PT_CONN_SORT");
        if (force && (force[0] == '0' || force[0] == '1') && force[1] == 0) sort_cv = force[0] - '0';
    }
    if (sort_cv < 0 && ws.sort_cfg == sort_cfg) sort_cv = ws.sort_yes;
    // the reordering is an optimisation: without memory for the second record buffer the render goes on in caller order
    if (sort_cv > 0 && ws.ensure_cv_sort(max_paths) != B2_OK) { cudaGetLastError(); sort_cv = 0; }
    bool sort_probe = false; int sort_probe_vertices = 0;     // counter copy of the first launch queued / its eye-vertex count
    StageTimer tm; tm.on = p.profile_stages != 0; tm.st = st;
    CudaEvent ev0, ev1, ev2, ev_wave, ev_probe;
    B2_CUDA(ev0.create()); B2_CUDA(ev1.create()); B2_CUDA(ev2.create()); B2_CUDA(ev_wave.create(cudaEventDisableTiming)); B2_CUDA(ev_probe.create(cudaEventDisableTiming));
    B2_CUDA(cudaEventRecord(ev0, st));
    B2_CUDA(cudaMemsetAsync(ws.film.p, 0, sizeof(float) * film_floats, st));
    B2_CUDA(cudaMemsetAsync(ws.counters.p, 0, sizeof(unsigned long long) * B2_CNT_N, st));
    B2_CUDA(cudaMemsetAsync(ws.ints.p, 0, sizeof(int) * 8, st));
    unsigned long long eye_samples = 0, eye_vertices = 0;
    // run_cpu_bdpt returns before touching the film when there are no lights (cpu_bdpt.cpp:178); here the film is zero
    const bool ref_set = p.strategies == B2_STRATEGIES_REFERENCE;
    const bool eye_pass = p.strategies != B2_STRATEGIES_LIGHT_TRACING, lt_pass = p.strategies == B2_STRATEGIES_FULL || p.strategies == B2_STRATEGIES_LIGHT_TRACING;
    const float w_mix = p.strategies == B2_STRATEGIES_FULL ? 0.5f : 1.0f;    // constant-weight combination of s = 1 and t = 1
    if (sc.n_lights > 0 && total > 0) {
        if (ref_set) B2_TRY(build_pool(c, bvh_depth, count, st, n_launches));
        B2_CUDA(cudaEventRecord(ev1, st));

        const size_t sb_t = stack_bytes(EXT_THREADS, bvh_depth), sb_c = stack_bytes(CONN_THREADS, bvh_depth);
        if (lt_pass) {                                       // t = 1: W*H*n_my light subpaths, splatted straight into the film sums
            const CamFrame cf = cam_frame(c.cdx, c.cdy);
            const float a0 = (float) sc.n_lights / (float) npix * w_mix;
            for (long long w0 = 0; w0 < total; w0 += (1ll << 24)) {
                const int n = (int) std::min<long long>(1ll << 24, total - w0);
                const uint32_t pix0 = (uint32_t) (w0 % npix); const int s0 = (int) (w0 / npix);
                tm.begin(StageTimer::CONNECT);
                if (count) k_light_trace<true><<<blocks_for(n, EXT_THREADS), EXT_THREADS, sb_t, st>>>(c, cf, n, pix0, s0, a0);
                else k_light_trace<false><<<blocks_for(n, EXT_THREADS), EXT_THREADS, sb_t, st>>>(c, cf, n, pix0, s0, a0);
                tm.end();
                n_launches++;
            }
            B2_CUDA(cudaGetLastError());
        }
        for (long long w0 = 0; eye_pass && w0 < total; w0 += max_paths) {
            int n_live = (int) std::min<long long>(max_paths, total - w0);
            c.live = ws.live_a.p; c.live_next = ws.live_b.p;
            n_batches++;
            eye_samples += (unsigned long long) n_live;
            B2_CUDA(cudaMemsetAsync(c.n_paths, 0, sizeof(int), st));
            bool primary = true;
            while (n_live > 0 && p.eye_depth > 0) {
                B2_CUDA(cudaMemsetAsync(c.n_live_next, 0, 3 * sizeof(int), st));      // n_live_next, n_cv, conn_cursor
                const unsigned gb = blocks_for(n_live, EXT_THREADS);
                tm.begin(primary ? StageTimer::PRIMARY : StageTimer::EXTEND);
                if (primary) {
                    const uint32_t pix0 = (uint32_t) (w0 % npix); const int s0 = (int) (w0 / npix);
                    if (count) k_extend<true, true><<<gb, EXT_THREADS, sb_t, st>>>(c, n_live, pix0, s0);
                    else k_extend<true, false><<<gb, EXT_THREADS, sb_t, st>>>(c, n_live, pix0, s0);
                } else {
                    if (count) k_extend<false, true><<<gb, EXT_THREADS, sb_t, st>>>(c, n_live, 0u, 0);
                    else k_extend<false, false><<<gb, EXT_THREADS, sb_t, st>>>(c, n_live, 0u, 0);
                }
                tm.end();
                // the size of the next wave is read back BEHIND the extend kernel and AHEAD of the connection stage: the host waits
                // on this event, not on the stream, so it learns n_live while k_connect still runs and queues the next extend
                // behind it -- no idle gap between waves (round 1 synchronised the stream here: 128 bubbles per C3 render)
                B2_CUDA(cudaMemcpyAsync(ws.h_ints, c.n_live_next, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
                B2_CUDA(cudaEventRecord(ev_wave, st));
                // persistent warps: one CTA wave over the SMs (work items claimed from a cursor on the device)
                const unsigned gc = (unsigned) std::min<long long>((long long) sc.n_sms * CONN_BLOCKS_PER_SM, std::max<long long>(1, ((long long) n_live * 8 + CONN_THREADS - 1) / CONN_THREADS));
                tm.begin(StageTimer::CONNECT);
                if (ref_set) {
                    const bool pool = conn_pool;
                    RenderCtx cc = c;
                    if (sort_cv < 0 && sort_probe) {            // second connection launch: the first one's counter is (nearly) there --
                        B2_CUDA(cudaEventSynchronize(ev_probe));  // the extend launch queued above keeps the GPU busy meanwhile
                        unsigned long long rays; memcpy(&rays, ws.h_ints + 6, sizeof(rays));
                        sort_cv = rays >= (unsigned long long) B2_CONN_SORT_MIN_RAYS * (unsigned long long) std::max(sort_probe_vertices, 1) ? 1 : 0;
                        ws.sort_cfg = sort_cfg; ws.sort_yes = sort_cv;
                        if (sort_cv > 0 && ws.ensure_cv_sort(max_paths) != B2_OK) { cudaGetLastError(); sort_cv = 0; }   // grows once per workspace
                    }
                    if (sort_cv > 0 && n_live >= B2_CONN_SORT_MIN && (!primary || B2_CONN_SORT_CV >= 2)) {
                        const uint32_t nu = (uint32_t) n_live;                 // n_cv <= n_live; the device count is read by the kernels
                        k_cv_keys<<<blocks_for(nu, 256), 256, 0, st>>>(c, nu, ws.cv_sort.keys.p, ws.cv_sort.idx.p);
                        B2_TRY(b2_sort_pairs(ws.cv_sort, nu, 31, st));
                        k_cv_gather<<<blocks_for((size_t) nu * B2_CV_F4, 256), 256, 0, st>>>(c, nu, ws.cv_sort.idx_sorted.p, ws.cv_sorted.p);
                        cc.cv = ws.cv_sorted.p;
                        n_launches += 8;                                       // keys, gather, CUB: histogram + scan + four 8-bit passes
                    }
                    if (pool) { if (count) k_connect<true, true><<<gc, CONN_THREADS, sb_c, st>>>(cc); else k_connect<false, true><<<gc, CONN_THREADS, sb_c, st>>>(cc); }
                    else { if (count) k_connect<true, false><<<gc, CONN_THREADS, sb_c, st>>>(cc); else k_connect<false, false><<<gc, CONN_THREADS, sb_c, st>>>(cc); }
                    if (sort_cv < 0 && !sort_probe) {
                        B2_CUDA(cudaMemcpyAsync(ws.h_ints + 6, ws.counters.p + B2_CNT_SHADOW, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
                        B2_CUDA(cudaEventRecord(ev_probe, st));
                    }
                } else {
                    const unsigned gn = (unsigned) std::min<long long>((long long) sc.n_sms * 8, blocks_for(n_live, EXT_THREADS));
                    if (count) k_nee<true><<<gn, EXT_THREADS, sb_t, st>>>(c, w_mix);
                    else k_nee<false><<<gn, EXT_THREADS, sb_t, st>>>(c, w_mix);
                }
                tm.end();
                n_launches += 2; n_waves++;
                B2_CUDA(cudaGetLastError());
                if (tm.on) { B2_CUDA(cudaStreamSynchronize(st)); tm.collect(); }
                B2_CUDA(cudaEventSynchronize(ev_wave));
                ws.closest_rays_host += (unsigned long long) n_live;
                n_live = ws.h_ints[0];
                eye_vertices += (unsigned long long) ws.h_ints[1];
                if (sort_cv < 0 && !sort_probe) { sort_probe = true; sort_probe_vertices = ws.h_ints[1]; }
                std::swap(c.live, c.live_next);
                primary = false;
            }
        }
    } else {
        B2_CUDA(cudaEventRecord(ev1, st));
    }
    if (p.film_mode == B2_FILM_SUM) k_resolve<<<blocks_for(film_floats, 256), 256, 0, st>>>(ws.film.p, d_film_out, film_floats, 1.0f);
    else if (p.film_mode == B2_FILM_ACCUMULATE) k_resolve_acc<<<blocks_for(film_floats, 256), 256, 0, st>>>(ws.film.p, d_film_out, film_floats, (float) p.spp);
    else k_resolve_div<<<blocks_for(film_floats, 256), 256, 0, st>>>(ws.film.p, d_film_out, film_floats, (float) p.spp);
    n_launches++;
    B2_CUDA(cudaGetLastError());
    B2_CUDA(cudaEventRecord(ev2, st));
    unsigned long long h_cnt[B2_CNT_N];
    B2_CUDA(cudaMemcpyAsync(h_cnt, ws.counters.p, sizeof(h_cnt), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        float ms_pool = 0.f, ms_total = 0.f;
        cudaEventElapsedTime(&ms_pool, ev0, ev1); cudaEventElapsedTime(&ms_total, ev0, ev2);
        stats->eye_samples = eye_pass ? eye_samples : (unsigned long long) total;
        stats->closest_rays = h_cnt[B2_CNT_CLOSEST] + ws.closest_rays_host;
        stats->shadow_rays = h_cnt[B2_CNT_SHADOW];
        stats->path_samples = h_cnt[B2_CNT_PATH_SAMPLES];
        stats->delta_cap_hits = h_cnt[B2_CNT_DELTA_CAP];
        stats->nodes_visited = h_cnt[B2_CNT_NODES];
        stats->tris_tested = h_cnt[B2_CNT_TRIS];
        stats->n_batches = n_batches; stats->n_waves = n_waves; stats->n_launches = n_launches;
        stats->ms_total = ms_total; stats->ms_light_pool = ms_pool;
        stats->pairs_tested = h_cnt[B2_CNT_PAIRS]; stats->eye_vertices = eye_vertices;
        tm.collect();
        stats->ms_primary = tm.ms[StageTimer::PRIMARY]; stats->ms_extend = tm.ms[StageTimer::EXTEND]; stats->ms_connect = tm.ms[StageTimer::CONNECT];
    }
    ws.closest_rays_host = 0;
    return B2_OK;
}

int b2_light_pool_impl(const DevScene &sc, int bvh_depth, RenderWorkspace &ws, const B2RenderParams &p, float *pool20_host, cudaStream_t st) {
    if (p.spl < 0 || p.light_depth <= 0) { b2_set_error("invalid spl / light_depth"); return B2_ERR_INVALID; }
    if ((long long) sc.n_lights * p.spl * p.light_depth >= (1ll << 27)) { b2_set_error("light pool larger than 2^27 vertices is not supported"); return B2_ERR_INVALID; }
    const int n_pool = sc.n_lights * p.spl * p.light_depth;
    if (n_pool == 0) return B2_OK;
    B2_TRY(ws.ensure(1, 1, n_pool, 3));
    B2_TRY(set_smem_limits(bvh_depth));
    B2Camera cam; memset(&cam, 0, sizeof(cam)); cam.look_at[2] = 1.f; cam.view_up[1] = 1.f; cam.fov = 50.f;
    B2RenderParams pp = p; pp.width = pp.height = 1;
    RenderCtx c; b2_setup_ctx(c, sc, cam, pp, ws, 1);
    B2_CUDA(cudaMemsetAsync(ws.counters.p, 0, sizeof(unsigned long long) * B2_CNT_N, st));
    uint64_t nl = 0;
    B2_TRY(build_pool(c, bvh_depth, false, st, nl));
    DevBuf<float> out;
    B2_TRY(out.alloc((size_t) 20 * n_pool));
    k_pool_export<<<blocks_for(n_pool, 128), 128, 0, st>>>(c, out.p);
    B2_CUDA(cudaGetLastError());
    B2_CUDA(cudaMemcpyAsync(pool20_host, out.p, sizeof(float) * 20 * (size_t) n_pool, cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    return B2_OK;
}
