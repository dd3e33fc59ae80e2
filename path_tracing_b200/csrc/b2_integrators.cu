// b2_integrators.cu -- the pt and ppm modes of the reference's command line on the wavefront / BVH infrastructure of the
// BDPT path (bodies in b2_pt.cuh).  Entered from b2_render_device_impl, so every render entry
// point of include/b2pt.h (host film, device film, progressive, multi-GPU) serves them through B2RenderParams.integrator.
//
//   pt  (run_cuda_pt, src/pt_cu.cu): the reference runs one thread per PIXEL looping over spp and depth.  Here one wave =
//       one bounce of every live path of a batch of up to 2^26 eye samples: k_pt_extend (closest hit, emission, BSDF
//       sample; survivors compacted into the next wave's queue with warp ballot + prefix) and k_pt_nee (one shadow ray
//       per vertex that takes direct light, read from a compact 80-byte record list), both adding to the film with
//       atomics.  Samples shard over GPUs as in BDPT.
//   ppm (run_cuda_ppm, src/ppm_cu.cu): eye pass -> hit points appended compactly; 63-bit cell keys radix-sorted (CUB) so
//       the hit points of a grid cell are one contiguous run; an open-addressing table maps cell -> run (the reference
//       threads a linked list per hash bucket through the unsorted hit points); one thread per photon walks the BVH and
//       adds its flux to the hit points of the 27 neighbouring cells with a 128-bit vector atomic; resolve per hit point.
#include "b2_render.hpp"
#include "b2_pt.cuh"
#include "b2_mis.cuh"
#include "b2_warp.cuh"
#include <algorithm>
#include <cub/cub.cuh>

namespace {

constexpr int PT_THREADS = 128;

// One bounce of every live path: closest hit, emission, BSDF sample; vertices that take direct light leave an 80-byte
// record in the compact list k_pt_nee consumes.  (A first version fused the shadow ray into this kernel: 5 400 SASS
// instructions, ncu: stall_no_instruction 4.6 - 8.2 per issue and 12 of 32 lanes active; profiles/r01_pt_summary.md.)
template <bool PRIMARY, bool COUNT>
__global__ void __launch_bounds__(PT_THREADS) k_pt_extend(RenderCtx c, int n_items, uint32_t pix0, int s0) {
    extern __shared__ uint2 s_stack[];
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    TravCounters tc{ 0u, 0u };
    uint32_t n_path = 0, n_cap = 0;
    PtState ps;
    EyeVtx ev;
    int path = -1;
    bool alive = false, want_nee = false;
    if (q < n_items) {
        if (PRIMARY) ps = pt_primary(c, pix0, s0, (uint32_t) q);
        else { path = c.live[q]; ps = pt_load(c, path); }
        const HitRec h = closest_hit_bvh<COUNT>(c.sc, ps.ro, ps.rd, s_stack + threadIdx.x, blockDim.x, &tc);
        float3 add;
        alive = pt_shade(c, ps, h, add, ev, want_nee, n_path, n_cap);
        if (add.x != 0.0f) atomicAdd(&c.film[3 * ps.pixel], add.x);
        if (add.y != 0.0f) atomicAdd(&c.film[3 * ps.pixel + 1], add.y);
        if (add.z != 0.0f) atomicAdd(&c.film[3 * ps.pixel + 2], add.z);
    }
    const int cvi = warp_alloc(c.n_cv, want_nee);
    if (want_nee) store_cv(c.cv, cvi, ev);
    if (PRIMARY) path = warp_alloc(c.n_paths, alive);                 // a path gets its state slot when it survives its first bounce
    const int li = warp_alloc(c.n_live_next, alive);
    if (alive) { pt_store(c, path, ps); c.live_next[li] = path; }
    warp_count_add(&c.counters[B2_CNT_PATH_SAMPLES], n_path);
    warp_count_add(&c.counters[B2_CNT_DELTA_CAP], n_cap);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}
// Next-event estimation of the vertices of this wave.  Most light samples fail the cosine / spot-cone tests before any ray is
// traced (85 % on the Cornell scene: ncu showed 6 - 8 of 32 lanes active when every lane ran prepare -> trace -> finish on its
// own record), so a warp alternates, like k_connect: while at least PT_NEE_REFILL lanes hold no shadow ray and records
// remain, the idle lanes pull records and prepare until they have a ray; otherwise every lane that holds a ray advances it by
// one wide node.  A lane whose ray is decided adds its contribution and becomes idle again.
constexpr int PT_NEE_REFILL = 12;
template <bool COUNT>
__global__ void __launch_bounds__(PT_THREADS) k_pt_nee(RenderCtx c) {
    extern __shared__ uint2 s_stack[];
    const int n_cv = *c.n_cv;
    TravCounters tc{ 0u, 0u };
    uint32_t n_shadow = 0, n_path = 0;
    const int stride = gridDim.x * blockDim.x;
    int q = blockIdx.x * blockDim.x + threadIdx.x, cur = -1;
    uint2 *stack = s_stack + threadIdx.x;
    bool busy = false;
    AnyTrav tv; NeeSample ns;
    for (;;) {
        const uint32_t busy_m = __ballot_sync(0xffffffffu, busy);
        const bool can = !busy && q < n_cv;                             // an idle lane with a record left to pull
        const int n_can = __popc(__ballot_sync(0xffffffffu, can));
        if (n_can >= PT_NEE_REFILL || (busy_m == 0u && n_can > 0)) {
            if (can) {                                                  // one record per idle lane and round
                const EyeVtx ev = load_cv(c.cv, q);
                ShadowRay sr;
                if (pt_nee_prepare(c, ev, sr, ns)) { any_init(tv, c.sc, sr.o, sr.d, sr.tmax); busy = true; cur = q; n_shadow++; }
                q += stride;
            }
        } else if (busy_m != 0u) {
            if (busy) {
                int r = any_step<COUNT, false, true>(c.sc, tv, stack, blockDim.x, &tc);
                if (r == B2_TRAV_EXHAUSTED && occluded_spheres(c.sc, tv.r.o, tv.r.d, 1e-3f, tv.tmax)) r = B2_TRAV_HIT;
                if (r != B2_TRAV_CONTINUE) {
                    busy = false;
                    if (r == B2_TRAV_EXHAUSTED) {
                        const EyeVtx ev = load_cv(c.cv, cur);
                        const float3 v = pt_nee_finish(c, ev, ns, n_path);
                        if (v.x != 0.0f) atomicAdd(&c.film[3 * ev.pixel], v.x);
                        if (v.y != 0.0f) atomicAdd(&c.film[3 * ev.pixel + 1], v.y);
                        if (v.z != 0.0f) atomicAdd(&c.film[3 * ev.pixel + 2], v.z);
                    }
                }
            }
        } else break;
    }
    warp_count_add(&c.counters[B2_CNT_SHADOW], n_shadow);
    warp_count_add(&c.counters[B2_CNT_PATH_SAMPLES], n_path);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}

// ---- B2_STRATEGIES_MIS (b2_mis.cuh): one thread per eye sample, its light subpath in local memory ----
#ifndef B2_MIS_THREADS
#define B2_MIS_THREADS 128
#endif
#ifndef B2_MIS_MIN_BLOCKS
#define B2_MIS_MIN_BLOCKS 12         // 40 registers: the kernel waits on instruction fetch and local memory, more resident warps hide both (tools/mis_ab.sh:
#endif                               // Cornell 256x256x16: 1 block 53.5 ms, 5: 52.5, 8: 47.5, 10: 46.8, 12: 45.0; 64 or 256 threads per block: no better)
constexpr int MIS_THREADS = B2_MIS_THREADS;
template <bool COUNT>
__global__ void __launch_bounds__(MIS_THREADS, B2_MIS_MIN_BLOCKS) k_bdpt_mis(RenderCtx c, CamFrame cf, int n_items, uint32_t pix0, int s0) {
    extern __shared__ uint2 s_stack[];
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    TravCounters tc{ 0u, 0u };
    MisCounters cnt{ 0u, 0u, 0u };
    const bool valid = q < n_items;                       // every lane runs the body: it votes per warp (b2_mis.cuh)
    uint32_t pixel;
    const uint32_t ds = udiv_magic(pix0 + (uint32_t) (valid ? q : 0), (uint32_t) (c.W * c.H), c.npix_magic, pixel);
    const uint32_t sidx = (uint32_t) (c.sample_offset + (s0 + (int) ds) * c.sample_stride);
    MisVtx lv[B2_MIS_MAXV];
    mis_sample_body<COUNT>(c, cf, valid, pixel, sidx, lv, s_stack + threadIdx.x, blockDim.x, &tc, cnt);
    warp_count_add(&c.counters[B2_CNT_CLOSEST], cnt.closest);
    warp_count_add(&c.counters[B2_CNT_SHADOW], cnt.shadow);
    warp_count_add(&c.counters[B2_CNT_PATH_SAMPLES], cnt.paths);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}

// ---- ppm ----
template <bool COUNT>
__global__ void __launch_bounds__(PT_THREADS) k_ppm_eye(RenderCtx c, PpmCtx pc, uint32_t sidx) {
    extern __shared__ uint2 s_stack[];
    const int pixel = blockIdx.x * blockDim.x + threadIdx.x;
    TravCounters tc{ 0u, 0u };
    uint32_t n_closest = 0, n_cap = 0;
    HitPoint hp;
    bool have = false;
    if (pixel < c.W * c.H) {
        float3 direct;
        have = ppm_eye_body<COUNT>(c, (uint32_t) pixel, sidx, s_stack + threadIdx.x, blockDim.x, &tc, hp, direct, n_closest, n_cap);
        c.film[3 * pixel] = direct.x; c.film[3 * pixel + 1] = direct.y; c.film[3 * pixel + 2] = direct.z;
    }
    const int k = warp_alloc(pc.n_hp, have);
    if (have) {
        float4 *o = pc.hp + (size_t) B2_HP_F4 * k;                   // pc.hp = the unsorted list in this kernel
        o[0] = mk4(hp.pos, 0.0f); o[1] = mk4(hp.n, u2f(hp.pixel)); o[2] = mk4(hp.wo, hp.mtl.roughness);
        o[3] = mk4(hp.thr, hp.mtl.metallic); o[4] = mk4(hp.mtl.base, hp.mtl.eta);
        int gx, gy, gz;
        ppm_cell_of(c.sc, pc.inv_cell, hp.pos, gx, gy, gz);
        pc.keys[k] = ppm_cell_key(gx, gy, gz);
        pc.order[k] = (uint32_t) k;
    }
    warp_count_add(&c.counters[B2_CNT_CLOSEST], n_closest);
    warp_count_add(&c.counters[B2_CNT_DELTA_CAP], n_cap);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}
// Sorted position k <- unsorted record order[k]; the first hit point of every cell claims a table slot {key + 1, start}.
__global__ void __launch_bounds__(256) k_ppm_reorder(const float4 *__restrict__ hp_in, PpmCtx pc, const unsigned long long *__restrict__ skeys,
                                                     const uint32_t *__restrict__ sorder, int n) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const float4 *src = hp_in + (size_t) B2_HP_F4 * sorder[k];
    float4 *dst = pc.hp + (size_t) B2_HP_F4 * k;
#pragma unroll
    for (int j = 0; j < B2_HP_F4; ++j) dst[j] = src[j];
    pc.flux[k] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    const unsigned long long key = skeys[k];
    if (k == 0 || skeys[k - 1] != key) {
        unsigned long long *tab = (unsigned long long *) pc.table;
        uint32_t slot = ppm_hash(key + 1ull) & pc.table_mask;
        while (atomicCAS(&tab[2 * (size_t) slot], 0ull, key + 1ull) != 0ull) slot = (slot + 1u) & pc.table_mask;   // keys are distinct here
        tab[2 * (size_t) slot + 1] = (unsigned long long) (uint32_t) k;
    }
}
// The last hit point of every cell completes its entry: {key + 1, start | count << 32}.
__global__ void __launch_bounds__(256) k_ppm_cell_ends(PpmCtx pc, const unsigned long long *__restrict__ skeys, int n) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const unsigned long long key = skeys[k];
    if (k + 1 < n && skeys[k + 1] == key) return;
    unsigned long long *tab = (unsigned long long *) pc.table;
    uint32_t slot = ppm_hash(key + 1ull) & pc.table_mask;
    while (tab[2 * (size_t) slot] != key + 1ull) slot = (slot + 1u) & pc.table_mask;
    const unsigned long long start = tab[2 * (size_t) slot + 1];
    tab[2 * (size_t) slot + 1] = start | ((unsigned long long) ((uint32_t) k + 1u - (uint32_t) start) << 32);
}
struct FluxAtomic {
    float4 *flux;
    __device__ __forceinline__ void operator()(uint32_t k, float3 e) const { atomicAdd(&flux[k], make_float4(e.x, e.y, e.z, 1.0f)); }   // red.global.add.v4.f32
};
template <bool COUNT>
__global__ void __launch_bounds__(PT_THREADS) k_ppm_photons(RenderCtx c, PpmCtx pc, int n_mine) {
    extern __shared__ uint2 s_stack[];
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    TravCounters tc{ 0u, 0u };
    uint32_t n_closest = 0, n_gather = 0, n_cap = 0;
    if (q < n_mine) {
        FluxAtomic fa{ pc.flux };
        ppm_photon_body<COUNT>(c, pc, c.sample_offset + q * c.sample_stride, s_stack + threadIdx.x, blockDim.x, &tc, fa, n_closest, n_gather, n_cap);
    }
    warp_count_add(&c.counters[B2_CNT_CLOSEST], n_closest);
    warp_count_add(&c.counters[B2_CNT_PATH_SAMPLES], n_gather);
    warp_count_add(&c.counters[B2_CNT_DELTA_CAP], n_cap);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}
__global__ void __launch_bounds__(256) k_ppm_resolve(RenderCtx c, PpmCtx pc, int n) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t pixel = f2u(pc.hp[(size_t) B2_HP_F4 * k + 1].w);
    const float4 f = pc.flux[k];
    const float3 direct = mk3(c.film[3 * pixel], c.film[3 * pixel + 1], c.film[3 * pixel + 2]);
    const float3 r = ppm_resolve(direct, mk3(f), true, pc.radius);
    c.film[3 * pixel] = r.x; c.film[3 * pixel + 1] = r.y; c.film[3 * pixel + 2] = r.z;
}

__global__ void __launch_bounds__(256) k_film_out(const float *__restrict__ sum, float *__restrict__ out, size_t n, float inv, int accumulate) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = accumulate ? out[i] + sum[i] * inv : sum[i] * inv;
}

int set_limits(int bvh_depth) {
    static int configured[64];
    int dev = 0;
    B2_CUDA(cudaGetDevice(&dev));
    if (dev < 64 && bvh_depth + 1 <= configured[dev]) return B2_OK;
    const int sb = (int) stack_bytes(PT_THREADS, bvh_depth);
    if (sb > 200 * 1024) { b2_set_error("BVH depth %d needs more shared memory for traversal stacks than a CTA has", bvh_depth); return B2_ERR_STATE; }
    B2_CUDA(cudaFuncSetAttribute(k_pt_extend<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sb));
    B2_CUDA(cudaFuncSetAttribute(k_pt_extend<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sb));
    B2_CUDA(cudaFuncSetAttribute(k_pt_extend<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sb));
    B2_CUDA(cudaFuncSetAttribute(k_pt_extend<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sb));
    B2_CUDA(cudaFuncSetAttribute(k_pt_nee<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sb));
    B2_CUDA(cudaFuncSetAttribute(k_pt_nee<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sb));
    B2_CUDA(cudaFuncSetAttribute(k_bdpt_mis<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) stack_bytes(MIS_THREADS, bvh_depth)));
    B2_CUDA(cudaFuncSetAttribute(k_bdpt_mis<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) stack_bytes(MIS_THREADS, bvh_depth)));
    B2_CUDA(cudaFuncSetAttribute(k_ppm_eye<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sb));
    B2_CUDA(cudaFuncSetAttribute(k_ppm_eye<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sb));
    B2_CUDA(cudaFuncSetAttribute(k_ppm_photons<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sb));
    B2_CUDA(cudaFuncSetAttribute(k_ppm_photons<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sb));
    if (dev < 64) configured[dev] = bvh_depth + 1;
    return B2_OK;
}

void fill_stats(B2Stats *stats, const unsigned long long *h_cnt, unsigned long long eye_samples, unsigned long long closest_host,
                uint64_t n_batches, uint64_t n_waves, uint64_t n_launches, float ms_total) {
    if (!stats) return;
    memset(stats, 0, sizeof(*stats));
    stats->eye_samples = eye_samples;
    stats->closest_rays = h_cnt[B2_CNT_CLOSEST] + closest_host;
    stats->shadow_rays = h_cnt[B2_CNT_SHADOW];
    stats->path_samples = h_cnt[B2_CNT_PATH_SAMPLES];
    stats->delta_cap_hits = h_cnt[B2_CNT_DELTA_CAP];
    stats->nodes_visited = h_cnt[B2_CNT_NODES]; stats->tris_tested = h_cnt[B2_CNT_TRIS];
    stats->n_batches = n_batches; stats->n_waves = n_waves; stats->n_launches = n_launches;
    stats->ms_total = ms_total;
}

}  // namespace

int RenderWorkspace::ensure_pt(int max_paths, size_t film_floats) {
    const size_t mp = (size_t) std::max(max_paths, 1);
    B2_TRY(ray_o.ensure(mp)); B2_TRY(ray_d.ensure(mp)); B2_TRY(thr.ensure(mp)); B2_TRY(last_pos.ensure(mp));
    B2_TRY(live_a.ensure(mp)); B2_TRY(live_b.ensure(mp)); B2_TRY(cv.ensure((size_t) B2_CV_F4 * mp));
    B2_TRY(film.ensure(film_floats));
    if (!ints.p) B2_TRY(ints.alloc(8));
    if (!counters.p) B2_TRY(counters.alloc(B2_CNT_N));
    if (!h_ints) B2_CUDA(cudaMallocHost((void **) &h_ints, 8 * sizeof(int)));
    return B2_OK;
}
int RenderWorkspace::ensure_ppm(int n_pix, size_t film_floats, size_t sort_tmp_bytes) {
    const size_t n = (size_t) std::max(n_pix, 1);
    B2_TRY(hp_in.ensure(B2_HP_F4 * n)); B2_TRY(hp.ensure(B2_HP_F4 * n)); B2_TRY(flux.ensure(n));
    B2_TRY(hp_keys.ensure(n)); B2_TRY(hp_keys_sorted.ensure(n)); B2_TRY(hp_order.ensure(n)); B2_TRY(hp_order_sorted.ensure(n));
    size_t cap = 1024;
    while (cap < 2 * n) cap <<= 1;
    B2_TRY(cell_table.ensure(cap));
    B2_TRY(sort_tmp.ensure(sort_tmp_bytes));
    B2_TRY(film.ensure(film_floats));
    if (!ints.p) B2_TRY(ints.alloc(8));
    if (!counters.p) B2_TRY(counters.alloc(B2_CNT_N));
    if (!h_ints) B2_CUDA(cudaMallocHost((void **) &h_ints, 8 * sizeof(int)));
    return B2_OK;
}

// ---------------------------------------------------------------------------------------------------------
// pt: run_cuda_pt (src/pt_cu_helper.cpp:62-72 -> pt_render_wrapper, src/pt_cu.cu:257-303)
// ---------------------------------------------------------------------------------------------------------
int b2_render_pt_impl(const DevScene &sc_in, int bvh_depth, RenderWorkspace &ws, const B2Camera &cam, const B2RenderParams &p,
                      float *d_film_out, cudaStream_t st, B2Stats *stats) {
    const int npix = p.width * p.height;
    const size_t film_floats = (size_t) 3 * npix;
    const int stride = p.sample_stride > 0 ? p.sample_stride : 1;
    int n_my = 0;
    for (long long s = p.sample_offset; s < p.spp; s += stride) n_my++;
    const long long total = (long long) npix * n_my;
    long long mp = p.max_batch_paths > 0 ? p.max_batch_paths : (1ll << 26);     // 72 B of state per path: 4.8 GB at 2^26
    mp = std::max<long long>(std::min<long long>(mp, std::max<long long>(total, 1)), 1);
    {   // halve the batch until its state (4 float4 rows, two queues, the 80-byte vertex records) fits half of the free memory
        size_t free_b = 0, total_b = 0;
        B2_CUDA(cudaMemGetInfo(&free_b, &total_b));
        const size_t per_path = 4 * sizeof(float4) + 2 * sizeof(int) + (size_t) B2_CV_F4 * sizeof(float4);
        const size_t have = (ws.ray_o.n + ws.ray_d.n + ws.thr.n + ws.last_pos.n + ws.cv.n) * sizeof(float4) + (ws.live_a.n + ws.live_b.n) * sizeof(int);
        while (mp > 4096 && per_path * (size_t) mp > (free_b + have) / 2) mp /= 2;
    }
    const int max_paths = (int) mp;
    B2_TRY(ws.ensure_pt(max_paths, film_floats));
    B2_TRY(set_limits(bvh_depth));
    DevScene sc = sc_in;
    sc.gpu_sem = 1;
    RenderCtx c; b2_setup_ctx(c, sc, cam, p, ws, max_paths);
    const bool count = p.count_traversal != 0;
    uint64_t n_launches = 0, n_batches = 0, n_waves = 0;
    unsigned long long eye_samples = 0, closest_host = 0;
    CudaEvent ev0, ev1, ev_wave;
    B2_CUDA(ev0.create()); B2_CUDA(ev1.create()); B2_CUDA(ev_wave.create(cudaEventDisableTiming));
    B2_CUDA(cudaEventRecord(ev0, st));
    B2_CUDA(cudaMemsetAsync(ws.film.p, 0, sizeof(float) * film_floats, st));
    B2_CUDA(cudaMemsetAsync(ws.counters.p, 0, sizeof(unsigned long long) * B2_CNT_N, st));
    B2_CUDA(cudaMemsetAsync(ws.ints.p, 0, sizeof(int) * 8, st));
    const size_t sb = stack_bytes(PT_THREADS, bvh_depth);
    for (long long w0 = 0; w0 < total && p.eye_depth > 0; w0 += max_paths) {
        int n_live = (int) std::min<long long>(max_paths, total - w0);
        c.live = ws.live_a.p; c.live_next = ws.live_b.p;
        n_batches++;
        eye_samples += (unsigned long long) n_live;
        B2_CUDA(cudaMemsetAsync(c.n_paths, 0, sizeof(int), st));
        bool primary = true;
        while (n_live > 0) {
            B2_CUDA(cudaMemsetAsync(c.n_live_next, 0, 2 * sizeof(int), st));             // n_live_next, n_cv
            const unsigned gb = blocks_for(n_live, PT_THREADS);
            if (primary) {
                const uint32_t pix0 = (uint32_t) (w0 % npix); const int s0 = (int) (w0 / npix);
                if (count) k_pt_extend<true, true><<<gb, PT_THREADS, sb, st>>>(c, n_live, pix0, s0);
                else k_pt_extend<true, false><<<gb, PT_THREADS, sb, st>>>(c, n_live, pix0, s0);
            } else {
                if (count) k_pt_extend<false, true><<<gb, PT_THREADS, sb, st>>>(c, n_live, 0u, 0);
                else k_pt_extend<false, false><<<gb, PT_THREADS, sb, st>>>(c, n_live, 0u, 0);
            }
            B2_CUDA(cudaMemcpyAsync(ws.h_ints, c.n_live_next, sizeof(int), cudaMemcpyDeviceToHost, st));
            B2_CUDA(cudaEventRecord(ev_wave, st));                 // the host waits for the wave size only, not for k_pt_nee (b2_render.cu)
            const unsigned gn = (unsigned) std::min<long long>((long long) sc.n_sms * 16, gb);
            if (count) k_pt_nee<true><<<gn, PT_THREADS, sb, st>>>(c);
            else k_pt_nee<false><<<gn, PT_THREADS, sb, st>>>(c);
            n_launches += 2; n_waves++;
            B2_CUDA(cudaGetLastError());
            B2_CUDA(cudaEventSynchronize(ev_wave));
            closest_host += (unsigned long long) n_live;
            n_live = ws.h_ints[0];
            std::swap(c.live, c.live_next);
            primary = false;
        }
    }
    const float inv = p.film_mode == B2_FILM_SUM ? 1.0f : 1.0f / (float) p.spp;
    k_film_out<<<blocks_for(film_floats, 256), 256, 0, st>>>(ws.film.p, d_film_out, film_floats, inv, p.film_mode == B2_FILM_ACCUMULATE);
    n_launches++;
    B2_CUDA(cudaGetLastError());
    B2_CUDA(cudaEventRecord(ev1, st));
    unsigned long long h_cnt[B2_CNT_N];
    B2_CUDA(cudaMemcpyAsync(h_cnt, ws.counters.p, sizeof(h_cnt), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev0, ev1);
    fill_stats(stats, h_cnt, eye_samples, closest_host, n_batches, n_waves, n_launches, ms);
    return B2_OK;
}

// ---------------------------------------------------------------------------------------------------------
// B2_STRATEGIES_MIS: every s/t strategy with real densities (b2_mis.cuh).  Batches of 2^22 eye samples, one launch each.
// ---------------------------------------------------------------------------------------------------------
int b2_render_mis_impl(const DevScene &sc, int bvh_depth, RenderWorkspace &ws, const B2Camera &cam, const B2RenderParams &p,
                       float *d_film_out, cudaStream_t st, B2Stats *stats) {
    const int npix = p.width * p.height;
    const size_t film_floats = (size_t) 3 * npix;
    const int stride = p.sample_stride > 0 ? p.sample_stride : 1;
    int n_my = 0;
    for (long long s = p.sample_offset; s < p.spp; s += stride) n_my++;
    const long long total = (long long) npix * n_my;
    if (p.eye_depth > B2_MIS_MAXV || p.light_depth - 1 > B2_MIS_MAXV) {          // a subpath stores at most B2_MIS_MAXV vertices, delta ones included
        b2_set_error("B2_STRATEGIES_MIS supports eye_depth <= %d and light_depth <= %d", B2_MIS_MAXV, B2_MIS_MAXV + 1); return B2_ERR_INVALID;
    }
    // the wavefront (b2_mis_wave.cu) is the product path; B2_MIS_MEGAKERNEL=1 keeps the one-thread-per-sample kernel below for A/B runs
    if (const char *e = getenv("B2_MIS_MEGAKERNEL"); !(e && atoi(e) != 0)) return b2_render_mis_wave_impl(sc, bvh_depth, ws, cam, p, d_film_out, st, stats);
    B2_TRY(ws.film.ensure(film_floats));
    if (!ws.ints.p) B2_TRY(ws.ints.alloc(8));
    if (!ws.counters.p) B2_TRY(ws.counters.alloc(B2_CNT_N));
    B2_TRY(set_limits(bvh_depth));
    RenderCtx c; b2_setup_ctx(c, sc, cam, p, ws, 1);
    const bool count = p.count_traversal != 0;
    uint64_t n_launches = 0, n_batches = 0;
    CudaEvent ev0, ev1, ev_wave;
    B2_CUDA(ev0.create()); B2_CUDA(ev1.create()); B2_CUDA(ev_wave.create(cudaEventDisableTiming));
    B2_CUDA(cudaEventRecord(ev0, st));
    B2_CUDA(cudaMemsetAsync(ws.film.p, 0, sizeof(float) * film_floats, st));
    B2_CUDA(cudaMemsetAsync(ws.counters.p, 0, sizeof(unsigned long long) * B2_CNT_N, st));
    const size_t sb = stack_bytes(MIS_THREADS, bvh_depth);
    const CamFrame cf = cam_frame(c.cdx, c.cdy);
    const long long batch = p.max_batch_paths > 0 ? p.max_batch_paths : (1ll << 22);
    for (long long w0 = 0; w0 < total && sc.n_lights > 0; w0 += batch) {
        const int n = (int) std::min<long long>(batch, total - w0);
        const uint32_t pix0 = (uint32_t) (w0 % npix); const int s0 = (int) (w0 / npix);
        if (count) k_bdpt_mis<true><<<blocks_for(n, MIS_THREADS), MIS_THREADS, sb, st>>>(c, cf, n, pix0, s0);
        else k_bdpt_mis<false><<<blocks_for(n, MIS_THREADS), MIS_THREADS, sb, st>>>(c, cf, n, pix0, s0);
        n_launches++; n_batches++;
    }
    B2_CUDA(cudaGetLastError());
    const float inv = p.film_mode == B2_FILM_SUM ? 1.0f : 1.0f / (float) p.spp;
    k_film_out<<<blocks_for(film_floats, 256), 256, 0, st>>>(ws.film.p, d_film_out, film_floats, inv, p.film_mode == B2_FILM_ACCUMULATE);
    n_launches++;
    B2_CUDA(cudaGetLastError());
    B2_CUDA(cudaEventRecord(ev1, st));
    unsigned long long h_cnt[B2_CNT_N];
    B2_CUDA(cudaMemcpyAsync(h_cnt, ws.counters.p, sizeof(h_cnt), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev0, ev1);
    fill_stats(stats, h_cnt, (unsigned long long) total, 0ull, n_batches, n_batches, n_launches, ms);
    return B2_OK;
}

// ---------------------------------------------------------------------------------------------------------
// ppm: run_cuda_ppm (src/ppm_cu_helper.cpp:62-70 -> ppm_render_wrapper, src/ppm_cu.cu:325-398)
// ---------------------------------------------------------------------------------------------------------
int b2_render_ppm_impl(const DevScene &sc_in, int bvh_depth, RenderWorkspace &ws, const B2Camera &cam, const B2RenderParams &p,
                       float *d_film_out, cudaStream_t st, B2Stats *stats) {
    if ((p.sample_stride > 1 || p.sample_offset > 0)) {
        b2_set_error("the ppm integrator does not shard: the photon flux of all photons is resolved with a clamp (ppm_cu.cu:293-315); render it on one GPU");
        return B2_ERR_INVALID;
    }
    if ((long long) sc_in.n_lights * p.spl > (1ll << 30)) { b2_set_error("too many photons (n_lights * spl > 2^30)"); return B2_ERR_INVALID; }
    const int npix = p.width * p.height;
    const size_t film_floats = (size_t) 3 * npix;
    size_t sort_bytes = 0;
    B2_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (unsigned long long *) nullptr, (unsigned long long *) nullptr,
                                            (uint32_t *) nullptr, (uint32_t *) nullptr, npix, 0, 63, st));
    B2_TRY(ws.ensure_ppm(npix, film_floats, sort_bytes));
    B2_TRY(set_limits(bvh_depth));
    DevScene sc = sc_in;
    sc.gpu_sem = 1;
    RenderCtx c; b2_setup_ctx(c, sc, cam, p, ws, 1);
    PpmCtx pc;
    pc.radius = p.ppm_radius > 0.0f ? p.ppm_radius : B2_PPM_RADIUS;
    pc.inv_cell = 1.0f / pc.radius;
    pc.hp = ws.hp_in.p; pc.flux = ws.flux.p; pc.direct = ws.film.p;
    pc.keys = ws.hp_keys.p; pc.order = ws.hp_order.p;
    pc.table = ws.cell_table.p; pc.table_mask = (uint32_t) (ws.cell_table.n - 1);
    pc.n_hp = ws.ints.p + 5;
    const bool count = p.count_traversal != 0;
    uint64_t n_launches = 0;
    CudaEvent ev0, ev1, ev_wave;
    B2_CUDA(ev0.create()); B2_CUDA(ev1.create()); B2_CUDA(ev_wave.create(cudaEventDisableTiming));
    B2_CUDA(cudaEventRecord(ev0, st));
    B2_CUDA(cudaMemsetAsync(ws.counters.p, 0, sizeof(unsigned long long) * B2_CNT_N, st));
    B2_CUDA(cudaMemsetAsync(ws.ints.p, 0, sizeof(int) * 8, st));
    B2_CUDA(cudaMemsetAsync(ws.cell_table.p, 0, sizeof(ulonglong2) * ws.cell_table.n, st));
    const size_t sb = stack_bytes(PT_THREADS, bvh_depth);
    // 1. eye pass (ppm_cu.cu:360-365): hit points, appended compactly
    if (count) k_ppm_eye<true><<<blocks_for(npix, PT_THREADS), PT_THREADS, sb, st>>>(c, pc, (uint32_t) p.sample_offset);
    else k_ppm_eye<false><<<blocks_for(npix, PT_THREADS), PT_THREADS, sb, st>>>(c, pc, (uint32_t) p.sample_offset);
    B2_CUDA(cudaMemcpyAsync(ws.h_ints, pc.n_hp, sizeof(int), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    B2_CUDA(cudaGetLastError());
    n_launches++;
    const int n_hp = ws.h_ints[0];
    const int n_photons = sc.n_lights * p.spl;                               // ppm_cu.cu:351
    if (n_hp > 0) {
        // 2. cell order + cell table (replaces reset_hash_grid / build_hash_grid_kernel, ppm_cu.cu:40-58, 368-378)
        B2_CUDA(cub::DeviceRadixSort::SortPairs(ws.sort_tmp.p, sort_bytes, ws.hp_keys.p, ws.hp_keys_sorted.p, ws.hp_order.p, ws.hp_order_sorted.p, n_hp, 0, 63, st));
        pc.hp = ws.hp.p;
        k_ppm_reorder<<<blocks_for(n_hp, 256), 256, 0, st>>>(ws.hp_in.p, pc, ws.hp_keys_sorted.p, ws.hp_order_sorted.p, n_hp);
        k_ppm_cell_ends<<<blocks_for(n_hp, 256), 256, 0, st>>>(pc, ws.hp_keys_sorted.p, n_hp);
        n_launches += 3;
        // 3. photons (ppm_cu.cu:381-386)
        if (n_photons > 0 && p.light_depth > 0) {
            if (count) k_ppm_photons<true><<<blocks_for(n_photons, PT_THREADS), PT_THREADS, sb, st>>>(c, pc, n_photons);
            else k_ppm_photons<false><<<blocks_for(n_photons, PT_THREADS), PT_THREADS, sb, st>>>(c, pc, n_photons);
            n_launches++;
        }
        // 4. resolve (ppm_cu.cu:389-392)
        k_ppm_resolve<<<blocks_for(n_hp, 256), 256, 0, st>>>(c, pc, n_hp);
        n_launches++;
    }
    k_film_out<<<blocks_for(film_floats, 256), 256, 0, st>>>(ws.film.p, d_film_out, film_floats, 1.0f, p.film_mode == B2_FILM_ACCUMULATE);
    n_launches++;
    B2_CUDA(cudaGetLastError());
    B2_CUDA(cudaEventRecord(ev1, st));
    unsigned long long h_cnt[B2_CNT_N];
    B2_CUDA(cudaMemcpyAsync(h_cnt, ws.counters.p, sizeof(h_cnt), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev0, ev1);
    fill_stats(stats, h_cnt, (unsigned long long) npix, 0ull, 1, 1, n_launches, ms);
    if (stats) { stats->eye_vertices = (unsigned long long) n_hp; stats->pairs_tested = (unsigned long long) n_photons; }
    return B2_OK;
}
