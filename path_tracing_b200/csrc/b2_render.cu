// b2_render.cu -- wavefront BDPT on the device: __global__ wrappers around the stage bodies of b2_bdpt.cuh and
// the host wave loop behind b2_render_bdpt / b2_render_bdpt_device / b2_light_pool (include/b2pt.h).
//
// Replaces run_cpu_bdpt (src/cpu_bdpt.cpp:173-488) / run_cuda_bdpt (include/bdpt_cu_helper.h:6).  Layout of one
// render (everything stays in HBM between stages; the scene tables and the light pool are L2-resident):
//
//   k_light_paths  one thread per light subpath -> pool (5 x float4 per vertex)         cpu_bdpt.cpp:218-325
//   k_pool_active  ordered compaction of the pool vertices that can contribute           :391
//   per batch of <= max_paths eye samples:
//     k_eye_gen    primary rays + path state                                             :349-361
//     per wave (one bounce of every live path):
//       k_closest  closest-hit traversal of the 8-wide BVH, shared-memory stacks         :369
//       k_shade    eye vertex record, BSDF sample, next ray; live-path compaction with
//                  warp ballot + prefix and one atomic per warp                          :372-385, :442-472
//       k_connect  eye vertex x pool: pre-tests for all pairs, survivors compacted in a
//                  shared-memory ring and finished at full occupancy (BSDFs, shadow
//                  any-hit, MIS), summed per eye vertex in shared memory                 :389-439
//     k_flush      per-path radiance -> film (atomicAdd)                                 :475-476
//   k_resolve      film / spp                                                            :479
#include "b2_render.hpp"
#include <algorithm>
#include <vector>

namespace {

constexpr int TRAV_THREADS = 128;     // closest-hit kernel block
constexpr int SHADE_THREADS = 256;
constexpr int CONN_THREADS = 256;     // connect kernel block: one eye vertex per thread in the pre-test phase
constexpr int CONN_CHUNK = 8;         // pool vertices pre-tested between two block barriers
constexpr int CONN_QCAP = CONN_THREADS * (CONN_CHUNK + 1);   // ring capacity: a full chunk on top of < one drain

__device__ __forceinline__ uint32_t lanemask_lt() { uint32_t m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m; }

// warp-reduce a per-thread count and add it to a global 64-bit counter with one atomic per warp
__device__ __forceinline__ void warp_count_add(unsigned long long *counter, uint32_t v) {
    v = __reduce_add_sync(0xffffffffu, v);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(counter, (unsigned long long) v);
}

template <bool COUNT>
__global__ void __launch_bounds__(TRAV_THREADS) k_light_paths(RenderCtx c, int n_paths) {
    extern __shared__ uint2 s_stack[];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    TravCounters tc{ 0u, 0u };
    uint32_t ncl = 0, ncap = 0;
    if (i < n_paths) light_path_body<COUNT>(c, i, s_stack + threadIdx.x, blockDim.x, &tc, ncl, ncap);
    warp_count_add(&c.counters[B2_CNT_CLOSEST], ncl);
    warp_count_add(&c.counters[B2_CNT_DELTA_CAP], ncap);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}

// Single block: ascending list of pool vertices with |throughput| >= 1e-6 (deterministic order).
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
        if (act) c.pool_active[s_base + off + __popc(m & lanemask_lt())] = i;
        __syncthreads();
        if (threadIdx.x == 0) s_base += tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) *c.n_pool_active = s_base;
}

__global__ void __launch_bounds__(256) k_eye_gen(RenderCtx c, int n, long long w0) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) eye_gen_body(c, i, w0);
}

template <bool COUNT>
__global__ void __launch_bounds__(TRAV_THREADS) k_closest(RenderCtx c, int n_live) {
    extern __shared__ uint2 s_stack[];
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    TravCounters tc{ 0u, 0u };
    if (q < n_live) {
        const int path = c.live[q];
        const HitRec h = closest_hit_bvh<COUNT>(c.sc, mk3(c.ray_o[path]), mk3(c.ray_d[path]), s_stack + threadIdx.x, blockDim.x, &tc);
        c.hit[q] = make_int2(h.prim, (int) __float_as_uint(h.t));
    }
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}

__global__ void __launch_bounds__(SHADE_THREADS) k_shade(RenderCtx c, int n_live) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t emit = 0, cap = 0;
    bool alive = false;
    int path = 0;
    if (q < n_live) { path = c.live[q]; alive = shade_body(c, q, emit, cap); }
    // live-path compaction: rank within the warp from the ballot, one atomic per warp for the base
    const uint32_t m = __ballot_sync(0xffffffffu, alive);
    if (m) {
        int base = 0;
        if ((threadIdx.x & 31) == 0) base = atomicAdd(c.n_live_next, __popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (alive) c.live_next[base + __popc(m & lanemask_lt())] = path;
    }
    warp_count_add(&c.counters[B2_CNT_PATH_SAMPLES], emit);
    warp_count_add(&c.counters[B2_CNT_DELTA_CAP], cap);
}

// Connection stage.  Phase A (every thread owns the eye vertex of queue slot q and walks the active pool list;
// the pool record loads are warp-uniform broadcasts): pairs that pass connect_pretest are pushed into a
// shared-memory ring as (owner thread << 24 | pool index).  Phase B: whenever the ring holds a full block of
// entries they are finished with every lane busy (connect_finish traces the shadow segment), and the result is
// added to the owner's accumulator in shared memory.  One global read-modify-write per eye vertex at the end.
template <bool COUNT>
__global__ void __launch_bounds__(CONN_THREADS) k_connect(RenderCtx c, int n_live) {
    extern __shared__ uint2 s_stack[];                    // CONN_THREADS x stack depth
    __shared__ uint32_t s_ring[CONN_QCAP];
    __shared__ float s_acc[3 * CONN_THREADS];
    __shared__ uint32_t s_tail;
    const int tid = threadIdx.x;
    const int q = blockIdx.x * CONN_THREADS + tid;
    const int qbase = blockIdx.x * CONN_THREADS;
    bool valid = false;
    float3 my_pos = mk3(0, 0, 0), my_n = mk3(0, 0, 0);
    if (q < n_live) {
        const float4 a = c.cv[B2_CV_F4 * q];
        if (__float_as_uint(a.w) != B2_CV_NONE) { valid = true; my_pos = mk3(a); my_n = mk3(c.cv[B2_CV_F4 * q + 1]); }
    }
    s_acc[tid] = 0.0f; s_acc[tid + CONN_THREADS] = 0.0f; s_acc[tid + 2 * CONN_THREADS] = 0.0f;
    if (tid == 0) s_tail = 0u;
    __syncthreads();
    if (__syncthreads_or(valid) == 0) return;            // no eye vertex in this block
    const int n_act = *c.n_pool_active;
    uint32_t head = 0u;                                   // block-uniform
    uint32_t n_shadow = 0, n_path = 0, n_pairs = 0;
    TravCounters tc{ 0u, 0u };

    auto drain = [&](uint32_t count) {                    // count <= CONN_THREADS entries starting at head
        if ((uint32_t) tid < count) {
            const uint32_t e = s_ring[(head + (uint32_t) tid) % CONN_QCAP];
            const int owner = (int) (e >> 24), lvi = (int) (e & 0xffffffu);
            const EyeVtx ev = load_cv(c.cv, qbase + owner);
            const float3 r = connect_finish<COUNT>(c, ev, lvi, s_stack + tid, CONN_THREADS, &tc, n_shadow, n_path);
            if (r.x != 0.0f) atomicAdd(&s_acc[owner], r.x);
            if (r.y != 0.0f) atomicAdd(&s_acc[owner + CONN_THREADS], r.y);
            if (r.z != 0.0f) atomicAdd(&s_acc[owner + 2 * CONN_THREADS], r.z);
        }
        head += count;
    };

    for (int a0 = 0; a0 < n_act; a0 += CONN_CHUNK) {
        const int a1 = min(a0 + CONN_CHUNK, n_act);
        for (int a = a0; a < a1; ++a) {
            const int lvi = c.pool_active[a];
            const bool pass = valid && connect_pretest(c, my_pos, my_n, lvi);
            if (valid) n_pairs++;
            const uint32_t m = __ballot_sync(0xffffffffu, pass);
            if (m) {
                uint32_t base = 0u;
                if ((tid & 31) == 0) base = atomicAdd(&s_tail, (uint32_t) __popc(m));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (pass) s_ring[(base + (uint32_t) __popc(m & lanemask_lt())) % CONN_QCAP] = ((uint32_t) tid << 24) | (uint32_t) lvi;
            }
        }
        __syncthreads();
        const uint32_t tail = s_tail;
        while (tail - head >= (uint32_t) CONN_THREADS) drain(CONN_THREADS);
        __syncthreads();                                  // ring slots of drained entries may be overwritten now
    }
    {
        const uint32_t tail = s_tail;
        if (tail != head) drain(tail - head);
    }
    __syncthreads();
    if (valid) {
        const float3 r = mk3(s_acc[tid], s_acc[tid + CONN_THREADS], s_acc[tid + 2 * CONN_THREADS]);
        if (r.x != 0.0f || r.y != 0.0f || r.z != 0.0f) {
            const int path = (int) __float_as_uint(c.cv[B2_CV_F4 * q + 1].w);
            float4 acc = c.acc[path];
            acc.x += r.x; acc.y += r.y; acc.z += r.z;
            c.acc[path] = acc;
        }
    }
    warp_count_add(&c.counters[B2_CNT_SHADOW], n_shadow);
    warp_count_add(&c.counters[B2_CNT_PATH_SAMPLES], n_path);
    warp_count_add(&c.counters[B2_CNT_PAIRS], n_pairs);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}

// cpu_bdpt.cpp:475-476: a sample whose radiance is not a valid colour contributes nothing
__global__ void __launch_bounds__(256) k_flush(RenderCtx c, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = c.acc[i];
    const float3 v = mk3(a);
    if (!valid_color(v)) return;
    const uint32_t pixel = c.pix[i];
    if (v.x != 0.0f) atomicAdd(&c.film[3 * pixel], v.x);
    if (v.y != 0.0f) atomicAdd(&c.film[3 * pixel + 1], v.y);
    if (v.z != 0.0f) atomicAdd(&c.film[3 * pixel + 2], v.z);
}

__global__ void __launch_bounds__(256) k_resolve(const float *__restrict__ sum, float *__restrict__ out, size_t n, float inv) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = sum[i] * inv;
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
        B2_TRY(acc.alloc(mp)); B2_TRY(pix.alloc(mp)); B2_TRY(sidx.alloc(mp)); B2_TRY(pstate.alloc(mp)); B2_TRY(hit.alloc(mp));
        B2_TRY(hist_pe.alloc((size_t) ed * mp)); B2_TRY(hist_pdf.alloc((size_t) ed * mp));
        B2_TRY(live_a.alloc(mp)); B2_TRY(live_b.alloc(mp)); B2_TRY(cv.alloc((size_t) B2_CV_F4 * mp));
        cap_paths = mp; cap_depth = ed;
    }
    if (n_pool > cap_pool) { B2_TRY(pool.alloc((size_t) B2_LV_F4 * n_pool)); B2_TRY(pool_active.alloc(n_pool)); cap_pool = n_pool; }
    if (film_floats > film.n) B2_TRY(film.alloc(film_floats));
    if (!ints.p) B2_TRY(ints.alloc(8));
    if (!counters.p) B2_TRY(counters.alloc(B2_CNT_N));
    if (!h_ints) B2_CUDA(cudaMallocHost((void **) &h_ints, 8 * sizeof(int)));
    return B2_OK;
}
void RenderWorkspace::release() {
    ray_o.release(); ray_d.release(); thr.release(); last_pos.release(); last_n.release(); acc.release(); pix.release(); sidx.release();
    pstate.release(); hit.release(); hist_pe.release(); hist_pdf.release(); live_a.release(); live_b.release(); cv.release();
    pool.release(); pool_active.release(); film.release(); ints.release(); counters.release();
    if (h_ints) { cudaFreeHost(h_ints); h_ints = nullptr; }
    cap_paths = cap_depth = cap_pool = 0;
}

static int validate_params(const B2RenderParams &p) {
    if (p.width <= 0 || p.height <= 0 || p.spp <= 0 || p.spl < 0 || p.eye_depth < 0 || p.light_depth <= 0) {
        b2_set_error("invalid render parameters (width/height/spp must be > 0, light_depth > 0, spl/eye_depth >= 0)"); return B2_ERR_INVALID;
    }
    if (p.eye_depth > 255) { b2_set_error("eye_depth > 255 is not supported"); return B2_ERR_INVALID; }
    if ((long long) p.width * p.height > (1ll << 30)) { b2_set_error("film too large"); return B2_ERR_INVALID; }
    if (p.sample_offset < 0 || p.sample_stride < 0) { b2_set_error("invalid sample_offset / sample_stride"); return B2_ERR_INVALID; }
    return B2_OK;
}

static void setup_ctx(RenderCtx &c, const DevScene &sc, const B2Camera &cam, const B2RenderParams &p, RenderWorkspace &ws, int max_paths) {
    memset(&c, 0, sizeof(c));
    c.sc = sc;
    camera_basis(cam.eye, cam.look_at, cam.view_up, cam.fov, p.width, p.height, c.cam_eye, c.cUL, c.cdx, c.cdy);
    c.W = p.width; c.H = p.height; c.spl = p.spl; c.eye_depth = p.eye_depth; c.light_depth = p.light_depth;
    c.max_delta = p.max_delta > 0 ? p.max_delta : 256;
    c.sample_offset = p.sample_offset; c.sample_stride = p.sample_stride > 0 ? p.sample_stride : 1;
    c.key = make_uint2((uint32_t) p.seed, (uint32_t) (p.seed >> 32));
    c.pool = ws.pool.p; c.n_pool = sc.n_lights * p.spl * p.light_depth;
    c.pool_active = ws.pool_active.p; c.n_pool_active = ws.ints.p + 0;
    c.max_paths = max_paths;
    c.ray_o = ws.ray_o.p; c.ray_d = ws.ray_d.p; c.thr = ws.thr.p; c.last_pos = ws.last_pos.p; c.last_n = ws.last_n.p;
    c.pix = ws.pix.p; c.sidx = ws.sidx.p; c.pstate = ws.pstate.p; c.hit = ws.hit.p;
    c.hist_pe = ws.hist_pe.p; c.hist_pdf = ws.hist_pdf.p;
    c.live = ws.live_a.p; c.live_next = ws.live_b.p; c.n_live_next = ws.ints.p + 1;
    c.cv = ws.cv.p; c.acc = ws.acc.p; c.film = ws.film.p; c.counters = ws.counters.p;
}

// Optional per-stage device timing: one CUDA event pair per launch, summed per stage after each wave's sync.
struct StageTimer {
    enum { GEN, CLOSEST, SHADE, CONNECT, FLUSH, N };
    bool on = false;
    cudaStream_t st = nullptr;
    std::vector<cudaEvent_t> pool;
    struct Rec { int stage; cudaEvent_t a, b; };
    std::vector<Rec> pending;
    float ms[N] = { 0, 0, 0, 0, 0 };
    cudaEvent_t get() { if (pool.empty()) { cudaEvent_t e; cudaEventCreate(&e); return e; } cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
    void begin(int stage) { if (!on) return; Rec r{ stage, get(), get() }; cudaEventRecord(r.a, st); pending.push_back(r); }
    void end() { if (!on) return; cudaEventRecord(pending.back().b, st); }
    void collect() {             // call after a stream synchronize
        for (Rec &r : pending) { float t = 0.f; if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess) ms[r.stage] += t; pool.push_back(r.a); pool.push_back(r.b); }
        pending.clear();
    }
    ~StageTimer() { for (cudaEvent_t e : pool) cudaEventDestroy(e); for (Rec &r : pending) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); } }
};

static size_t stack_bytes(int threads, int bvh_depth) { return (size_t) threads * (size_t) std::max(bvh_depth + 1, 2) * sizeof(uint2); }

// Light pool generation (shared by render and b2_light_pool).  Leaves *n_pool_active in ws.ints[0].
static int build_pool(RenderCtx &c, int bvh_depth, bool count, cudaStream_t st, uint64_t &n_launches) {
    if (c.n_pool <= 0) { B2_CUDA(cudaMemsetAsync(c.n_pool_active, 0, sizeof(int), st)); return B2_OK; }
    B2_CUDA(cudaMemsetAsync(c.pool, 0, sizeof(float4) * B2_LV_F4 * (size_t) c.n_pool, st));    // value-initialised pool, cpu_bdpt.cpp:203
    const int n_paths = c.sc.n_lights * c.spl;
    const size_t sb = stack_bytes(TRAV_THREADS, bvh_depth);
    if (count) k_light_paths<true><<<blocks_for(n_paths, TRAV_THREADS), TRAV_THREADS, sb, st>>>(c, n_paths);
    else k_light_paths<false><<<blocks_for(n_paths, TRAV_THREADS), TRAV_THREADS, sb, st>>>(c, n_paths);
    k_pool_active<<<1, 1024, 0, st>>>(c);
    n_launches += 2;
    B2_CUDA(cudaGetLastError());
    return B2_OK;
}

static int set_smem_limits(int bvh_depth) {
    static int configured_depth = -1;
    if (bvh_depth <= configured_depth) return B2_OK;
    const size_t sb_t = stack_bytes(TRAV_THREADS, bvh_depth), sb_c = stack_bytes(CONN_THREADS, bvh_depth);
    if (sb_c + sizeof(uint32_t) * CONN_QCAP + sizeof(float) * 3 * CONN_THREADS + 64 > 227 * 1024) {
        b2_set_error("BVH depth %d needs more shared memory for traversal stacks than a CTA has", bvh_depth); return B2_ERR_STATE;
    }
    B2_CUDA(cudaFuncSetAttribute(k_light_paths<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_t));
    B2_CUDA(cudaFuncSetAttribute(k_light_paths<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_t));
    B2_CUDA(cudaFuncSetAttribute(k_closest<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_t));
    B2_CUDA(cudaFuncSetAttribute(k_closest<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_t));
    B2_CUDA(cudaFuncSetAttribute(k_connect<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_c));
    B2_CUDA(cudaFuncSetAttribute(k_connect<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb_c));
    configured_depth = bvh_depth;
    return B2_OK;
}

int b2_render_device_impl(const DevScene &sc, int bvh_depth, RenderWorkspace &ws, const B2Camera &cam, const B2RenderParams &p,
                          float *d_film_out, cudaStream_t st, B2Stats *stats) {
    B2_TRY(validate_params(p));
    const int npix = p.width * p.height;
    const size_t film_floats = (size_t) 3 * npix;
    const int stride = p.sample_stride > 0 ? p.sample_stride : 1;
    int n_my = 0;
    for (long long s = p.sample_offset; s < p.spp; s += stride) n_my++;
    const long long total = (long long) npix * n_my;
    long long mp = p.max_batch_paths > 0 ? p.max_batch_paths : (1ll << 22);
    mp = std::max<long long>(std::min<long long>(mp, std::max<long long>(total, 1)), 1);
    const int max_paths = (int) mp;
    const int n_pool = sc.n_lights * p.spl * p.light_depth;
    B2_TRY(ws.ensure(max_paths, std::max(p.eye_depth, 1), std::max(n_pool, 1), film_floats));
    B2_TRY(set_smem_limits(bvh_depth));
    RenderCtx c; setup_ctx(c, sc, cam, p, ws, max_paths);
    const bool count = p.count_traversal != 0;
    uint64_t n_launches = 0, n_batches = 0, n_waves = 0;
    StageTimer tm; tm.on = p.profile_stages != 0; tm.st = st;
    cudaEvent_t ev0, ev1, ev2;
    B2_CUDA(cudaEventCreate(&ev0)); B2_CUDA(cudaEventCreate(&ev1)); B2_CUDA(cudaEventCreate(&ev2));
    B2_CUDA(cudaEventRecord(ev0, st));
    B2_CUDA(cudaMemsetAsync(ws.film.p, 0, sizeof(float) * film_floats, st));
    B2_CUDA(cudaMemsetAsync(ws.counters.p, 0, sizeof(unsigned long long) * B2_CNT_N, st));
    B2_CUDA(cudaMemsetAsync(ws.ints.p, 0, sizeof(int) * 8, st));
    unsigned long long eye_samples = 0;
    // run_cpu_bdpt returns before touching the film when there are no lights (cpu_bdpt.cpp:178); here the film is zero
    if (sc.n_lights > 0 && total > 0) {
        B2_TRY(build_pool(c, bvh_depth, count, st, n_launches));
        B2_CUDA(cudaEventRecord(ev1, st));
        const size_t sb_t = stack_bytes(TRAV_THREADS, bvh_depth), sb_c = stack_bytes(CONN_THREADS, bvh_depth);
        for (long long w0 = 0; w0 < total; w0 += max_paths) {
            int n_live = (int) std::min<long long>(max_paths, total - w0);
            const int n_batch = n_live;
            c.live = ws.live_a.p; c.live_next = ws.live_b.p;
            tm.begin(StageTimer::GEN);
            k_eye_gen<<<blocks_for(n_live, 256), 256, 0, st>>>(c, n_live, w0);
            tm.end();
            n_launches++; n_batches++;
            eye_samples += (unsigned long long) n_live;
            while (n_live > 0 && p.eye_depth > 0) {
                B2_CUDA(cudaMemsetAsync(c.n_live_next, 0, sizeof(int), st));
                tm.begin(StageTimer::CLOSEST);
                if (count) k_closest<true><<<blocks_for(n_live, TRAV_THREADS), TRAV_THREADS, sb_t, st>>>(c, n_live);
                else k_closest<false><<<blocks_for(n_live, TRAV_THREADS), TRAV_THREADS, sb_t, st>>>(c, n_live);
                tm.end();
                tm.begin(StageTimer::SHADE);
                k_shade<<<blocks_for(n_live, SHADE_THREADS), SHADE_THREADS, 0, st>>>(c, n_live);
                tm.end();
                B2_CUDA(cudaMemcpyAsync(ws.h_ints, c.n_live_next, sizeof(int), cudaMemcpyDeviceToHost, st));
                tm.begin(StageTimer::CONNECT);
                if (count) k_connect<true><<<blocks_for(n_live, CONN_THREADS), CONN_THREADS, sb_c, st>>>(c, n_live);
                else k_connect<false><<<blocks_for(n_live, CONN_THREADS), CONN_THREADS, sb_c, st>>>(c, n_live);
                tm.end();
                n_launches += 3; n_waves++;
                B2_CUDA(cudaStreamSynchronize(st));
                B2_CUDA(cudaGetLastError());
                tm.collect();
                ws.closest_rays_host += (unsigned long long) n_live;
                n_live = ws.h_ints[0];
                std::swap(c.live, c.live_next);
            }
            tm.begin(StageTimer::FLUSH);
            k_flush<<<blocks_for(n_batch, 256), 256, 0, st>>>(c, n_batch);
            tm.end();
            n_launches++;
        }
    } else {
        B2_CUDA(cudaEventRecord(ev1, st));
    }
    if (p.film_mode == B2_FILM_SUM) k_resolve<<<blocks_for(film_floats, 256), 256, 0, st>>>(ws.film.p, d_film_out, film_floats, 1.0f);
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
        stats->eye_samples = eye_samples;
        stats->closest_rays = h_cnt[B2_CNT_CLOSEST] + ws.closest_rays_host;
        stats->shadow_rays = h_cnt[B2_CNT_SHADOW];
        stats->path_samples = h_cnt[B2_CNT_PATH_SAMPLES];
        stats->delta_cap_hits = h_cnt[B2_CNT_DELTA_CAP];
        stats->nodes_visited = h_cnt[B2_CNT_NODES];
        stats->tris_tested = h_cnt[B2_CNT_TRIS];
        stats->n_batches = n_batches; stats->n_waves = n_waves; stats->n_launches = n_launches;
        stats->ms_total = ms_total; stats->ms_light_pool = ms_pool;
        stats->pairs_tested = h_cnt[B2_CNT_PAIRS];
        tm.collect();
        stats->ms_gen = tm.ms[StageTimer::GEN]; stats->ms_closest = tm.ms[StageTimer::CLOSEST]; stats->ms_shade = tm.ms[StageTimer::SHADE];
        stats->ms_connect = tm.ms[StageTimer::CONNECT]; stats->ms_flush = tm.ms[StageTimer::FLUSH];
    }
    ws.closest_rays_host = 0;
    cudaEventDestroy(ev0); cudaEventDestroy(ev1); cudaEventDestroy(ev2);
    return B2_OK;
}

int b2_light_pool_impl(const DevScene &sc, int bvh_depth, RenderWorkspace &ws, const B2RenderParams &p, float *pool20_host, cudaStream_t st) {
    if (p.spl < 0 || p.light_depth <= 0) { b2_set_error("invalid spl / light_depth"); return B2_ERR_INVALID; }
    const int n_pool = sc.n_lights * p.spl * p.light_depth;
    if (n_pool == 0) return B2_OK;
    B2_TRY(ws.ensure(1, 1, n_pool, 3));
    B2_TRY(set_smem_limits(bvh_depth));
    B2Camera cam; memset(&cam, 0, sizeof(cam)); cam.look_at[2] = 1.f; cam.view_up[1] = 1.f; cam.fov = 50.f;
    B2RenderParams pp = p; pp.width = pp.height = 1;
    RenderCtx c; setup_ctx(c, sc, cam, pp, ws, 1);
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
