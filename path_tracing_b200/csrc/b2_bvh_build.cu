// b2_bvh_build.cu -- GPU BVH build: kernels wrapping the bodies of b2_bvh_build.cuh plus the level loop.
// Sorting and the two per-level exclusive scans use CUB (shipped with CUDA; plain library primitives off
// the hot path).  See the header for the pipeline.
#include <cub/cub.cuh>
#include <mutex>
#include "b2_bvh_build.cuh"
#include "b2_cuda_util.hpp"
#include "b2_device_scene.hpp"

namespace {

__device__ __forceinline__ uint32_t f2ord(float f) { uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__host__ __device__ __forceinline__ float ord2f(uint32_t u) {
    uint32_t b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    float f; memcpy(&f, &b, 4); return f;
}

// 1. padded triangle boxes + centroid bounds (warp shuffle reduction, then one atomic per warp and axis)
__global__ void k_tri_bounds(const BuildTri *__restrict__ in, int n, float pad, float *__restrict__ tlo, float *__restrict__ thi, uint32_t *cbounds) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    float c[3] = { 3e38f, 3e38f, 3e38f }, C[3] = { -3e38f, -3e38f, -3e38f };
    if (i < n) {
        float lo[3], hi[3];
        tri_bounds(in[i], pad, lo, hi);
        for (int a = 0; a < 3; ++a) { tlo[3 * i + a] = lo[a]; thi[3 * i + a] = hi[a]; c[a] = C[a] = 0.5f * (lo[a] + hi[a]); }
    }
    for (int a = 0; a < 3; ++a) {
        for (int o = 16; o > 0; o >>= 1) { c[a] = fminf(c[a], __shfl_xor_sync(0xffffffffu, c[a], o)); C[a] = fmaxf(C[a], __shfl_xor_sync(0xffffffffu, C[a], o)); }
        if ((threadIdx.x & 31) == 0) { atomicMin(&cbounds[a], f2ord(c[a])); atomicMax(&cbounds[3 + a], f2ord(C[a])); }
    }
}
// ---- small scenes: PLOC passes and the collapse of one build in ONE CTA (no host round trip per pass / level) ----
// Same bodies, same order of decisions, hence the same tree as the multi-launch path (tests/test_bvh_builders.py compares
// both with the host mirror).  The 48-triangle scene of the C2 bench took 0.73 ms of host time for ~14 synchronisations.
constexpr int SMALL_N = 1024;
struct SmallBuildInfo { int n_nodes, depth, passes, tris_done, error; };
__global__ void __launch_bounds__(SMALL_N) k_build_small(BinTree bt, const float *__restrict__ tlo, const float *__restrict__ thi, int radius, int max_leaf,
                                                          int *cid_a, int *cid_b, int *nn, CollapseItem *items_a, CollapseItem *items_b, CollapseOut *outs,
                                                          uint4 *nodes, int *perm, SmallBuildInfo *info) {
    typedef cub::BlockScan<unsigned long long, SMALL_N> Scan64;
    typedef cub::BlockScan<int, SMALL_N> Scan32;
    __shared__ union { typename Scan64::TempStorage s64; typename Scan32::TempStorage s32; } tmp;
    const int i = threadIdx.x, n = bt.n;
    int passes = 0, error = 0;
    if (n >= 2) {
        if (i < n) { ploc_leaf(bt, tlo, thi, i); cid_a[i] = (n - 1) + i; }
        __syncthreads();
        int *cur = cid_a, *nxt = cid_b;
        int m = n, created = 0;
        while (m > 1) {
            if (i < m) nn[i] = ploc_nearest(bt, cur, m, i, radius);
            __syncthreads();
            int r = 2;
            unsigned long long f = 0ull, ex = 0ull, total = 0ull;
            if (i < m) { r = ploc_role(nn, i); f = (r == 1 ? 1ull : 0ull) | (r != 2 ? 1ull << 32 : 0ull); }
            Scan64(tmp.s64).ExclusiveSum(f, ex, total);
            if (i < m && r != 2) {
                const int pos = (int) (ex >> 32);
                if (r == 1) {
                    const int node = n - 2 - (created + (int) (uint32_t) ex);
                    ploc_merge(bt, cur[i], cur[nn[i]], node);
                    nxt[pos] = node;
                } else nxt[pos] = cur[i];
            }
            __syncthreads();
            const int merged = (int) (uint32_t) total, alive = (int) (total >> 32);
            if (merged <= 0 || alive != m - merged) { error = 1; break; }              // uniform: `total` is the block aggregate
            created += merged; m = alive;
            int *t = cur; cur = nxt; nxt = t;
            passes++;
        }
    } else if (i == 0) {
        for (int a = 0; a < 3; ++a) { bt.lo[a] = tlo[a]; bt.hi[a] = thi[a]; }
    }
    __syncthreads();
    CollapseItem *cur = items_a, *nxt = items_b;
    if (i == 0) cur[0].bin_node = 0;
    __syncthreads();
    int m = error ? 0 : 1, depth = 0;
    uint32_t level_start = 0, tris_done = 0;
    while (m > 0) {
        if (m > SMALL_N) { error = 2; break; }
        int ni = 0, nt = 0, io = 0, to = 0, itot = 0, ttot = 0;
        if (i < m) collapse_pick(bt, cur[i].bin_node, max_leaf, outs[i], ni, nt);
        Scan32(tmp.s32).ExclusiveSum(ni, io, itot);
        __syncthreads();
        Scan32(tmp.s32).ExclusiveSum(nt, to, ttot);
        const uint32_t next_start = level_start + (uint32_t) m;
        if (i < m) collapse_emit(bt, cur[i].bin_node, outs[i], max_leaf, next_start + (uint32_t) io, tris_done + (uint32_t) to,
                                 nodes + (size_t) (level_start + i) * B2_NODE_UINT4, nxt, io, perm);
        __syncthreads();
        level_start = next_start; tris_done += (uint32_t) ttot;
        CollapseItem *t = cur; cur = nxt; nxt = t;
        m = itot;
        depth++;
    }
    if (i == 0) { info->n_nodes = (int) level_start; info->depth = depth; info->passes = passes; info->tris_done = (int) tris_done; info->error = error; }
}
// Morton keys with the centroid bounds read on the device (k_tri_bounds left them in cb as ordered uints)
__global__ void k_morton_dev(int n, const float *__restrict__ tlo, const float *__restrict__ thi, const uint32_t *__restrict__ cb, uint64_t *keys, int *vals) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float cmn[3], cmx[3];
    for (int a = 0; a < 3; ++a) { cmn[a] = ord2f(cb[a]); cmx[a] = ord2f(cb[3 + a]); }
    const float3 cmin = make_float3(cmn[0], cmn[1], cmn[2]);
    const float3 cinv = make_float3(__fdiv_rn(2097152.0f, fmaxf(__fsub_rn(cmx[0], cmn[0]), 1e-30f)), __fdiv_rn(2097152.0f, fmaxf(__fsub_rn(cmx[1], cmn[1]), 1e-30f)),
                                    __fdiv_rn(2097152.0f, fmaxf(__fsub_rn(cmx[2], cmn[2]), 1e-30f)));
    keys[i] = morton63(&tlo[3 * i], &thi[3 * i], cmin, cinv); vals[i] = i;
}
__global__ void k_karras(BinTree bt) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < bt.n - 1) karras_node(bt, i);
}
__global__ void k_refit(BinTree bt, const float *__restrict__ tlo, const float *__restrict__ thi) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < bt.n) refit_from_leaf(bt, tlo, thi, j);
}
__global__ void k_single_leaf_box(BinTree bt, const float *tlo, const float *thi) {
    for (int a = 0; a < 3; ++a) { bt.lo[a] = tlo[a]; bt.hi[a] = thi[a]; }
}
// ---- PLOC passes (bodies in b2_bvh_build.cuh) ----
__global__ void k_ploc_leaves(BinTree bt, const float *__restrict__ tlo, const float *__restrict__ thi, int *cid) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < bt.n) { ploc_leaf(bt, tlo, thi, j); cid[j] = (bt.n - 1) + j; }
}
__global__ void k_ploc_nearest(BinTree bt, const int *__restrict__ cid, int m, int radius, int *nn) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) nn[i] = ploc_nearest(bt, cid, m, i, radius);
}
// flags: low word = 1 for a left partner (one new node), high word = 1 for a cluster that survives the pass
__global__ void k_ploc_flags(const int *__restrict__ nn, int m, unsigned long long *flags) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) { const int r = ploc_role(nn, i); flags[i] = (r == 1 ? 1ull : 0ull) | (r != 2 ? 1ull << 32 : 0ull); }
}
__global__ void k_ploc_apply(BinTree bt, const int *__restrict__ cid, const int *__restrict__ nn, const unsigned long long *__restrict__ scan, int m,
                             int created, int *cid_out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const int r = ploc_role(nn, i);
    if (r == 2) return;
    const int pos = (int) (scan[i] >> 32);
    if (r == 1) {
        const int node = bt.n - 2 - (created + (int) (uint32_t) scan[i]);      // the last merge of the build creates node 0, the root
        ploc_merge(bt, cid[i], cid[nn[i]], node);
        cid_out[pos] = node;
    } else cid_out[pos] = cid[i];
}
__global__ void k_collapse_pick(BinTree bt, const CollapseItem *__restrict__ items, int m, int max_leaf, CollapseOut *outs, int *nint, int *ntri) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < m) collapse_pick(bt, items[k].bin_node, max_leaf, outs[k], nint[k], ntri[k]);
}
__global__ void k_collapse_emit(BinTree bt, const CollapseItem *__restrict__ items, const CollapseOut *__restrict__ outs, int m, int max_leaf,
                                const int *__restrict__ ioff, const int *__restrict__ toff, uint32_t level_start, uint32_t next_start, uint32_t tris_done,
                                uint4 *nodes, CollapseItem *next_items, int *perm) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < m)
        collapse_emit(bt, items[k].bin_node, outs[k], max_leaf, next_start + (uint32_t) ioff[k], tris_done + (uint32_t) toff[k],
                      nodes + (size_t) (level_start + k) * B2_NODE_UINT4, next_items, ioff[k], perm);
}
__global__ void k_pack_tris(const BuildTri *__restrict__ in, const int *__restrict__ sorted, const int *__restrict__ perm, int n, float4 *tri, int32_t *obj) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n) pack_tri(in, sorted, perm, s, tri, obj);
}

// Build temporaries come out of ONE device allocation (bump-allocated; kept per device between builds while it is small):
// two dozen cudaMalloc / cudaFree pairs cost more than the kernels of a small build (48 triangles: 0.9 ms per build, most of
// it allocation), and the end-to-end step of the bench creates and builds its scene every time.
struct BuildArena {
    char *base = nullptr; size_t cap = 0, off = 0; int dev = 0;
    static constexpr size_t KEEP = (size_t) 256 << 20;
    static std::mutex &mtx() { static std::mutex m; return m; }
    static BuildArena *cache() { static BuildArena c[64]; return c; }
    int acquire(size_t need) {
        cudaGetDevice(&dev);
        {
            std::lock_guard<std::mutex> lk(mtx());
            BuildArena &c = cache()[dev & 63];
            if (c.base) { base = c.base; cap = c.cap; c.base = nullptr; c.cap = 0; }
        }
        if (cap < need) {
            if (base) cudaFree(base);
            base = nullptr; cap = 0;
            cudaError_t e = cudaMalloc((void **) &base, need);
            if (e != cudaSuccess) { base = nullptr; b2_set_error("cudaMalloc(%zu bytes) for the BVH build failed: %s", need, cudaGetErrorString(e)); return B2_ERR_OOM; }
            cap = need;
        }
        off = 0;
        return B2_OK;
    }
    void *take(size_t bytes) {
        const size_t o = (off + 255) & ~(size_t) 255;
        if (!base || o + bytes > cap) return nullptr;
        off = o + bytes;
        return base + o;
    }
    ~BuildArena() {
        if (!base) return;
        if (cap <= KEEP) {
            std::lock_guard<std::mutex> lk(mtx());
            BuildArena &c = cache()[dev & 63];
            if (!c.base) { c.base = base; c.cap = cap; base = nullptr; }
        }
        if (base) cudaFree(base);
        base = nullptr;
    }
};
thread_local BuildArena *t_arena = nullptr;
// DevBuf's interface over the arena (falls back to its own cudaMalloc if the arena is exhausted)
template <class T>
struct ArenaBuf {
    T *p = nullptr; size_t n = 0; bool own = false;
    ArenaBuf() = default;
    ArenaBuf(const ArenaBuf &) = delete;
    ArenaBuf &operator=(const ArenaBuf &) = delete;
    ~ArenaBuf() { release(); }
    void release() { if (own && p) cudaFree(p); p = nullptr; n = 0; own = false; }
    int alloc(size_t count) {
        release();
        if (count == 0) count = 1;
        if (t_arena) p = (T *) t_arena->take(count * sizeof(T));
        if (!p) {
            cudaError_t e = cudaMalloc((void **) &p, count * sizeof(T));
            if (e != cudaSuccess) { p = nullptr; b2_set_error("cudaMalloc(%zu bytes) failed: %s", count * sizeof(T), cudaGetErrorString(e)); return B2_ERR_OOM; }
            own = true;
        }
        n = count;
        return B2_OK;
    }
    int ensure(size_t count) { return (count <= n && p) ? B2_OK : alloc(count); }
};

}  // namespace

// SAH cost of the chosen wide tree, root area normalised, in units of one wide-node test (B2_SAH_C_NODE = 1): the root's
// C_internal of the cost tables (exact boxes).  One 40-byte read.
static int read_sah_cost(const BinTree &bt, cudaStream_t st, float &out) {
    out = 0.0f;
    if (bt.n < 2 || !bt.sah) return B2_OK;
    float c = 0.f, lo[3], hi[3];
    B2_CUDA(cudaMemcpyAsync(&c, bt.sah, sizeof(float), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaMemcpyAsync(lo, bt.lo, sizeof(lo), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaMemcpyAsync(hi, bt.hi, sizeof(hi), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2], a = dx * dy + dy * dz + dz * dx;
    out = a > 0.0f ? c / a : 0.0f;
    return B2_OK;
}

int b2_build_bvh_device(const BuildTri *d_in, int n, float pad, int max_leaf, int builder, int ploc_radius, cudaStream_t st, DeviceBvh &out) {
    out.release();
    if (n <= 0) return B2_OK;
    cudaEvent_t ev0, ev1;
    B2_CUDA(cudaEventCreate(&ev0)); B2_CUDA(cudaEventCreate(&ev1));
    B2_CUDA(cudaEventRecord(ev0, st));
    const unsigned T = 256;
    size_t sort_bytes = 0;
    B2_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (uint64_t *) nullptr, (uint64_t *) nullptr, (int *) nullptr, (int *) nullptr, n, 0, 63, st));
    BuildArena arena;                                    // ~410 B of temporaries per triangle, all live at once
    B2_TRY(arena.acquire((size_t) 500 * (size_t) n + sort_bytes + ((size_t) 1 << 20)));
    struct ArenaScope { ArenaScope(BuildArena *a) { t_arena = a; } ~ArenaScope() { t_arena = nullptr; } } arena_scope(&arena);
    ArenaBuf<float> tlo, thi, lo, hi;
    ArenaBuf<uint32_t> cb, visit;
    ArenaBuf<uint64_t> keys, skeys;
    ArenaBuf<int> vals, sorted, left, right, parent, first, last, leaf_parent, perm, nint, ntri, ioff, toff, count, cid_a, cid_b, nn;
    ArenaBuf<unsigned long long> pflags, pscan;
    ArenaBuf<unsigned char> tmp;
    ArenaBuf<CollapseItem> items_a, items_b;
    ArenaBuf<CollapseOut> outs;
    ArenaBuf<uint4> nodes_tmp;
    B2_TRY(tlo.alloc(3 * (size_t) n)); B2_TRY(thi.alloc(3 * (size_t) n));
    B2_TRY(cb.alloc(6));
    uint32_t cb_init[6] = { 0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u };
    B2_CUDA(cudaMemcpyAsync(cb.p, cb_init, sizeof(cb_init), cudaMemcpyHostToDevice, st));
    k_tri_bounds<<<blocks_for(n, T), T, 0, st>>>(d_in, n, pad, tlo.p, thi.p, cb.p);
    B2_TRY(keys.alloc(n)); B2_TRY(skeys.alloc(n)); B2_TRY(vals.alloc(n)); B2_TRY(sorted.alloc(n));
    k_morton_dev<<<blocks_for(n, T), T, 0, st>>>(n, tlo.p, thi.p, cb.p, keys.p, vals.p);      // centroid bounds stay on the device: no host read here
    size_t tmp_bytes = 0;
    B2_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys.p, skeys.p, vals.p, sorted.p, n, 0, 63, st));
    B2_TRY(tmp.alloc(tmp_bytes));
    B2_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, keys.p, skeys.p, vals.p, sorted.p, n, 0, 63, st));
    keys.release(); vals.release();

    B2_TRY(left.alloc(n)); B2_TRY(right.alloc(n));
    B2_TRY(lo.alloc(6 * (size_t) n)); B2_TRY(hi.alloc(6 * (size_t) n));
    int ploc_passes = 0;
    ArenaBuf<float> sah; ArenaBuf<unsigned char> dec;           // cost tables of the SAH-optimal collapse: 8 floats + 8 bytes per binary node
    B2_TRY(sah.alloc(16 * (size_t) n)); B2_TRY(dec.alloc(16 * (size_t) n));
    BinTree bt{ n, skeys.p, sorted.p, left.p, right.p, nullptr, nullptr, nullptr, nullptr, lo.p, hi.p, nullptr, nullptr, sah.p, dec.p, max_leaf };
    const bool small = n <= SMALL_N && (builder != B2_BUILDER_LBVH || n < 2);
    if (small) {
        // one CTA runs every PLOC pass and every collapse level (k_build_small); one host read for the whole build
        B2_TRY(count.alloc(2 * (size_t) n)); B2_TRY(cid_a.alloc(n)); B2_TRY(cid_b.alloc(n)); B2_TRY(nn.alloc(n));
        bt.count = count.p;
        B2_TRY(perm.alloc(n));
        const size_t node_cap = (size_t) n + 1;
        B2_TRY(nodes_tmp.alloc(node_cap * B2_NODE_UINT4));
        B2_TRY(items_a.alloc(node_cap)); B2_TRY(items_b.alloc(node_cap)); B2_TRY(outs.alloc(node_cap));
        ArenaBuf<SmallBuildInfo> info;
        B2_TRY(info.alloc(1));
        k_build_small<<<1, SMALL_N, 0, st>>>(bt, tlo.p, thi.p, ploc_radius, max_leaf, cid_a.p, cid_b.p, nn.p, items_a.p, items_b.p, outs.p, nodes_tmp.p, perm.p, info.p);
        B2_CUDA(cudaGetLastError());
        SmallBuildInfo hi_;
        B2_CUDA(cudaMemcpyAsync(&hi_, info.p, sizeof(hi_), cudaMemcpyDeviceToHost, st));
        B2_CUDA(cudaStreamSynchronize(st));
        if (hi_.error) { b2_set_error("small BVH build failed (internal error %d)", hi_.error); return B2_ERR_STATE; }
        if (hi_.tris_done != n) { b2_set_error("BVH collapse placed %d of %d triangles (internal error)", hi_.tris_done, n); return B2_ERR_STATE; }
        if (hi_.depth + 2 > B2_STACK_CAP) { b2_set_error("BVH depth %d exceeds the traversal stack (%d)", hi_.depth, B2_STACK_CAP); return B2_ERR_STATE; }
        const size_t n_nodes = (size_t) hi_.n_nodes;
        B2_TRY(out.nodes.alloc(n_nodes * B2_NODE_UINT4));
        B2_CUDA(cudaMemcpyAsync(out.nodes.p, nodes_tmp.p, n_nodes * B2_NODE_UINT4 * sizeof(uint4), cudaMemcpyDeviceToDevice, st));
        B2_TRY(out.tri.alloc(3 * (size_t) n)); B2_TRY(out.tri_obj.alloc(n));
        k_pack_tris<<<blocks_for(n, T), T, 0, st>>>(d_in, sorted.p, perm.p, n, out.tri.p, out.tri_obj.p);
        B2_CUDA(cudaGetLastError());
        B2_CUDA(cudaEventRecord(ev1, st));
        B2_CUDA(cudaStreamSynchronize(st));
        float ms = 0.f;
        cudaEventElapsedTime(&ms, ev0, ev1);
        cudaEventDestroy(ev0); cudaEventDestroy(ev1);
        out.n_tris = n; out.n_nodes = (int) n_nodes; out.depth = hi_.depth; out.build_ms = ms; out.build_passes = hi_.passes;
        B2_TRY(read_sah_cost(bt, st, out.sah_cost));
        return B2_OK;
    }
    if (n < 2) {
        k_single_leaf_box<<<1, 1, 0, st>>>(bt, tlo.p, thi.p);
    } else if (builder == B2_BUILDER_LBVH) {
        B2_TRY(parent.alloc(n)); B2_TRY(first.alloc(n)); B2_TRY(last.alloc(n)); B2_TRY(leaf_parent.alloc(n)); B2_TRY(visit.alloc(n));
        B2_CUDA(cudaMemsetAsync(visit.p, 0, sizeof(uint32_t) * n, st));
        B2_CUDA(cudaMemsetAsync(leaf_parent.p, 0xff, sizeof(int) * n, st));
        bt.parent = parent.p; bt.first = first.p; bt.last = last.p; bt.leaf_parent = leaf_parent.p; bt.visit = visit.p;
        k_karras<<<blocks_for(n - 1, T), T, 0, st>>>(bt);
        k_refit<<<blocks_for(n, T), T, 0, st>>>(bt, tlo.p, thi.p);
    } else {
        // PLOC: nearest neighbour within `radius` positions of the Morton order, mutual pairs merge, survivors are compacted in
        // order (one 64-bit exclusive scan gives both the new node index and the new position); one host read per pass
        B2_TRY(count.alloc(2 * (size_t) n)); B2_TRY(cid_a.alloc(n)); B2_TRY(cid_b.alloc(n)); B2_TRY(nn.alloc(n)); B2_TRY(pflags.alloc(n)); B2_TRY(pscan.alloc(n));
        bt.count = count.p;
        size_t pscan_bytes = 0;
        B2_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, pscan_bytes, pflags.p, pscan.p, n, st));
        B2_TRY(tmp.ensure(pscan_bytes));
        k_ploc_leaves<<<blocks_for(n, T), T, 0, st>>>(bt, tlo.p, thi.p, cid_a.p);
        int *cur_c = cid_a.p, *nxt_c = cid_b.p;
        int m = n, created = 0;
        while (m > 1) {
            k_ploc_nearest<<<blocks_for(m, 128), 128, 0, st>>>(bt, cur_c, m, ploc_radius, nn.p);
            k_ploc_flags<<<blocks_for(m, T), T, 0, st>>>(nn.p, m, pflags.p);
            B2_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, pscan_bytes, pflags.p, pscan.p, m, st));
            unsigned long long tail[2];
            B2_CUDA(cudaMemcpyAsync(&tail[0], pflags.p + (m - 1), sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
            B2_CUDA(cudaMemcpyAsync(&tail[1], pscan.p + (m - 1), sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
            k_ploc_apply<<<blocks_for(m, T), T, 0, st>>>(bt, cur_c, nn.p, pscan.p, m, created, nxt_c);
            B2_CUDA(cudaStreamSynchronize(st));
            const unsigned long long tot = tail[0] + tail[1];
            const int merged = (int) (uint32_t) tot, alive = (int) (tot >> 32);
            if (merged <= 0 || alive != m - merged) { b2_set_error("PLOC pass made no progress (internal error: m %d merged %d alive %d)", m, merged, alive); return B2_ERR_STATE; }
            created += merged; m = alive;
            int *t = cur_c; cur_c = nxt_c; nxt_c = t;
            ploc_passes++;
        }
    }
    B2_CUDA(cudaGetLastError());

    // level-synchronous collapse
    B2_TRY(perm.alloc(n));
    size_t node_cap = (size_t) n + 1;
    B2_TRY(nodes_tmp.alloc(node_cap * B2_NODE_UINT4));
    B2_TRY(items_a.alloc(node_cap)); B2_TRY(items_b.alloc(node_cap)); B2_TRY(outs.alloc(node_cap));
    B2_TRY(nint.alloc(node_cap)); B2_TRY(ntri.alloc(node_cap)); B2_TRY(ioff.alloc(node_cap)); B2_TRY(toff.alloc(node_cap));
    size_t scan_bytes = 0;
    B2_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, nint.p, ioff.p, (int) node_cap, st));
    B2_TRY(tmp.ensure(scan_bytes));
    CollapseItem root_item; root_item.bin_node = 0;
    B2_CUDA(cudaMemcpyAsync(items_a.p, &root_item, sizeof(root_item), cudaMemcpyHostToDevice, st));
    CollapseItem *cur = items_a.p, *nxt = items_b.p;
    int m = 1, depth = 0;
    uint32_t level_start = 0, tris_done = 0;
    while (m > 0) {
        if ((size_t) level_start + m > node_cap) { b2_set_error("BVH collapse overflow (internal error)"); return B2_ERR_STATE; }
        k_collapse_pick<<<blocks_for(m, 128), 128, 0, st>>>(bt, cur, m, max_leaf, outs.p, nint.p, ntri.p);
        B2_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, scan_bytes, nint.p, ioff.p, m, st));
        B2_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, scan_bytes, ntri.p, toff.p, m, st));
        int tail[4];
        B2_CUDA(cudaMemcpyAsync(&tail[0], nint.p + (m - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
        B2_CUDA(cudaMemcpyAsync(&tail[1], ioff.p + (m - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
        B2_CUDA(cudaMemcpyAsync(&tail[2], ntri.p + (m - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
        B2_CUDA(cudaMemcpyAsync(&tail[3], toff.p + (m - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
        B2_CUDA(cudaStreamSynchronize(st));
        const int itot = tail[0] + tail[1], ttot = tail[2] + tail[3];
        const uint32_t next_start = level_start + (uint32_t) m;
        k_collapse_emit<<<blocks_for(m, 128), 128, 0, st>>>(bt, cur, outs.p, m, max_leaf, ioff.p, toff.p, level_start, next_start, tris_done, nodes_tmp.p, nxt, perm.p);
        B2_CUDA(cudaGetLastError());
        level_start = next_start; tris_done += (uint32_t) ttot;
        CollapseItem *t = cur; cur = nxt; nxt = t;
        m = itot;
        depth++;
    }
    if (tris_done != (uint32_t) n) { b2_set_error("BVH collapse placed %u of %d triangles (internal error)", tris_done, n); return B2_ERR_STATE; }
    if (depth + 2 > B2_STACK_CAP) { b2_set_error("BVH depth %d exceeds the traversal stack (%d)", depth, B2_STACK_CAP); return B2_ERR_STATE; }
    const size_t n_nodes = level_start;
    B2_TRY(out.nodes.alloc(n_nodes * B2_NODE_UINT4));
    B2_CUDA(cudaMemcpyAsync(out.nodes.p, nodes_tmp.p, n_nodes * B2_NODE_UINT4 * sizeof(uint4), cudaMemcpyDeviceToDevice, st));
    B2_TRY(out.tri.alloc(3 * (size_t) n)); B2_TRY(out.tri_obj.alloc(n));
    k_pack_tris<<<blocks_for(n, T), T, 0, st>>>(d_in, sorted.p, perm.p, n, out.tri.p, out.tri_obj.p);
    B2_CUDA(cudaGetLastError());
    B2_CUDA(cudaEventRecord(ev1, st));
    B2_CUDA(cudaStreamSynchronize(st));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev0, ev1);
    cudaEventDestroy(ev0); cudaEventDestroy(ev1);
    out.n_tris = n; out.n_nodes = (int) n_nodes; out.depth = depth; out.build_ms = ms; out.build_passes = ploc_passes;
    B2_TRY(read_sah_cost(bt, st, out.sah_cost));
    return B2_OK;
}
