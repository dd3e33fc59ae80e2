// b2_noalloc.cu -- b2_light_pool_device_noalloc: the light-vertex pool of a render (cpu_bdpt.cpp:202-325, the same
// light_path_body as b2_light_pool / the renderer) for callers that must not allocate: code under CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
// The entry point calls no cudaMalloc / cudaFree / allocating library routine.
//
// b2_light_pool keeps its scratch (the 80-byte pool records the subpath kernel writes, and its counters) in the scene's
// RenderWorkspace, which it allocates on demand.  Here that scratch is ONE file-scope device array of a fixed size:
// B2_NOALLOC_MAX_POOL_VERTICES records (2.6 MB) + counters.  Every GPU that loads libb2pt.so holds it until exit.
// Consequences, all part of the contract (include/b2pt.h):
//   * pools larger than the maximum are rejected with B2_ERR_INVALID and a message, never truncated
//   * every call shares the array and finds the previous call's contents in it: the part a call uses is cleared first
//     (the reference value-initialises its pool, cpu_bdpt.cpp:203), on the caller's stream
//   * two calls that overlap on the GPU (different streams, or different scene handles on one device) race on the array
// In host code `g_scratch` is only a host-side stand-in for the symbol: its device address is asked for once per device,
// with that device current (cudaGetSymbolAddress), and that pointer is what the kernels get.
#include <mutex>
#include "b2_scene_handle.hpp"
#include "b2_warp.cuh"

namespace {

constexpr int NA_THREADS = 128;
struct NoAllocScratch {
    float4 pool[(size_t) B2_LV_F4 * B2_NOALLOC_MAX_POOL_VERTICES];
    unsigned long long counters[B2_CNT_N];
};
static __device__ __align__(256) NoAllocScratch g_scratch;

__global__ void __launch_bounds__(NA_THREADS) k_light_paths_static(RenderCtx c, int n_paths) {
    extern __shared__ uint2 s_stack[];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    TravCounters tc{ 0u, 0u };
    uint32_t ncl = 0, ncap = 0;
    if (i < n_paths) light_path_body<false>(c, i, s_stack + threadIdx.x, blockDim.x, &tc, ncl, ncap);
    warp_count_add(&c.counters[B2_CNT_CLOSEST], ncl);
    warp_count_add(&c.counters[B2_CNT_DELTA_CAP], ncap);
}
// pool record -> the 20 floats per vertex b2_light_pool returns
__global__ void __launch_bounds__(NA_THREADS) k_pool_export_static(const float4 *__restrict__ pool, int n_pool, float *__restrict__ out20) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pool) return;
    const LightVtx v = load_lv(pool, i);
    float *o = out20 + 20 * (size_t) i;
    o[0] = v.pos.x; o[1] = v.pos.y; o[2] = v.pos.z; o[3] = v.n.x; o[4] = v.n.y; o[5] = v.n.z; o[6] = v.thr.x; o[7] = v.thr.y; o[8] = v.thr.z;
    o[9] = v.mtl.base.x; o[10] = v.mtl.base.y; o[11] = v.mtl.base.z; o[12] = v.mtl.roughness; o[13] = v.mtl.metallic; o[14] = v.mtl.eta;
    o[15] = (v.flags & B2_LV_SRC) ? 1.f : 0.f; o[16] = (v.flags & B2_LV_PARALLEL) ? 1.f : 0.f; o[17] = v.cutoff; o[18] = v.pdf_fwd; o[19] = v.pdf_rev;
}

// device address of g_scratch, once per device (the symbol has one instance per device that loaded the module)
std::mutex g_mtx;
NoAllocScratch *g_dev_ptr[64];
int scratch_ptr(int device, NoAllocScratch **out) {
    if (device < 0 || device >= 64) { b2_set_error("device %d out of range", device); return B2_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(g_mtx);
    if (!g_dev_ptr[device]) {
        void *p = nullptr;
        B2_CUDA(cudaGetSymbolAddress(&p, g_scratch));             // `device` is current (the caller set it)
        if (!p) { b2_set_error("cudaGetSymbolAddress returned NULL for the static scratch"); return B2_ERR_CUDA; }
        g_dev_ptr[device] = (NoAllocScratch *) p;
    }
    *out = g_dev_ptr[device];
    return B2_OK;
}

}  // namespace

extern "C" int b2_light_pool_device_noalloc(B2Scene *s, const B2RenderParams *params, float *d_pool20, void *cuda_stream) {
    B2_TRY(b2_check_built(s));
    if (!params || !d_pool20) { b2_set_error("NULL argument"); return B2_ERR_INVALID; }
    const B2RenderParams &p = *params;
    if (p.spl < 0 || p.light_depth <= 0) { b2_set_error("invalid spl / light_depth"); return B2_ERR_INVALID; }
    const long long n_pool_ll = (long long) s->hs.n_lights * p.spl * p.light_depth;
    if (n_pool_ll > B2_NOALLOC_MAX_POOL_VERTICES) {
        b2_set_error("light pool of %lld vertices (n_lights %d x spl %d x light_depth %d) exceeds the static scratch of b2_light_pool_device_noalloc "
                     "(%d vertices): use b2_light_pool", n_pool_ll, s->hs.n_lights, p.spl, p.light_depth, B2_NOALLOC_MAX_POOL_VERTICES);
        return B2_ERR_INVALID;
    }
    const int n_pool = (int) n_pool_ll;
    if (n_pool == 0) return B2_OK;
    const size_t smem = stack_bytes(NA_THREADS, s->bvh.depth);
    if (smem > 48 * 1024) { b2_set_error("BVH depth %d needs %zu bytes of traversal stack per CTA, more than this entry point supports (48 KB)", s->bvh.depth, smem); return B2_ERR_STATE; }
    std::lock_guard<std::mutex> lk(s->mtx);
    B2_CUDA(cudaSetDevice(s->device));
    NoAllocScratch *sp = nullptr;
    B2_TRY(scratch_ptr(s->device, &sp));
    cudaStream_t st = cuda_stream ? (cudaStream_t) cuda_stream : s->stream;
    RenderCtx c; memset(&c, 0, sizeof(c));
    c.sc = s->view();
    c.sc.ggx_fix = p.strategies != B2_STRATEGIES_REFERENCE;
    c.spl = p.spl; c.light_depth = p.light_depth; c.eye_depth = 1; c.W = c.H = 1;
    c.max_delta = p.max_delta > 0 ? p.max_delta : 256;
    c.key = make_uint2((uint32_t) p.seed, (uint32_t) (p.seed >> 32));
    c.pool = sp->pool; c.n_pool = n_pool; c.counters = sp->counters;
    // the array keeps the last call's records: clear what this call uses (value-initialised pool, cpu_bdpt.cpp:203)
    B2_CUDA(cudaMemsetAsync(sp->pool, 0, sizeof(float4) * B2_LV_F4 * (size_t) n_pool, st));
    B2_CUDA(cudaMemsetAsync(sp->counters, 0, sizeof(sp->counters), st));
    const int n_paths = s->hs.n_lights * p.spl;
    k_light_paths_static<<<blocks_for(n_paths, NA_THREADS), NA_THREADS, smem, st>>>(c, n_paths);
    k_pool_export_static<<<blocks_for(n_pool, NA_THREADS), NA_THREADS, 0, st>>>(sp->pool, n_pool, d_pool20);
    B2_CUDA(cudaGetLastError());
    if (!cuda_stream) B2_CUDA(cudaStreamSynchronize(st));      // own stream: blocking; caller's stream: asynchronous on it
    return B2_OK;
}
