// b2_multi.cu -- b2_render_bdpt_multi: one node, G GPUs, one host thread per GPU inside ONE process (C++ callers such
// as pt_cli --gpus G).  Scene, BVH and light pool are replicated (one B2Scene per device, same inputs -> identical
// deterministic builds and identical Philox streams); GPU g renders the sample indices s = g (mod G) of every pixel
// into a raw-sum film; the films are summed onto GPU 0 with one NCCL reduce over NVLink and resolved there.
// NCCL is loaded with dlopen at the first multi-GPU call, so single-GPU users of libb2pt.so need no NCCL at all.
// (Python callers run one process per GPU under torchrun and use torch.distributed instead: path_tracing_b200/multi.py.)
#include <dlfcn.h>
#include <algorithm>
#include <cstring>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>
#include "b2_scene_handle.hpp"

namespace {

// minimal NCCL surface (ABI-stable since NCCL 2.x): opaque communicator, enums as ints
typedef struct ncclComm *ncclComm_t;
enum { kNcclSuccess = 0, kNcclFloat32 = 7, kNcclSum = 0 };
struct Nccl {
    void *h = nullptr;
    int (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Reduce)(const void *, void *, size_t, int, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    std::once_flag once;
    bool ok = false;
    // one attempt per process, whichever thread comes first; a library that lacks one of the symbols stays "not loaded"
    bool load() { std::call_once(once, [this]() { ok = load_once(); }); return ok; }
    bool load_once() {
        for (const char *n : { "libnccl.so.2", "libnccl.so" }) { h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
        if (!h) return false;
        CommInitAll = (decltype(CommInitAll)) dlsym(h, "ncclCommInitAll");
        CommDestroy = (decltype(CommDestroy)) dlsym(h, "ncclCommDestroy");
        Reduce = (decltype(Reduce)) dlsym(h, "ncclReduce");
        GroupStart = (decltype(GroupStart)) dlsym(h, "ncclGroupStart");
        GroupEnd = (decltype(GroupEnd)) dlsym(h, "ncclGroupEnd");
        GetErrorString = (decltype(GetErrorString)) dlsym(h, "ncclGetErrorString");
        return CommInitAll && CommDestroy && Reduce && GroupStart && GroupEnd;
    }
};
Nccl g_nccl;

// Communicators are kept per device list for the life of the process: ncclCommInitAll costs ~3 s on an 8-GPU node, which
// a progressive render (one call per pass) or a second frame should not pay again.
struct CommSet { std::vector<int> devs; std::vector<ncclComm_t> comms; };
std::mutex g_comm_mtx;
std::vector<CommSet> g_comm_cache;
int get_comms(const std::vector<int> &devs, std::vector<ncclComm_t> &out) {
    std::lock_guard<std::mutex> lk(g_comm_mtx);
    for (const CommSet &c : g_comm_cache) if (c.devs == devs) { out = c.comms; return kNcclSuccess; }
    CommSet c; c.devs = devs; c.comms.assign(devs.size(), nullptr);
    const int rc = g_nccl.CommInitAll(c.comms.data(), (int) devs.size(), devs.data());
    if (rc != kNcclSuccess) return rc;
    g_comm_cache.push_back(c);
    out = c.comms;
    return kNcclSuccess;
}

__global__ void k_scale(float *film, size_t n, float inv) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) film[i] *= inv;
}

}  // namespace

// Communicator set-up ahead of the first multi-GPU render (ncclCommInitAll: ~3 s on an 8-GPU node), so that it is not part
// of that call's wall time.  Optional: b2_render_bdpt_multi creates the communicators on demand.
extern "C" int b2_multi_init(const int *devices, int n_gpus) {
    if (!devices || n_gpus <= 0) { b2_set_error("NULL / invalid argument"); return B2_ERR_INVALID; }
    if (n_gpus == 1) return B2_OK;
    for (int g = 0; g < n_gpus; ++g) for (int k = 0; k < g; ++k) if (devices[k] == devices[g]) { b2_set_error("device %d listed twice", devices[g]); return B2_ERR_INVALID; }
    if (!g_nccl.load()) { b2_set_error("multi-GPU render needs NCCL (libnccl.so.2 could not be loaded or lacks a symbol)"); return B2_ERR_STATE; }
    std::vector<int> devs(devices, devices + n_gpus);
    std::vector<ncclComm_t> comms;
    const int nrc = get_comms(devs, comms);
    if (nrc != kNcclSuccess) { b2_set_error("ncclCommInitAll failed: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(nrc) : "?"); return B2_ERR_CUDA; }
    return B2_OK;
}

extern "C" int b2_render_bdpt_multi(B2Scene **scenes, int n_gpus, const B2Camera *cam, const B2RenderParams *params, float *film_rgb_host, B2Stats *stats) {
    if (!scenes || n_gpus <= 0 || !cam || !params || !film_rgb_host) { b2_set_error("NULL / invalid argument"); return B2_ERR_INVALID; }
    for (int g = 0; g < n_gpus; ++g) {
        B2_TRY(b2_check_built(scenes[g]));
        for (int k = 0; k < g; ++k) if (scenes[k]->device == scenes[g]->device) { b2_set_error("two scene handles on device %d", scenes[g]->device); return B2_ERR_INVALID; }
    }
    if (params->width <= 0 || params->height <= 0 || params->spp <= 0) { b2_set_error("invalid film size / spp"); return B2_ERR_INVALID; }
    if (params->film_mode != B2_FILM_MEAN && params->film_mode != B2_FILM_SUM) { b2_set_error("film_mode must be B2_FILM_MEAN or B2_FILM_SUM here (B2_FILM_ACCUMULATE: device entry point only)"); return B2_ERR_INVALID; }
    if (n_gpus == 1) return b2_render_bdpt(scenes[0], cam, params, film_rgb_host, stats);
    if (!g_nccl.load()) { b2_set_error("multi-GPU render needs NCCL (libnccl.so.2 could not be loaded or lacks a symbol)"); return B2_ERR_STATE; }
    const size_t nf = (size_t) 3 * params->width * params->height;
    std::vector<int> devs(n_gpus);
    for (int g = 0; g < n_gpus; ++g) devs[g] = scenes[g]->device;
    std::vector<ncclComm_t> comms;
    int nrc = get_comms(devs, comms);
    if (nrc != kNcclSuccess) { b2_set_error("ncclCommInitAll failed: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(nrc) : "?"); return B2_ERR_CUDA; }
    std::vector<float *> d_film(n_gpus, nullptr);
    std::vector<int> rc(n_gpus, B2_OK);
    std::vector<std::string> err(n_gpus);
    std::vector<B2Stats> st(n_gpus);
    // every scene stays locked across BOTH phases (the reduce and the read-back use its film buffer and its stream), taken
    // in device order so that two concurrent multi-GPU calls over the same scenes cannot deadlock
    std::vector<int> order(n_gpus);
    for (int g = 0; g < n_gpus; ++g) order[g] = g;
    std::sort(order.begin(), order.end(), [&](int a, int b) { return devs[a] < devs[b]; });
    std::vector<std::unique_ptr<std::lock_guard<std::mutex>>> locks;
    for (int k = 0; k < n_gpus; ++k) locks.emplace_back(new std::lock_guard<std::mutex>(scenes[order[k]]->mtx));
    // phase 1: every GPU renders its shard (threads: the wave loop blocks on its own stream)
    std::vector<std::thread> th;
    for (int g = 0; g < n_gpus; ++g)
        th.emplace_back([&, g]() {
            B2Scene *s = scenes[g];
            if (cudaSetDevice(s->device) != cudaSuccess || s->ws_ref().ensure_out(nf) != B2_OK) { rc[g] = B2_ERR_OOM; err[g] = "film buffer allocation failed"; return; }
            d_film[g] = s->ws_ref().film_out.p;
            B2RenderParams p = *params;
            const int stride = p.sample_stride > 0 ? p.sample_stride : 1;
            p.sample_offset = params->sample_offset + g * stride; p.sample_stride = stride * n_gpus; p.film_mode = B2_FILM_SUM;
            L2Window l2w(s, s->stream);
            rc[g] = b2_render_device_impl(s->view(), s->bvh.depth, s->ws_ref(), *cam, p, d_film[g], s->stream, &st[g]);
            if (rc[g] != B2_OK) err[g] = b2_last_error();
        });
    for (auto &t : th) t.join();
    int bad = B2_OK;
    for (int g = 0; g < n_gpus; ++g) if (rc[g] != B2_OK) { bad = rc[g]; b2_set_error("GPU %d: %s", devs[g], err[g].c_str()); break; }
    // phase 2: one reduce of the raw-sum films onto GPU 0 (NVLink / NVSwitch), resolve, read back
    cudaError_t e = cudaSuccess;
    if (bad == B2_OK) {
        g_nccl.GroupStart();
        for (int g = 0; g < n_gpus; ++g) {
            e = cudaSetDevice(devs[g]);
            if (e != cudaSuccess) break;
            nrc = g_nccl.Reduce(d_film[g], d_film[g], nf, kNcclFloat32, kNcclSum, 0, comms[g], scenes[g]->stream);
            if (nrc != kNcclSuccess) break;
        }
        const int nrc2 = g_nccl.GroupEnd();
        if (e != cudaSuccess) { b2_set_error("cudaSetDevice failed before the reduce: %s", cudaGetErrorString(e)); bad = B2_ERR_CUDA; }
        else if (nrc != kNcclSuccess || nrc2 != kNcclSuccess) { b2_set_error("ncclReduce failed: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(nrc != kNcclSuccess ? nrc : nrc2) : "?"); bad = B2_ERR_CUDA; }
    }
    if (bad == B2_OK) {
        for (int g = 1; g < n_gpus && e == cudaSuccess; ++g) { e = cudaSetDevice(devs[g]); if (e == cudaSuccess) e = cudaStreamSynchronize(scenes[g]->stream); }
        if (e == cudaSuccess) e = cudaSetDevice(devs[0]);
        if (e == cudaSuccess) {
            if (params->film_mode != B2_FILM_SUM) k_scale<<<blocks_for(nf, 256), 256, 0, scenes[0]->stream>>>(d_film[0], nf, 1.0f / (float) params->spp);
            // a page-locked caller buffer takes the film straight from the device (as in b2_render_bdpt); a pageable one goes
            // through the pinned staging buffer of scene 0
            cudaPointerAttributes attr;
            const bool pinned = cudaPointerGetAttributes(&attr, film_rgb_host) == cudaSuccess && attr.type == cudaMemoryTypeHost;
            cudaGetLastError();
            float *dst = pinned ? film_rgb_host : scenes[0]->ws_ref().h_film;
            e = cudaMemcpyAsync(dst, d_film[0], nf * sizeof(float), cudaMemcpyDeviceToHost, scenes[0]->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(scenes[0]->stream);
            if (e == cudaSuccess && !pinned) memcpy(film_rgb_host, dst, nf * sizeof(float));
        }
        if (e != cudaSuccess) { b2_set_error("reduce / film read-back failed: %s", cudaGetErrorString(e)); bad = B2_ERR_CUDA; }
    }
    for (int g = 0; g < n_gpus; ++g) { cudaSetDevice(devs[g]); }
    if (stats && bad == B2_OK) {
        *stats = st[0];
        for (int g = 1; g < n_gpus; ++g) {
            stats->eye_samples += st[g].eye_samples; stats->closest_rays += st[g].closest_rays; stats->shadow_rays += st[g].shadow_rays;
            stats->path_samples += st[g].path_samples; stats->delta_cap_hits += st[g].delta_cap_hits; stats->nodes_visited += st[g].nodes_visited;
            stats->tris_tested += st[g].tris_tested; stats->n_launches += st[g].n_launches; stats->pairs_tested += st[g].pairs_tested;
            stats->eye_vertices += st[g].eye_vertices; stats->n_waves += st[g].n_waves; stats->n_batches += st[g].n_batches;
            if (st[g].ms_total > stats->ms_total) stats->ms_total = st[g].ms_total;       // device time = max over GPUs
        }
    }
    return bad;
}
