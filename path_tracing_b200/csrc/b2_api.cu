// b2_api.cu -- the extern "C" entry points of include/b2pt.h that touch the device: scene upload + BVH build,
// ray queries, the BDPT render, the light pool and the table kernels.  Host-only entry points (scene grammar
// parser, film writers) live in b2_host.cpp.  There is no CPU execution path in this library: every compute
// call needs a CUDA device and fails with B2_ERR_NO_DEVICE otherwise.
#include <cstdarg>
#include <cstring>
#include <mutex>
#include "b2_device_scene.hpp"
#include "b2_host_scene.hpp"
#include "b2_queries.hpp"
#include "b2_render.hpp"
#include "b2_scene_handle.hpp"
#include <vector>

static thread_local char g_err[512] = "";
void b2_set_error(const char *fmt, ...) {
    va_list ap; va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// ---- per-device workspace cache (one spare workspace per device) ----
static std::mutex g_ws_mtx;
static RenderWorkspace *g_ws_cache[64];
RenderWorkspace &B2Scene::ws_ref() {
    if (!ws_ptr) {
        std::lock_guard<std::mutex> lk(g_ws_mtx);
        if (device >= 0 && device < 64 && g_ws_cache[device]) { ws_ptr = g_ws_cache[device]; g_ws_cache[device] = nullptr; }
        else ws_ptr = new RenderWorkspace;
    }
    return *ws_ptr;
}
void b2_workspace_release(B2Scene *s) {
    if (!s->ws_ptr) return;
    std::lock_guard<std::mutex> lk(g_ws_mtx);
    if (s->device >= 0 && s->device < 64 && !g_ws_cache[s->device]) g_ws_cache[s->device] = s->ws_ptr;
    else delete s->ws_ptr;
    s->ws_ptr = nullptr;
}

static int need_device(int device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) { cudaGetLastError(); b2_set_error("no CUDA device available (%s); libb2pt has no CPU fallback", e != cudaSuccess ? cudaGetErrorString(e) : "device count 0"); return B2_ERR_NO_DEVICE; }
    if (device < 0 || device >= n) { b2_set_error("device %d out of range (0..%d)", device, n - 1); return B2_ERR_INVALID; }
    B2_CUDA(cudaSetDevice(device));
    return B2_OK;
}
// Page-locked staging memory for the triangle upload records, kept for the life of the process: one block, handed to one
// scene at a time (HostScene's TriBuf hooks).  b2_scene_create returns it as soon as the records are on the device, so eight
// scenes created one after the other (pt_cli --gpus 8) share it; a scene that finds it taken gets its own cudaHostAlloc.
namespace {
std::mutex g_stage_mtx;
void *g_stage = nullptr; size_t g_stage_cap = 0; bool g_stage_busy = false;
void *stage_alloc(size_t bytes) {
    {
        std::lock_guard<std::mutex> lk(g_stage_mtx);
        if (!g_stage_busy) {
            if (g_stage_cap < bytes) {
                if (g_stage) cudaFreeHost(g_stage);
                g_stage = nullptr; g_stage_cap = 0;
                const size_t cap = bytes + bytes / 8;              // a little room: the next scene of a session is rarely the same size
                if (cudaHostAlloc(&g_stage, cap, cudaHostAllocPortable) == cudaSuccess) g_stage_cap = cap; else { g_stage = nullptr; cudaGetLastError(); }
            }
            if (g_stage) { g_stage_busy = true; return g_stage; }
        }
    }
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }   // nullptr: TriBuf falls back to malloc
    return p;
}
void stage_free(void *p) {
    {
        std::lock_guard<std::mutex> lk(g_stage_mtx);
        if (p == g_stage) { g_stage_busy = false; return; }
    }
    cudaFreeHost(p);
}
int upload_tris(DevBuf<BuildTri> &d, const TriBuf &h, cudaStream_t st) {
    B2_TRY(d.alloc(h.size()));
    if (h.size()) B2_CUDA(cudaMemcpyAsync(d.p, h.data(), h.size() * sizeof(BuildTri), cudaMemcpyHostToDevice, st));
    return B2_OK;
}
}  // namespace

template <class T> static int upload(DevBuf<T> &d, const std::vector<T> &h, cudaStream_t st) {
    B2_TRY(d.alloc(h.size()));
    if (!h.empty()) B2_CUDA(cudaMemcpyAsync(d.p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, st));
    return B2_OK;
}

extern "C" {

const char *b2_last_error(void) { return g_err; }
const char *b2_version(void) { return "b2pt 0.2.6 (sm_100a)"; }     // bump with every kernel change: profiles/*_traffic.json are tied to it
int b2_device_count(int *count) {
    if (!count) { b2_set_error("count is NULL"); return B2_ERR_INVALID; }
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { cudaGetLastError(); n = 0; }
    *count = n;
    return B2_OK;
}

int b2_scene_create(const B2Tri *tris, size_t n_tris, const B2Sphere *spheres, size_t n_spheres, const B2Material *materials, size_t n_materials,
                    const B2Light *lights, size_t n_lights, int device, B2Scene **out) {
    if (!out) { b2_set_error("out is NULL"); return B2_ERR_INVALID; }
    *out = nullptr;
    if ((n_tris && !tris) || (n_spheres && !spheres) || (n_materials && !materials) || (n_lights && !lights)) { b2_set_error("NULL array with non-zero count"); return B2_ERR_INVALID; }
    if (n_tris >= (1u << 28) || n_spheres >= (1u << 28) || n_lights >= (1u << 20)) { b2_set_error("scene too large"); return B2_ERR_INVALID; }
    B2_TRY(need_device(device));
    B2Scene *s = new B2Scene;
    s->device = device;
    if (cudaDeviceGetAttribute(&s->n_sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || s->n_sms <= 0) s->n_sms = 148;   // B200
    const char *err = nullptr;
    TriBuf::alloc_hook() = stage_alloc; TriBuf::free_hook() = stage_free;
    if (!flatten_scene(tris, n_tris, spheres, n_spheres, materials, n_materials, lights, n_lights, s->hs, &err)) { b2_set_error("%s", err ? err : "bad scene"); delete s; return B2_ERR_INVALID; }
    int rc = B2_OK;
    do {
        if (cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) != cudaSuccess) { b2_set_error("cudaStreamCreate failed"); rc = B2_ERR_CUDA; break; }
        if ((rc = upload_tris(s->d_in, s->hs.tris, s->stream)) != B2_OK) break;
        if ((rc = upload(s->sph, s->hs.sph, s->stream)) != B2_OK) break;
        if ((rc = upload(s->lights, s->hs.lights, s->stream)) != B2_OK) break;
        if ((rc = upload(s->mats, s->hs.mats, s->stream)) != B2_OK) break;
        if ((rc = upload(s->groups, s->hs.groups, s->stream)) != B2_OK) break;
        if ((rc = s->isect_counters.alloc(3)) != B2_OK) break;
        if (cudaMemsetAsync(s->isect_counters.p, 0, 3 * sizeof(unsigned long long), s->stream) != cudaSuccess || cudaStreamSynchronize(s->stream) != cudaSuccess) {
            b2_set_error("scene upload failed: %s", cudaGetErrorString(cudaGetLastError())); rc = B2_ERR_CUDA; break;
        }
    } while (0);
    if (rc != B2_OK) { b2_scene_destroy(s); return rc; }           // (drains the stream before the records are freed)
    s->hs.tris.release();                                      // the records are on the device (stream drained above): staging block back to the cache
    *out = s;
    return B2_OK;
}

void b2_scene_destroy(B2Scene *s) {
    if (!s) return;
    cudaSetDevice(s->device);
    if (s->stream) { cudaStreamSynchronize(s->stream); cudaStreamDestroy(s->stream); }
    b2_workspace_release(s);
    // the stream has drained: the scene's device blocks go to the block cache for the next scene (DevBlockCache, b2_cuda_util.hpp)
    s->d_in.release_cached(); s->sph.release_cached(); s->lights.release_cached(); s->mats.release_cached(); s->groups.release_cached();
    s->isect_counters.release_cached(); s->bvh.nodes.release_cached(); s->bvh.tri.release_cached(); s->bvh.tri_obj.release_cached();
    delete s;
}

int b2_scene_build_bvh(B2Scene *s, const B2BvhParams *params) {
    if (!s) { b2_set_error("scene is NULL"); return B2_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(s->mtx);
    B2_CUDA(cudaSetDevice(s->device));
    int max_leaf = (params && params->max_leaf_tris > 0) ? params->max_leaf_tris : 3;
    if (max_leaf > 3) { b2_set_error("max_leaf_tris must be 1..3"); return B2_ERR_INVALID; }
    const float pad_rel = (params && params->pad_rel > 0.0f) ? params->pad_rel : 3.8146973e-6f;   // 2^-18
    const HostScene &hs = s->hs;
    const float ext = std::max({ hs.scene_max.x - hs.scene_min.x, hs.scene_max.y - hs.scene_min.y, hs.scene_max.z - hs.scene_min.z });
    const float pad = pad_rel * std::max(hs.max_abs, ext) + 1e-30f;
    const int builder = params ? params->builder : B2_BUILDER_PLOC;
    if (builder != B2_BUILDER_PLOC && builder != B2_BUILDER_LBVH) { b2_set_error("unknown BVH builder %d", builder); return B2_ERR_INVALID; }
    const int radius = (params && params->ploc_radius > 0) ? std::min(params->ploc_radius, 64) : 2;   // measured on C3 / C5: radius 2 > 3 > 4 > 8 (profiles/r01_ploc_summary.md)
    B2_TRY(b2_build_bvh_device(s->d_in.p, (int) hs.n_tris, pad, max_leaf, builder, radius, s->stream, s->bvh));
    s->built = true; s->max_leaf = max_leaf;
    // persisting-L2 carve-out for the top of the node array: at most half of the L2 and at most what the device allows;
    // B2_L2_PERSIST_MB overrides (0 = no window) for A/B measurements (profiles/r02_l2_window.md)
    {
        int max_persist = 0, max_window = 0, l2 = 0;
        cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, s->device);
        cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, s->device);
        cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, s->device);
        size_t budget = std::min<size_t>((size_t) std::max(max_persist, 0), (size_t) std::max(l2, 0) / 2);
        if (const char *e = getenv("B2_L2_PERSIST_MB")) budget = std::min<size_t>((size_t) std::max(max_persist, 0), (size_t) std::max(atol(e), 0l) << 20);
        const size_t node_bytes = (size_t) s->bvh.n_nodes * sizeof(uint4) * B2_NODE_UINT4;
        s->l2_window_bytes = 0; s->l2_hit_ratio = 1.0f;
        if (budget > 0 && node_bytes > 0 && max_window > 0) {
            size_t cur = 0;
            cudaDeviceGetLimit(&cur, cudaLimitPersistingL2CacheSize);
            if (cur < budget && cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, budget) != cudaSuccess) { cudaGetLastError(); budget = cur; }
            s->l2_window_bytes = std::min<size_t>(std::min<size_t>(node_bytes, budget), (size_t) max_window);
        }
    }
    return B2_OK;
}

int b2_scene_bvh_info(const B2Scene *s, B2BvhInfo *info) {
    if (!s || !info) { b2_set_error("NULL argument"); return B2_ERR_INVALID; }
    if (!s->built) { b2_set_error("BVH not built: call b2_scene_build_bvh first"); return B2_ERR_STATE; }
    memset(info, 0, sizeof(*info));
    info->n_tris = (uint64_t) s->bvh.n_tris; info->n_nodes = (uint64_t) s->bvh.n_nodes;
    info->node_bytes = (uint64_t) s->bvh.n_nodes * sizeof(uint4) * B2_NODE_UINT4; info->tri_bytes = (uint64_t) s->bvh.n_tris * 48u;
    info->depth = s->bvh.depth; info->build_ms = s->bvh.build_ms; info->sah_cost = s->bvh.sah_cost; info->build_passes = s->bvh.build_passes;
    return B2_OK;
}

int b2_check_built(B2Scene *s) {
    if (!s) { b2_set_error("scene is NULL"); return B2_ERR_INVALID; }
    if (!s->built) { b2_set_error("BVH not built: call b2_scene_build_bvh first"); return B2_ERR_STATE; }
    return B2_OK;
}

int b2_intersect_device(B2Scene *s, const float *d_org, const float *d_dir, size_t n, int mode, int32_t *d_prim_id, float *d_t, void *stream) {
    B2_TRY(b2_check_built(s));
    if (n && (!d_org || !d_dir || !d_prim_id)) { b2_set_error("NULL ray / result buffer"); return B2_ERR_INVALID; }
    B2_CUDA(cudaSetDevice(s->device));
    L2Window l2w(s, (cudaStream_t) stream);
    return b2_intersect_launch(s->view(), s->bvh.depth, d_org, d_dir, n, mode, d_prim_id, d_t, s->isect_counters.p, (cudaStream_t) stream, &s->ray_sort);
}

int b2_intersect(B2Scene *s, const float *org, const float *dir, size_t n, int mode, int32_t *prim_id, float *t) {
    B2_TRY(b2_check_built(s));
    if (n == 0) return B2_OK;
    if (!org || !dir || !prim_id) { b2_set_error("NULL ray / result buffer"); return B2_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(s->mtx);
    B2_CUDA(cudaSetDevice(s->device));
    DevBuf<float> d_o, d_d, d_t; DevBuf<int32_t> d_id;
    B2_TRY(d_o.alloc(3 * n)); B2_TRY(d_d.alloc(3 * n)); B2_TRY(d_id.alloc(n));
    const bool want_t = t && !(mode & B2_ANY);
    if (want_t) B2_TRY(d_t.alloc(n));
    B2_CUDA(cudaMemcpyAsync(d_o.p, org, 3 * n * sizeof(float), cudaMemcpyHostToDevice, s->stream));
    B2_CUDA(cudaMemcpyAsync(d_d.p, dir, 3 * n * sizeof(float), cudaMemcpyHostToDevice, s->stream));
    L2Window l2w(s, s->stream);
    B2_TRY(b2_intersect_launch(s->view(), s->bvh.depth, d_o.p, d_d.p, n, mode, d_id.p, want_t ? d_t.p : nullptr, s->isect_counters.p, s->stream, &s->ray_sort));
    B2_CUDA(cudaMemcpyAsync(prim_id, d_id.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, s->stream));
    if (want_t) B2_CUDA(cudaMemcpyAsync(t, d_t.p, n * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
    B2_CUDA(cudaStreamSynchronize(s->stream));
    return B2_OK;
}

int b2_intersect_counters(B2Scene *s, uint64_t out[3], int reset) {
    if (!s || !out) { b2_set_error("NULL argument"); return B2_ERR_INVALID; }
    B2_CUDA(cudaSetDevice(s->device));
    B2_CUDA(cudaDeviceSynchronize());
    unsigned long long h[3];
    B2_CUDA(cudaMemcpy(h, s->isect_counters.p, sizeof(h), cudaMemcpyDeviceToHost));
    out[0] = h[0]; out[1] = h[1]; out[2] = h[2];
    if (reset) B2_CUDA(cudaMemset(s->isect_counters.p, 0, sizeof(h)));
    return B2_OK;
}

int b2_render_bdpt_device(B2Scene *s, const B2Camera *cam, const B2RenderParams *params, float *d_film_rgb, void *stream, B2Stats *stats) {
    B2_TRY(b2_check_built(s));
    if (!cam || !params || !d_film_rgb) { b2_set_error("NULL argument"); return B2_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(s->mtx);
    B2_CUDA(cudaSetDevice(s->device));
    L2Window l2w(s, stream ? (cudaStream_t) stream : s->stream);
    return b2_render_device_impl(s->view(), s->bvh.depth, s->ws_ref(), *cam, *params, d_film_rgb, stream ? (cudaStream_t) stream : s->stream, stats);
}

int b2_render_bdpt(B2Scene *s, const B2Camera *cam, const B2RenderParams *params, float *film_rgb_host, B2Stats *stats) {
    B2_TRY(b2_check_built(s));
    if (!cam || !params || !film_rgb_host) { b2_set_error("NULL argument"); return B2_ERR_INVALID; }
    if (params->width <= 0 || params->height <= 0) { b2_set_error("invalid film size"); return B2_ERR_INVALID; }
    // B2_FILM_ACCUMULATE adds into the caller's DEVICE film; here the film is a scratch buffer that still holds the previous call's image
    if (params->film_mode != B2_FILM_MEAN && params->film_mode != B2_FILM_SUM) { b2_set_error("film_mode must be B2_FILM_MEAN or B2_FILM_SUM here (B2_FILM_ACCUMULATE: device entry point only)"); return B2_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(s->mtx);
    B2_CUDA(cudaSetDevice(s->device));
    const size_t nf = (size_t) 3 * params->width * params->height;
    RenderWorkspace &ws = s->ws_ref();
    B2_TRY(ws.ensure_out(nf));
    L2Window l2w(s, s->stream);
    B2_TRY(b2_render_device_impl(s->view(), s->bvh.depth, ws, *cam, *params, ws.film_out.p, s->stream, stats));
    // a caller buffer that is already page-locked (cudaHostAlloc / cudaHostRegister, e.g. a pinned torch tensor) takes the film
    // straight from the device; a pageable one goes through the pinned staging buffer in slices, so that the host copy of
    // slice k overlaps the PCIe transfer of slice k + 1 (one 12.6 MB memcpy after the whole transfer cost 1.2 ms of the 8.7 ms C2 step)
    cudaPointerAttributes attr;
    const bool pinned = cudaPointerGetAttributes(&attr, film_rgb_host) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError();                                        // a pageable pointer is not an error for us
    if (pinned) {
        B2_CUDA(cudaMemcpyAsync(film_rgb_host, ws.film_out.p, nf * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
        B2_CUDA(cudaStreamSynchronize(s->stream));
        return B2_OK;
    }
    const size_t slice = (size_t) 1 << 19;                     // 2 MB of floats
    const size_t n_slices = (nf + slice - 1) / slice;
    std::vector<cudaEvent_t> done(n_slices);
    for (size_t k = 0; k < n_slices; ++k) {
        const size_t o = k * slice, m = std::min(slice, nf - o);
        B2_CUDA(cudaMemcpyAsync(ws.h_film + o, ws.film_out.p + o, m * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
        B2_CUDA(cudaEventCreateWithFlags(&done[k], cudaEventDisableTiming));
        B2_CUDA(cudaEventRecord(done[k], s->stream));
    }
    for (size_t k = 0; k < n_slices; ++k) {
        const size_t o = k * slice, m = std::min(slice, nf - o);
        B2_CUDA(cudaEventSynchronize(done[k]));
        memcpy(film_rgb_host + o, ws.h_film + o, m * sizeof(float));
        cudaEventDestroy(done[k]);
    }
    return B2_OK;
}

int b2_render_bdpt_progressive(B2Scene *s, const B2Camera *cam, const B2RenderParams *params, int32_t passes, float *film_rgb_host,
                               float *rms, B2PassCallback on_pass, void *user, B2Stats *stats) {
    B2_TRY(b2_check_built(s));
    if (!cam || !params || !film_rgb_host || passes <= 0) { b2_set_error("NULL / invalid argument"); return B2_ERR_INVALID; }
    if (params->width <= 0 || params->height <= 0) { b2_set_error("invalid film size"); return B2_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(s->mtx);
    B2_CUDA(cudaSetDevice(s->device));
    const size_t nf = (size_t) 3 * params->width * params->height;
    RenderWorkspace &ws = s->ws_ref();
    B2_TRY(ws.ensure_out(nf));
    const bool per_pass = rms != nullptr || on_pass != nullptr;       // otherwise the film only crosses PCIe once, at the end
    std::vector<uint8_t> cur8, last8;
    if (rms) { cur8.resize(nf); last8.resize(nf); }
    B2Stats tot; memset(&tot, 0, sizeof(tot));
    for (int32_t k = 0; k < passes; ++k) {
        B2RenderParams p = *params;
        p.seed = params->seed + (uint64_t) k;
        p.film_mode = k == 0 ? B2_FILM_MEAN : B2_FILM_ACCUMULATE;
        B2Stats st;
        L2Window l2w(s, s->stream);
        B2_TRY(b2_render_device_impl(s->view(), s->bvh.depth, ws, *cam, p, ws.film_out.p, s->stream, &st));
        tot.eye_samples += st.eye_samples; tot.closest_rays += st.closest_rays; tot.shadow_rays += st.shadow_rays; tot.path_samples += st.path_samples;
        tot.delta_cap_hits += st.delta_cap_hits; tot.nodes_visited += st.nodes_visited; tot.tris_tested += st.tris_tested;
        tot.n_batches += st.n_batches; tot.n_waves += st.n_waves; tot.n_launches += st.n_launches; tot.pairs_tested += st.pairs_tested;
        tot.eye_vertices += st.eye_vertices; tot.ms_total += st.ms_total; tot.ms_light_pool += st.ms_light_pool;
        tot.ms_primary += st.ms_primary; tot.ms_extend += st.ms_extend; tot.ms_connect += st.ms_connect;
        if (per_pass || k == passes - 1) {
            B2_CUDA(cudaMemcpyAsync(ws.h_film, ws.film_out.p, nf * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
            B2_CUDA(cudaStreamSynchronize(s->stream));
            const float inv = 1.0f / (float) (k + 1);
            for (size_t i = 0; i < nf; ++i) film_rgb_host[i] = ws.h_film[i] * inv;
            if (rms) {
                b2_film_to_rgb8(film_rgb_host, params->width, params->height, cur8.data());
                rms[k] = 0.0f;
                if (k > 0) b2_frame_rms(cur8.data(), last8.data(), nf, &rms[k]);
                cur8.swap(last8);
            }
            if (on_pass) on_pass(k, film_rgb_host, params->width, params->height, user);
        }
    }
    if (stats) *stats = tot;
    return B2_OK;
}

int b2_light_pool(B2Scene *s, const B2RenderParams *params, float *pool20_host) {
    B2_TRY(b2_check_built(s));
    if (!params || !pool20_host) { b2_set_error("NULL argument"); return B2_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(s->mtx);
    B2_CUDA(cudaSetDevice(s->device));
    return b2_light_pool_impl(s->view(), s->bvh.depth, s->ws_ref(), *params, pool20_host, s->stream);
}

int b2_bsdf_evaluate(size_t n, const float *mtl6, const float *wo, const float *wi, const float *N, float *out3) {
    B2_TRY(need_device(0));
    return b2_tables_bsdf_evaluate(n, mtl6, wo, wi, N, out3);
}
int b2_bsdf_pdf(size_t n, const float *mtl6, const float *wo, const float *wi, const float *N, float *out) {
    B2_TRY(need_device(0));
    return b2_tables_bsdf_pdf(n, mtl6, wo, wi, N, out);
}
int b2_bsdf_sample(size_t n, const float *mtl6, const float *wo, const float *N, const float *u3, const float *cur_eta,
                   float *wi_out, float *f_out, float *pdf_out, int32_t *delta_out, float *eta_out) {
    B2_TRY(need_device(0));
    return b2_tables_bsdf_sample(n, mtl6, wo, N, u3, cur_eta, wi_out, f_out, pdf_out, delta_out, eta_out);
}
int b2_mis_weight(size_t n, const float *eye17, const int32_t *n_eye, const int32_t *s_idx, const float *lit20, const int32_t *n_lit,
                  const int32_t *t_idx, const float *d_vec, const float *dist2, const float *cam_pos, int32_t max_verts, float *out) {
    B2_TRY(need_device(0));
    return b2_tables_mis_weight(n, eye17, n_eye, s_idx, lit20, n_lit, t_idx, d_vec, dist2, cam_pos, max_verts, out);
}
int b2_philox(size_t n, const uint32_t *ctr4, const uint32_t *key2, uint32_t *out4, float *unit_out4) {
    B2_TRY(need_device(0));
    return b2_tables_philox(n, ctr4, key2, out4, unit_out4);
}

}  // extern "C"
