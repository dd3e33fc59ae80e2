// b2_render.hpp -- host-side declarations of the wavefront renderer (b2_render.cu) used by the C ABI (b2_api.cu).
#pragma once
#include "b2_cuda_util.hpp"
#include "b2_bdpt.cuh"
#include "b2_queries.hpp"

// Device buffers of the wavefront state; owned by a B2Scene and kept across calls (grown on demand).
struct RenderWorkspace {
    DevBuf<float4> ray_o, ray_d, thr, last_pos, last_n, hist_pe, cv, pool, pre;
    DevBuf<float2> hist_pdf;
    DevBuf<int> live_a, live_b, ints;                // ints: n_pool_active, n_live_next, n_cv, n_paths
    DevBuf<float> film, film_out;                    // running sums; resolved film of the host-buffer entry point
    float *h_film = nullptr; size_t h_film_n = 0;    // pinned staging for the film read-back
    int ensure_out(size_t film_floats);
    DevBuf<unsigned long long> counters;
    int *h_ints = nullptr;                           // pinned read-back slot
    int cap_paths = 0, cap_depth = 0, cap_pool = 0;
    unsigned long long closest_rays_host = 0;
    int ensure(int max_paths, int eye_depth, int n_pool, size_t film_floats);
    // the pt / ppm integrators (b2_integrators.cu): path state without the BDPT history, hit points and their cell table
    int ensure_pt(int max_paths, size_t film_floats);
    DevBuf<float4> hp_in, hp, flux;
    DevBuf<unsigned long long> hp_keys, hp_keys_sorted;
    DevBuf<uint32_t> hp_order, hp_order_sorted;
    DevBuf<ulonglong2> cell_table;
    DevBuf<unsigned char> sort_tmp;
    int ensure_ppm(int n_pix, size_t film_floats, size_t sort_tmp_bytes);
    // B2_STRATEGIES_MIS as a wavefront (b2_mis_wave.cu): walk state, stored light vertices, eye history and records; counters + queues; pairs
    DevBuf<float4> mw_f4; DevBuf<int> mw_i32; DevBuf<uint2> mw_pairs;
    int ensure_mis_wave(int slots, int n_lights, size_t film_floats);
    // connection stage on position-sorted eye vertices (B2_CONN_SORT_CV): the reordered records and the sort's scratch
    DevBuf<float4> cv_sorted; RaySortScratch cv_sort;
    int ensure_cv_sort(int max_paths);
    unsigned long long sort_cfg = 0; int sort_yes = 0;       // the last render's decision and the sampling parameters it was taken for
    void release();
    ~RenderWorkspace() { release(); }
};

void b2_setup_ctx(RenderCtx &c, const DevScene &sc, const B2Camera &cam, const B2RenderParams &p, RenderWorkspace &ws, int max_paths);
int b2_render_pt_impl(const DevScene &sc, int bvh_depth, RenderWorkspace &ws, const B2Camera &cam, const B2RenderParams &p,
                      float *d_film_out, cudaStream_t st, B2Stats *stats);
int b2_render_ppm_impl(const DevScene &sc, int bvh_depth, RenderWorkspace &ws, const B2Camera &cam, const B2RenderParams &p,
                       float *d_film_out, cudaStream_t st, B2Stats *stats);
int b2_render_mis_impl(const DevScene &sc, int bvh_depth, RenderWorkspace &ws, const B2Camera &cam, const B2RenderParams &p,
                       float *d_film_out, cudaStream_t st, B2Stats *stats);      // B2_STRATEGIES_MIS (b2_mis.cuh)
int b2_render_mis_wave_impl(const DevScene &sc, int bvh_depth, RenderWorkspace &ws, const B2Camera &cam, const B2RenderParams &p,
                            float *d_film_out, cudaStream_t st, B2Stats *stats); // ... as a wavefront (b2_mis_wave.cu), the default

// d_film_out: W*H*3 floats on the device.  Blocks until `st` has drained the render (the counters are read back).
int b2_render_device_impl(const DevScene &sc, int bvh_depth, RenderWorkspace &ws, const B2Camera &cam, const B2RenderParams &p,
                          float *d_film_out, cudaStream_t st, B2Stats *stats);
int b2_light_pool_impl(const DevScene &sc, int bvh_depth, RenderWorkspace &ws, const B2RenderParams &p, float *pool20_host, cudaStream_t st);
