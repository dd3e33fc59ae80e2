// tests/hostsim/hostsim_render.hpp -- TEST-ONLY sequential host driver of the wavefront BDPT stage bodies
// (path_tracing_b200/csrc/b2_bdpt.cuh).  Same wave loop as b2_render.cu, one "thread" at a time.
#pragma once
#include <vector>
#include "../../path_tracing_b200/csrc/b2_bdpt.cuh"

struct HsRenderOut { unsigned long long counters[B2_CNT_N]; };

inline void hs_setup_ctx(RenderCtx &c, const DevScene &sc, const B2Camera &cam, const B2RenderParams &p) {
    memset(&c, 0, sizeof(c));
    c.sc = sc;
    camera_basis(cam.eye, cam.look_at, cam.view_up, cam.fov, p.width, p.height, c.cam_eye, c.cUL, c.cdx, c.cdy);
    c.W = p.width; c.H = p.height; c.spl = p.spl; c.eye_depth = p.eye_depth; c.light_depth = p.light_depth;
    c.max_delta = p.max_delta > 0 ? p.max_delta : 256;
    c.sample_offset = p.sample_offset; c.sample_stride = p.sample_stride > 0 ? p.sample_stride : 1;
    c.key = make_uint2((uint32_t) p.seed, (uint32_t) (p.seed >> 32));
}

inline void hs_build_pool(RenderCtx &c, std::vector<float4> &pool, std::vector<int> &active, unsigned long long *counters) {
    const int n_paths = c.sc.n_lights * c.spl;
    c.n_pool = n_paths * c.light_depth;
    pool.assign((size_t) B2_LV_F4 * std::max(c.n_pool, 1), make_float4(0, 0, 0, 0));
    c.pool = pool.data();
    uint2 stack[B2_STACK_CAP]; TravCounters tc{ 0, 0 };
    uint32_t ncl = 0, ncap = 0;
    for (int i = 0; i < n_paths; ++i) light_path_body<true>(c, i, stack, 1, &tc, ncl, ncap);
    counters[B2_CNT_CLOSEST] += ncl; counters[B2_CNT_DELTA_CAP] += ncap; counters[B2_CNT_NODES] += tc.nodes; counters[B2_CNT_TRIS] += tc.tris;
    active.clear();
    for (int i = 0; i < c.n_pool; ++i) if (pool_vertex_active(c.pool, i)) active.push_back(i);
}

inline void hs_render(const DevScene &sc, const B2Camera &cam, const B2RenderParams &p, float *film_out, HsRenderOut &out) {
    RenderCtx c; hs_setup_ctx(c, sc, cam, p);
    memset(&out, 0, sizeof(out));
    const int npix = p.width * p.height;
    std::vector<float> film((size_t) 3 * npix, 0.0f);
    c.film = film.data(); c.counters = out.counters;
    int n_my = 0;
    for (int s = c.sample_offset; s < p.spp; s += c.sample_stride) n_my++;
    if (sc.n_lights > 0 && n_my > 0) {
        std::vector<float4> pool; std::vector<int> active;
        hs_build_pool(c, pool, active, out.counters);
        const int max_paths = p.max_batch_paths > 0 ? p.max_batch_paths : (1 << 16);
        c.max_paths = max_paths;
        std::vector<float4> ray_o(max_paths), ray_d(max_paths), thr(max_paths), last_pos(max_paths), last_n(max_paths), acc(max_paths), hist_pe((size_t) p.eye_depth * max_paths), cv((size_t) B2_CV_F4 * max_paths);
        std::vector<float2> hist_pdf((size_t) p.eye_depth * max_paths);
        std::vector<uint32_t> pix(max_paths), sidx(max_paths), pstate(max_paths);
        std::vector<int2> hit(max_paths);
        std::vector<int> live(max_paths), live_next(max_paths);
        c.ray_o = ray_o.data(); c.ray_d = ray_d.data(); c.thr = thr.data(); c.last_pos = last_pos.data(); c.last_n = last_n.data(); c.acc = acc.data();
        c.hist_pe = hist_pe.data(); c.hist_pdf = hist_pdf.data(); c.cv = cv.data(); c.pix = pix.data(); c.sidx = sidx.data(); c.pstate = pstate.data();
        c.hit = hit.data(); c.live = live.data(); c.live_next = live_next.data();
        uint2 stack[B2_STACK_CAP]; TravCounters tc{ 0, 0 };
        const long long total = (long long) npix * n_my;
        for (long long w0 = 0; w0 < total; w0 += max_paths) {
            int n_live = (int) std::min<long long>(max_paths, total - w0);
            const int n_batch = n_live;
            for (int i = 0; i < n_live; ++i) eye_gen_body(c, i, w0);
            out.counters[B2_CNT_EYE_SAMPLES] += n_live;
            while (n_live > 0 && p.eye_depth > 0) {
                for (int q = 0; q < n_live; ++q) {
                    const int path = c.live[q];
                    HitRec h = closest_hit_bvh<true>(c.sc, mk3(c.ray_o[path]), mk3(c.ray_d[path]), stack, 1, &tc);
                    c.hit[q] = make_int2(h.prim, (int) f2u(h.t));
                }
                out.counters[B2_CNT_CLOSEST] += n_live;
                int n_next = 0; uint32_t emit = 0, cap = 0;
                for (int q = 0; q < n_live; ++q) if (shade_body(c, q, emit, cap)) c.live_next[n_next++] = c.live[q];
                out.counters[B2_CNT_PATH_SAMPLES] += emit; out.counters[B2_CNT_DELTA_CAP] += cap;
                for (int q = 0; q < n_live; ++q) {
                    if (f2u(c.cv[B2_CV_F4 * q].w) == B2_CV_NONE) continue;
                    const EyeVtx ev = load_cv(c.cv, q);
                    float3 sum = mk3(0, 0, 0);
                    uint32_t nsh = 0, npa = 0;
                    for (int a : active) {
                        out.counters[B2_CNT_PAIRS]++;
                        if (!connect_pretest(c, ev.pos, ev.n, a)) continue;
                        sum = sum + connect_finish<true>(c, ev, a, stack, 1, &tc, nsh, npa);
                    }
                    out.counters[B2_CNT_SHADOW] += nsh; out.counters[B2_CNT_PATH_SAMPLES] += npa;
                    c.acc[ev.path] = mk4(mk3(c.acc[ev.path]) + sum, 0.0f);
                }
                std::swap(c.live, c.live_next);
                n_live = n_next;
            }
            for (int i = 0; i < n_batch; ++i) {
                const float3 v = mk3(c.acc[i]);
                if (!valid_color(v)) continue;
                film[3 * (size_t) c.pix[i]] += v.x; film[3 * (size_t) c.pix[i] + 1] += v.y; film[3 * (size_t) c.pix[i] + 2] += v.z;
            }
        }
        out.counters[B2_CNT_NODES] += tc.nodes; out.counters[B2_CNT_TRIS] += tc.tris;
    }
    for (size_t i = 0; i < film.size(); ++i) film_out[i] = p.film_mode == B2_FILM_SUM ? film[i] : film[i] / (float) p.spp;
}
