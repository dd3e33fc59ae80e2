// tests/hostsim/hostsim_render.hpp -- TEST-ONLY sequential host driver of the wavefront BDPT stage bodies
// (path_tracing_b200/csrc/b2_bdpt.cuh).  Same wave loop as b2_render.cu, one "thread" at a time.
#pragma once
#include <vector>
#include "../../path_tracing_b200/csrc/b2_bdpt.cuh"

struct HsRenderOut { unsigned long long counters[B2_CNT_N]; unsigned long long pretest_mismatch; };

inline void hs_setup_ctx(RenderCtx &c, const DevScene &sc, const B2Camera &cam, const B2RenderParams &p) {
    memset(&c, 0, sizeof(c));
    c.sc = sc;
    camera_basis(cam.eye, cam.look_at, cam.view_up, cam.fov, p.width, p.height, c.cam_eye, c.cUL, c.cdx, c.cdy);
    c.W = p.width; c.H = p.height; c.spl = p.spl; c.eye_depth = p.eye_depth; c.light_depth = p.light_depth;
    c.w_magic = udiv_magic_of((uint32_t) p.width); c.npix_magic = udiv_magic_of((uint32_t) p.width * (uint32_t) p.height);
    c.max_delta = p.max_delta > 0 ? p.max_delta : 256;
    c.sample_offset = p.sample_offset; c.sample_stride = p.sample_stride > 0 ? p.sample_stride : 1;
    c.key = make_uint2((uint32_t) p.seed, (uint32_t) (p.seed >> 32));
}

// factored: the MIS weight with its two walks hoisted out of the pair loop, as k_pool_active / k_connect run it (mis_light_sum in
// the pool record, mis_eye_sum once per eye vertex); false: the reference's sequential walks per pair
inline void hs_build_pool(RenderCtx &c, std::vector<float4> &pool, std::vector<float4> &pre, unsigned long long *counters, bool factored = false) {
    const int n_paths = c.sc.n_lights * c.spl;
    c.n_pool = n_paths * c.light_depth;
    pool.assign((size_t) B2_LV_F4 * std::max(c.n_pool, 1), make_float4(0, 0, 0, 0));
    c.pool = pool.data();
    uint2 stack[B2_STACK_CAP]; TravCounters tc{ 0, 0 };
    uint32_t ncl = 0, ncap = 0;
    for (int i = 0; i < n_paths; ++i) light_path_body<true>(c, i, stack, 1, &tc, ncl, ncap);
    counters[B2_CNT_CLOSEST] += ncl; counters[B2_CNT_DELTA_CAP] += ncap; counters[B2_CNT_NODES] += tc.nodes; counters[B2_CNT_TRIS] += tc.tris;
    pre.clear();
    for (int i = 0; i < c.n_pool; ++i)
        if (pool_vertex_active(c.pool, i)) {
            float4 r[B2_PRE_F4]; make_pretest_rec(c, i, r); pre.insert(pre.end(), r, r + B2_PRE_F4);
            if (factored) {
                const LightVtx lv = load_lv(c.pool, i);
                const int t_idx = i % c.light_depth;
                const float ls = mis_light_sum(RenderLightAcc{ c, lv, i - t_idx }, t_idx);
                c.pool[B2_LV_F4 * i + 4].w = std::isfinite(ls) ? ls : 3.0e38f;
            }
        }
    c.pre = pre.data();
}

// Same wave structure as b2_render.cu: extend (primary or from stored state) -> compact eye-vertex list -> connect.
inline void hs_render(const DevScene &sc, const B2Camera &cam, const B2RenderParams &p, float *film_out, HsRenderOut &out, bool factored = false) {
    RenderCtx c; hs_setup_ctx(c, sc, cam, p);
    memset(&out, 0, sizeof(out));
    const int npix = p.width * p.height;
    std::vector<float> film((size_t) 3 * npix, 0.0f);
    c.film = film.data(); c.counters = out.counters;
    int n_my = 0;
    for (int s = c.sample_offset; s < p.spp; s += c.sample_stride) n_my++;
    if (sc.n_lights > 0 && n_my > 0) {
        std::vector<float4> pool, pre;
        hs_build_pool(c, pool, pre, out.counters, factored);
        const int n_act = (int) (pre.size() / B2_PRE_F4);
        const int max_paths = p.max_batch_paths > 0 ? p.max_batch_paths : (1 << 16);
        c.max_paths = max_paths;
        std::vector<float4> ray_o(max_paths), ray_d(max_paths), thr(max_paths), last_pos(max_paths), last_n(max_paths), hist_pe((size_t) std::max(p.eye_depth, 1) * max_paths), cv((size_t) B2_CV_F4 * max_paths);
        std::vector<float2> hist_pdf((size_t) std::max(p.eye_depth, 1) * max_paths);
        std::vector<int> live(max_paths), live_next(max_paths);
        c.ray_o = ray_o.data(); c.ray_d = ray_d.data(); c.thr = thr.data(); c.last_pos = last_pos.data(); c.last_n = last_n.data();
        c.hist_pe = hist_pe.data(); c.hist_pdf = hist_pdf.data(); c.cv = cv.data(); c.live = live.data(); c.live_next = live_next.data();
        uint2 stack[B2_STACK_CAP]; TravCounters tc{ 0, 0 };
        const long long total = (long long) npix * n_my;
        for (long long w0 = 0; w0 < total; w0 += max_paths) {
            int n_live = (int) std::min<long long>(max_paths, total - w0);
            out.counters[B2_CNT_EYE_SAMPLES] += n_live;
            int n_paths = 0;
            bool primary = true;
            while (n_live > 0 && p.eye_depth > 0) {
                int n_next = 0, n_cv = 0;
                for (int q = 0; q < n_live; ++q) {
                    PathState ps; int path = -1;
                    if (primary) ps = primary_path(c, (uint32_t) (w0 % ((long long) c.W * c.H)), (int) (w0 / ((long long) c.W * c.H)), (uint32_t) q); else { path = c.live[q]; ps = load_path(c, path); }
                    const HitRec h = closest_hit_bvh<true>(c.sc, ps.ro, ps.rd, stack, 1, &tc);
                    ShadeOut so; shade_step(c, ps, h, so);
                    if (so.has_cv) {
                        if (primary) path = n_paths++;
                        so.ev.path = path;
                        store_cv(c.cv, n_cv++, so.ev);
                        if (so.store_hist) {
                            c.hist_pe[(size_t) so.ev.depth * c.max_paths + path] = mk4(so.ev.pos, so.ev.mtl.eta);
                            c.hist_pdf[(size_t) so.ev.depth * c.max_paths + path] = make_float2(so.pdf_fwd, so.pdf_rev);
                        }
                        if (so.alive) { store_path(c, path, ps); c.live_next[n_next++] = path; }
                    }
                    if (so.emitter_hit) {
                        out.counters[B2_CNT_PATH_SAMPLES]++;
                        if (valid_color(so.emitted)) { film[3 * (size_t) ps.pixel] += so.emitted.x; film[3 * (size_t) ps.pixel + 1] += so.emitted.y; film[3 * (size_t) ps.pixel + 2] += so.emitted.z; }
                    }
                    if (so.delta_cap) out.counters[B2_CNT_DELTA_CAP]++;
                }
                out.counters[B2_CNT_CLOSEST] += n_live;
                for (int q = 0; q < n_cv; ++q) {
                    const EyeVtx ev = load_cv(c.cv, q);
                    float3 sum = mk3(0, 0, 0);
                    uint32_t nsh = 0, npa = 0;
                    float es = factored ? mis_eye_sum(RenderEyeAcc{ c, ev }, ev.depth) : -1.0f;
                    if (factored && !std::isfinite(es)) es = 3.0e38f;
                    for (int a = 0; a < n_act; ++a) {
                        out.counters[B2_CNT_PAIRS]++;
                        const float4 r0 = c.pre[B2_PRE_F4 * a], r1 = c.pre[B2_PRE_F4 * a + 1], r2 = c.pre[B2_PRE_F4 * a + 2];
                        const int lvi = (int) f2u(r1.w);
                        const bool pass = pretest_rec(ev.pos, ev.n, r0, r1, r2);
                        if (pass != connect_pretest(c, ev.pos, ev.n, lvi)) out.pretest_mismatch++;
                        if (!pass) continue;
                        if (!factored) { sum = sum + connect_finish<true>(c, ev, lvi, stack, 1, &tc, nsh, npa); continue; }
                        ShadowRay sr; float3 pc;
                        if (!connect_prepare(c, ev, lvi, sr, pc)) continue;
                        nsh++;
                        if (occluded_bvh<true>(c.sc, sr.o, sr.d, sr.tmax, stack, 1, &tc)) continue;
                        npa++;
                        sum = sum + connect_weight(c, ev, lvi, pc, es, true);
                    }
                    out.counters[B2_CNT_SHADOW] += nsh; out.counters[B2_CNT_PATH_SAMPLES] += npa;
                    film[3 * (size_t) ev.pixel] += sum.x; film[3 * (size_t) ev.pixel + 1] += sum.y; film[3 * (size_t) ev.pixel + 2] += sum.z;
                }
                std::swap(c.live, c.live_next);
                n_live = n_next;
                primary = false;
            }
        }
        out.counters[B2_CNT_NODES] += tc.nodes; out.counters[B2_CNT_TRIS] += tc.tris;
    }
    for (size_t i = 0; i < film.size(); ++i) film_out[i] = p.film_mode == B2_FILM_SUM ? film[i] : film[i] / (float) p.spp;
}
