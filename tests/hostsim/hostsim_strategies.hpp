// tests/hostsim/hostsim_strategies.hpp -- TEST-ONLY sequential host drivers of the strategies beyond the reference set
// (path_tracing_b200/csrc/b2_strategies.cuh, b2_mis.cuh): next-event estimation, light tracing, 1/2 + 1/2, and the full
// s/t set with the balance heuristic.  One "thread" at a time; the film is a plain float array.
#pragma once
#include <vector>
#include "hostsim_render.hpp"
#include "../../path_tracing_b200/csrc/b2_mis.cuh"

inline void hs_render_strategies(const DevScene &sc_in, const B2Camera &cam, const B2RenderParams &p, float *film_out, HsRenderOut &out) {
    DevScene sc = sc_in; sc.ggx_fix = 1;
    RenderCtx c; hs_setup_ctx(c, sc, cam, p);
    memset(&out, 0, sizeof(out));
    const int npix = p.width * p.height;
    std::vector<float> film((size_t) 3 * npix, 0.0f);
    c.film = film.data(); c.counters = out.counters;
    const CamFrame cf = cam_frame(c.cdx, c.cdy);
    uint2 stack[B2_STACK_CAP]; TravCounters tc{ 0, 0 };
    int n_my = 0;
    for (int s = c.sample_offset; s < p.spp; s += c.sample_stride) n_my++;
    const long long total = sc.n_lights > 0 ? (long long) npix * n_my : 0;
    const bool lt = p.strategies == B2_STRATEGIES_FULL || p.strategies == B2_STRATEGIES_LIGHT_TRACING;
    const bool nee = p.strategies == B2_STRATEGIES_FULL || p.strategies == B2_STRATEGIES_NEE;
    const float w_mix = p.strategies == B2_STRATEGIES_FULL ? 0.5f : 1.0f;
    uint32_t ncl = 0, nsh = 0, npa = 0;
    MisCounters mc{ 0u, 0u, 0u };
    std::vector<MisVtx> lv(B2_MIS_MAXV);
    for (long long w = 0; w < total; ++w) {
        const uint32_t pixel = (uint32_t) (w % npix);
        const uint32_t sidx = (uint32_t) (c.sample_offset + (int) (w / npix) * c.sample_stride);
        if (p.strategies == B2_STRATEGIES_MIS) { mis_sample_body<true>(c, cf, true, pixel, sidx, lv.data(), stack, 1, &tc, mc); continue; }
        if (lt) light_trace_body<true>(c, cf, (unsigned long long) sidx * (unsigned long long) npix + pixel, (float) sc.n_lights / (float) npix * w_mix, stack, 1, &tc, ncl, nsh, npa);
        if (nee && p.eye_depth > 0) {
            PathState ps = primary_path(c, (uint32_t) (w % npix), (int) (w / npix), 0u);
            for (;;) {
                const HitRec h = closest_hit_bvh<true>(c.sc, ps.ro, ps.rd, stack, 1, &tc);
                ncl++;
                ShadeOut o;
                shade_step(c, ps, h, o);
                if (o.emitter_hit) film_add(c, ps.pixel, o.emitted);
                if (o.has_cv) film_add(c, o.ev.pixel, nee_body<true>(c, o.ev, w_mix, stack, 1, &tc, nsh, npa));
                if (!o.alive) break;
            }
        }
    }
    out.counters[B2_CNT_EYE_SAMPLES] = (unsigned long long) total;
    out.counters[B2_CNT_CLOSEST] = ncl + mc.closest; out.counters[B2_CNT_SHADOW] = nsh + mc.shadow; out.counters[B2_CNT_PATH_SAMPLES] = npa + mc.paths;
    out.counters[B2_CNT_NODES] = tc.nodes; out.counters[B2_CNT_TRIS] = tc.tris;
    for (size_t i = 0; i < film.size(); ++i) film_out[i] = p.film_mode == B2_FILM_SUM ? film[i] : film[i] / (float) p.spp;
}
