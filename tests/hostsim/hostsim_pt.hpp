// tests/hostsim/hostsim_pt.hpp -- TEST-ONLY sequential host drivers of the pt / ppm bodies (path_tracing_b200/csrc/b2_pt.cuh),
// same structure as b2_integrators.cu with one "thread" at a time and a std::map in place of the device cell table.
#pragma once
#include <algorithm>
#include <map>
#include <vector>
#include "hostsim_render.hpp"
#include "../../path_tracing_b200/csrc/b2_pt.cuh"

inline void hs_render_pt(const DevScene &sc_in, const B2Camera &cam, const B2RenderParams &p, float *film_out, HsRenderOut &out) {
    DevScene sc = sc_in; sc.gpu_sem = 1;
    RenderCtx c; hs_setup_ctx(c, sc, cam, p);
    memset(&out, 0, sizeof(out));
    const int npix = p.width * p.height;
    std::vector<double> film((size_t) 3 * npix, 0.0);
    uint2 stack[B2_STACK_CAP]; TravCounters tc{ 0, 0 };
    uint32_t nsh = 0, npa = 0, ncap = 0;
    int n_my = 0;
    for (int s = c.sample_offset; s < p.spp; s += c.sample_stride) n_my++;
    const long long total = (long long) npix * n_my;
    for (long long w = 0; w < total && p.eye_depth > 0; ++w) {
        PtState ps = pt_primary(c, (uint32_t) (w % npix), (int) (w / npix), 0u);
        out.counters[B2_CNT_EYE_SAMPLES]++;
        for (;;) {
            const HitRec h = closest_hit_bvh<true>(c.sc, ps.ro, ps.rd, stack, 1, &tc);
            out.counters[B2_CNT_CLOSEST]++;
            float3 add;
            const bool alive = pt_bounce<true>(c, ps, h, stack, 1, &tc, add, nsh, npa, ncap);
            film[3 * (size_t) ps.pixel] += add.x; film[3 * (size_t) ps.pixel + 1] += add.y; film[3 * (size_t) ps.pixel + 2] += add.z;
            if (!alive) break;
        }
    }
    out.counters[B2_CNT_SHADOW] = nsh; out.counters[B2_CNT_PATH_SAMPLES] = npa; out.counters[B2_CNT_DELTA_CAP] = ncap;
    out.counters[B2_CNT_NODES] = tc.nodes; out.counters[B2_CNT_TRIS] = tc.tris;
    for (size_t i = 0; i < film.size(); ++i) film_out[i] = (float) (p.film_mode == B2_FILM_SUM ? film[i] : film[i] / (double) p.spp);
}

struct HsFluxAdd { std::vector<double> *flux; void operator()(uint32_t k, float3 e) const { (*flux)[3 * k] += e.x; (*flux)[3 * k + 1] += e.y; (*flux)[3 * k + 2] += e.z; } };
inline void hs_render_ppm(const DevScene &sc_in, const B2Camera &cam, const B2RenderParams &p, float *film_out, HsRenderOut &out) {
    DevScene sc = sc_in; sc.gpu_sem = 1;
    RenderCtx c; hs_setup_ctx(c, sc, cam, p);
    memset(&out, 0, sizeof(out));
    const int npix = p.width * p.height;
    PpmCtx pc; memset(&pc, 0, sizeof(pc));
    pc.radius = p.ppm_radius > 0.0f ? p.ppm_radius : B2_PPM_RADIUS; pc.inv_cell = 1.0f / pc.radius;
    uint2 stack[B2_STACK_CAP]; TravCounters tc{ 0, 0 };
    uint32_t ncl = 0, ncap = 0, nga = 0;
    std::vector<float3> direct(npix, mk3(0, 0, 0));
    struct Rec { unsigned long long key; HitPoint hp; };
    std::vector<Rec> recs;
    for (int pixel = 0; pixel < npix; ++pixel) {
        HitPoint hp;
        if (ppm_eye_body<true>(c, (uint32_t) pixel, (uint32_t) p.sample_offset, stack, 1, &tc, hp, direct[pixel], ncl, ncap)) {
            int gx, gy, gz; ppm_cell_of(c.sc, pc.inv_cell, hp.pos, gx, gy, gz);
            recs.push_back(Rec{ ppm_cell_key(gx, gy, gz), hp });
        }
    }
    std::stable_sort(recs.begin(), recs.end(), [](const Rec &a, const Rec &b) { return a.key < b.key; });
    const int n_hp = (int) recs.size();
    std::vector<float4> hpv((size_t) B2_HP_F4 * std::max(n_hp, 1));
    size_t cap = 1024; while (cap < 2 * (size_t) std::max(n_hp, 1)) cap <<= 1;
    std::vector<ulonglong2> table(cap, ulonglong2{ 0ull, 0ull });
    pc.hp = hpv.data(); pc.table = table.data(); pc.table_mask = (uint32_t) (cap - 1);
    for (int k = 0; k < n_hp; ++k) {
        const HitPoint &hp = recs[k].hp;
        float4 *o = pc.hp + (size_t) B2_HP_F4 * k;
        o[0] = mk4(hp.pos, 0.0f); o[1] = mk4(hp.n, u2f(hp.pixel)); o[2] = mk4(hp.wo, hp.mtl.roughness); o[3] = mk4(hp.thr, hp.mtl.metallic); o[4] = mk4(hp.mtl.base, hp.mtl.eta);
        const unsigned long long key = recs[k].key + 1ull;
        uint32_t slot = ppm_hash(key) & pc.table_mask;
        while (table[slot].x != 0ull && table[slot].x != key) slot = (slot + 1u) & pc.table_mask;
        if (table[slot].x == 0ull) { table[slot].x = key; table[slot].y = (unsigned long long) (uint32_t) k; }
        table[slot].y += 1ull << 32;
    }
    std::vector<double> flux((size_t) 3 * std::max(n_hp, 1), 0.0);
    HsFluxAdd fa{ &flux };
    const int n_photons = sc.n_lights * p.spl;
    for (int i = 0; i < n_photons && n_hp > 0 && p.light_depth > 0; ++i) ppm_photon_body<true>(c, pc, i, stack, 1, &tc, fa, ncl, nga, ncap);
    for (int i = 0; i < npix; ++i) { film_out[3 * i] = direct[i].x; film_out[3 * i + 1] = direct[i].y; film_out[3 * i + 2] = direct[i].z; }
    for (int k = 0; k < n_hp; ++k) {
        const uint32_t pixel = recs[k].hp.pixel;
        const float3 r = ppm_resolve(direct[pixel], mk3((float) flux[3 * k], (float) flux[3 * k + 1], (float) flux[3 * k + 2]), true, pc.radius);
        film_out[3 * pixel] = r.x; film_out[3 * pixel + 1] = r.y; film_out[3 * pixel + 2] = r.z;
    }
    out.counters[B2_CNT_CLOSEST] = ncl; out.counters[B2_CNT_PATH_SAMPLES] = nga; out.counters[B2_CNT_DELTA_CAP] = ncap;
    out.counters[B2_CNT_SHADOW] = (unsigned long long) n_hp;
    out.counters[B2_CNT_NODES] = tc.nodes; out.counters[B2_CNT_TRIS] = tc.tris;
}
