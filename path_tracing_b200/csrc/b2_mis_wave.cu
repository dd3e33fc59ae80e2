// b2_mis_wave.cu -- B2_STRATEGIES_MIS as a WAVEFRONT.
//
// The estimator is the one of b2_mis.cuh (mis_sample_body: one light subpath per eye sample, every s/t strategy, balance
// heuristic over real area densities); that body stays as the one-thread-per-sample restatement the host compilation
// (tests/hostsim) and the A/B switch B2_MIS_MEGAKERNEL=1 run.  Here the same arithmetic, the same Philox streams and the
// same acceptance tests are cut into stages whose items are dense:
//
//   k_mw_light   one bounce of every LIVE light subpath: closest hit, vertex record to HBM, a (slot, vertex) record if the
//                vertex projects into the film (m = n strategy), BSDF sample; survivors compacted with one atomic per warp
//   k_mw_splat   once per batch, over the compact list of stored light vertices inside the camera frustum: BSDF, densities,
//                balance weight, shadow ray to the pinhole, film splat
//   k_mw_eye     one bounce of every LIVE eye subpath: closest hit, emitter, vertex record for the connection stage, the
//                geometric tests of the vertex against every light and every stored light vertex of ITS sample (surviving
//                pairs appended to a compact list), BSDF sample; survivors compacted
//   k_mw_connect one thread per surviving pair: both BSDFs, the density arrays of the joined path, balance weight, shadow
//                ray, film atomic
//
// b2_mis.cuh ran one thread per sample through all of this under warp votes: 7-9 of 32 lanes active, 9 900 SASS instructions
// stalling on instruction fetch, the light subpath in local memory (profiles/r01_mis_summary.md).  Subpath vertices are now
// coalesced HBM rows ([vertex][row][slot]), a dead path costs nothing from the next bounce on, and each kernel is a fraction
// of the code.  Films are equal to the megakernel's up to the order of the film atomics (tests/test_gpu_strategies.py).
#include <algorithm>
#include <cstdlib>
#include "b2_render.hpp"
#include "b2_mis.cuh"
#include "b2_warp.cuh"

namespace {

constexpr int MW_THREADS = 128;
constexpr int MW_LV_ROWS = 6;      // float4 rows per stored light vertex
constexpr int MW_CV_ROWS = 7;      // float4 rows per eye-vertex record of the current bounce
// ints[]: device-side counters of a batch
enum { MW_N_LIVE_NEXT = 0, MW_N_CV = 1, MW_N_PAIRS = 2, MW_S_MAX = 3, MW_N_INTS = 8 };

// light vertex rows: 0 {pos, pdf_fwd}  1 {n, pdf_rev}  2 {wo, delta | nd << 1}  3 {beta, roughness}  4 {base, metallic}  5 {eta, -, -, -}
// eye-vertex record rows: 0 {pos, d2prev}  1 {n, t | nd_e << 8}  2 {wo, slot}  3 {thr, pixel}  4 {base, roughness}  5 {prev_n, metallic}  6 {eta, -, -, -}
struct MisWave {
    float4 *l_ro, *l_rd, *l_alpha; int4 *l_int;              // light walk: {ro, eta} {rd, last_pdf} {alpha, first_off} {S, nd, n_delta, prev_delta}
    float4 *lv; int *l_count;                                // stored light vertices [(k * MW_LV_ROWS + row) * N + slot], their number per slot
    float4 *e_ro, *e_rd, *e_thr, *e_ppos, *e_pn; int4 *e_int; // eye walk: {ro, eta} {rd, last_pdf} {thr, -} {prev_pos, -} {prev_n, -} {T, nd_e, n_delta, prev_delta}
    float4 *e_hist;                                          // [t * N + slot] {e_fwd, e_rev, delta, -}
    float4 *cv;                                              // [row * N + q]
    int *live, *live_next;
    uint2 *pairs;                                            // {q, candidate}
    int *ints;
    int N;
};

__device__ __forceinline__ float4 &lvrow(const MisWave &w, int k, int row, int slot) { return w.lv[((size_t) k * MW_LV_ROWS + row) * w.N + slot]; }
__device__ __forceinline__ MisVtx load_lv_vertex(const MisWave &w, int k, int slot) {
    const float4 r0 = lvrow(w, k, 0, slot), r1 = lvrow(w, k, 1, slot), r2 = lvrow(w, k, 2, slot), r3 = lvrow(w, k, 3, slot), r4 = lvrow(w, k, 4, slot), r5 = lvrow(w, k, 5, slot);
    MisVtx v;
    v.pos = mk3(r0); v.pdf_fwd = r0.w; v.n = mk3(r1); v.pdf_rev = r1.w; v.wo = mk3(r2);
    const uint32_t fl = f2u(r2.w); v.delta = (int) (fl & 1u); v.nd = (int) (fl >> 1);
    v.beta = mk3(r3); v.mtl.roughness = r3.w; v.mtl.base = mk3(r4); v.mtl.metallic = r4.w; v.mtl.eta = r5.x;
    return v;
}
__device__ __forceinline__ void store_lv_vertex(const MisWave &w, int k, int slot, const MisVtx &v) {
    lvrow(w, k, 0, slot) = mk4(v.pos, v.pdf_fwd); lvrow(w, k, 1, slot) = mk4(v.n, v.pdf_rev);
    lvrow(w, k, 2, slot) = mk4(v.wo, u2f((uint32_t) v.delta | ((uint32_t) v.nd << 1)));
    lvrow(w, k, 3, slot) = mk4(v.beta, v.mtl.roughness); lvrow(w, k, 4, slot) = mk4(v.mtl.base, v.mtl.metallic);
    lvrow(w, k, 5, slot) = mk4(mk3(v.mtl.eta, 0.0f, 0.0f), 0.0f);
}
// pL / pE / dl entries 1..m of a joined path from the stored light vertices 0..m-1 of `slot`
__device__ __forceinline__ void light_densities(const MisWave &w, int slot, int m, float *pL, float *pE, unsigned char *dl) {
    for (int i = 1; i <= m; ++i) {
        pL[i] = lvrow(w, i - 1, 0, slot).w; pE[i] = lvrow(w, i - 1, 1, slot).w;
        dl[i] = (unsigned char) (f2u(lvrow(w, i - 1, 2, slot).w) & 1u);
    }
}
__device__ __forceinline__ void sample_of_slot(const RenderCtx &c, uint32_t pix0, int s0, int slot, uint32_t &pixel, uint32_t &sidx) {
    const uint32_t ds = udiv_magic(pix0 + (uint32_t) slot, (uint32_t) (c.W * c.H), c.npix_magic, pixel);
    sidx = (uint32_t) (c.sample_offset + (s0 + (int) ds) * c.sample_stride);
}

// ---------------- light subpaths: one bounce (the light walk of mis_sample_body, b2_mis.cuh) ----------------
template <bool COUNT>
__global__ void __launch_bounds__(MW_THREADS) k_mw_light(RenderCtx c, CamFrame cf, MisWave w, int n_items, int step, uint32_t pix0, int s0) {
    extern __shared__ uint2 s_stack[];
    uint2 *stack = s_stack + threadIdx.x;
    const int item = blockIdx.x * blockDim.x + threadIdx.x;
    const DevScene &sc = c.sc;
    const int n_cap = c.eye_depth, l_cap = c.light_depth - 1;
    TravCounters tc{ 0u, 0u };
    MisCounters cnt{ 0u, 0u, 0u };
    bool alive = item < n_items;
    int slot = 0, S = 0, nd = 0, n_delta = 0;
    bool prev_delta = false;
    uint32_t pixel = 0u, sidx = 0u;
    float3 ro = mk3(0, 0, 0), rd = mk3(0, 0, 1), alpha = mk3(0, 0, 0);
    float eta = 1.0f, last_pdf = 0.0f, first_off = 0.0f;
    if (alive) {
        slot = step == 0 ? item : w.live[item];
        sample_of_slot(c, pix0, s0, slot, pixel, sidx);
        if (step == 0) {
            alive = l_cap > 0 && n_cap > 0;
            if (alive) {
                const uint4 o = philox4x32_10(make_uint4(pixel, sidx, 0u, B2_RNG_MIS_LIGHT), c.key);
                const float r1 = u32_to_unit(o.x), r2 = u32_to_unit(o.y), r3 = u32_to_unit(o.z);
                int li = (int) (r3 * (float) sc.n_lights);
                if (li > sc.n_lights - 1) li = sc.n_lights - 1;
                const SpotLight L = spot_light(sc, li);
                const float omega = 2.0f * B2_PI * (1.0f - L.cos_cut);
                alive = !L.parallel && omega > 0.0f;
                float3 u_vec = (fabsf(L.dir.x) > 0.9f) ? mk3(0, 1, 0) : mk3(1, 0, 0);
                const float3 v_vec = normalize3(cross3(L.dir, u_vec));
                u_vec = normalize3(cross3(v_vec, L.dir));
                const float ct = 1.0f - r1 * (1.0f - L.cos_cut), st = fast_sqrt(fmaxf(0.0f, 1.0f - ct * ct));
                float sphi, cphi;
                sincos_2pi(r2, sphi, cphi);
                rd = normalize3(u_vec * (st * cphi) + v_vec * (st * sphi) + L.dir * ct);
                ro = L.pos + rd * L.r;
                alpha = mk3(sc.lights[4 * li + 2]) * (float) sc.n_lights;
                last_pdf = alive ? 1.0f / omega : 0.0f; first_off = L.r;
            }
            if (!alive) w.l_count[slot] = 0;
        } else {
            const float4 a = w.l_ro[slot], b = w.l_rd[slot], d = w.l_alpha[slot];
            const int4 i4 = w.l_int[slot];
            ro = mk3(a); eta = a.w; rd = mk3(b); last_pdf = b.w; alpha = mk3(d); first_off = d.w;
            S = i4.x; nd = i4.y; n_delta = i4.z; prev_delta = i4.w != 0;
        }
    }
    const bool was_alive = alive;
    HitRec h; h.prim = B2_PRIM_MISS; h.t = 0.0f;
    if (alive) { h = mis_closest<COUNT>(sc, ro, rd, stack, blockDim.x, &tc); cnt.closest++; }
    Surface s; s.is_light = false; s.pos = s.n = mk3(0, 0, 0); s.mtl.base = mk3(0, 0, 0); s.mtl.roughness = s.mtl.metallic = s.mtl.eta = 0.0f;
    if (alive && h.prim == B2_PRIM_MISS) alive = false;
    if (alive) { s = mis_surface(sc, ro, rd, h); if (s.is_light) alive = false; }
    MisVtx v; v.delta = 0; v.nd = 0; v.pos = v.n = v.wo = v.beta = mk3(0, 0, 0); v.mtl = s.mtl; v.pdf_fwd = v.pdf_rev = 0.0f;
    if (alive) {
        v.pos = s.pos; v.n = s.n; v.wo = rd * -1.0f; v.beta = alpha; v.mtl = s.mtl; v.delta = is_delta_material(s.mtl) ? 1 : 0;
        if (!v.delta && nd + 1 > l_cap) alive = false;
    }
    // ---- m = n: connect to the pinhole and splat ----
    bool splat = false;
    uint32_t px = 0; float cos_c = 1.0f, d2c = 1.0f, cos_y = 0.0f; float3 to_eye = mk3(0, 0, 1);
    if (alive) {
        nd += v.delta ? 0 : 1; v.nd = nd;
        const float dist = h.t + first_off;
        v.pdf_fwd = prev_delta ? 1.0f : area_pdf(last_pdf, s.n, rd, dist * dist);
        v.pdf_rev = 0.0f;
        store_lv_vertex(w, S, slot, v); S++;
        if (!v.delta && nd <= n_cap && camera_project(c, cf, v.pos, px, cos_c, d2c, to_eye)) { cos_y = dot3(v.n, to_eye); splat = cos_y > 0.0f; }
    }
    // the camera connection itself (BSDF, densities, weight, shadow ray, splat) runs in k_mw_splat over a compact list of
    // (slot, vertex) records once the light phase is over: few lanes of a warp have a vertex inside the frustum
    {
        const int ri = warp_alloc(&w.ints[MW_N_PAIRS], splat);
        if (splat) w.pairs[ri] = make_uint2((uint32_t) slot, (uint32_t) (S - 1));
    }
    if (alive && ((!v.delta && nd == l_cap) || S >= B2_MIS_MAXV)) alive = false;
    BsdfSample bs; bs.wi = mk3(0, 0, 0); bs.f = mk3(0, 0, 0); bs.pdf = 0.0f; bs.new_eta = eta; bs.is_delta = false;
    if (alive) {
        float u_rr, u1, u2;
        rng_draw3(c.key, B2_RNG_MIS_LIGHT, pixel, sidx, (uint32_t) step, u_rr, u1, u2);
        mis_bsdf_sample(s.mtl, v.wo, s.n, u_rr, u1, u2, eta, bs);
        if (bs.pdf <= 0.0f) alive = false;
    }
    if (alive) {
        if (S >= 2) {                                                           // density of the previous vertex seen from this one
            const float3 ppos = mk3(lvrow(w, S - 2, 0, slot)), pn = mk3(lvrow(w, S - 2, 1, slot));
            const float3 dv = ppos - v.pos;
            lvrow(w, S - 2, 1, slot).w = bs.is_delta ? 1.0f : area_pdf(bsdf_pdf(s.mtl, bs.wi, v.wo, s.n), pn, v.wo, dot3(dv, dv));
        }
        first_off = 0.0f;
        rd = bs.wi;
        if (bs.is_delta) {
            alpha = alpha * bs.f;
            eta = bs.new_eta;
            ro = s.pos + s.n * (dot3(bs.wi, s.n) < 0.0f ? -B2_EPSILON : B2_EPSILON);
            prev_delta = true;
            if (++n_delta >= c.max_delta) alive = false;
        } else {
            const float3 f_adj = bsdf_evaluate(s.mtl, bs.wi, v.wo, s.n);        // transport order: wo = towards the eye side
            alpha = alpha * f_adj * fabsf(dot3(s.n, bs.wi)) / bs.pdf;
            if (!valid_color(alpha) || (alpha.x <= 0.0f && alpha.y <= 0.0f && alpha.z <= 0.0f)) alive = false;
            ro = s.pos + s.n * B2_EPSILON;
            last_pdf = bs.pdf; prev_delta = false;
        }
    }
    if (was_alive) {
        if (alive) {
            w.l_ro[slot] = mk4(ro, eta); w.l_rd[slot] = mk4(rd, last_pdf); w.l_alpha[slot] = mk4(alpha, first_off);
            w.l_int[slot] = make_int4(S, nd, n_delta, prev_delta ? 1 : 0);
        } else {
            w.l_count[slot] = S;                                                 // the subpath is complete: S stored vertices
            atomicMax(&w.ints[MW_S_MAX], S);
        }
    }
    const int li = warp_alloc(&w.ints[MW_N_LIVE_NEXT], alive);
    if (alive) w.live_next[li] = slot;
    warp_count_add(&c.counters[B2_CNT_CLOSEST], cnt.closest);
    warp_count_add(&c.counters[B2_CNT_SHADOW], cnt.shadow);
    warp_count_add(&c.counters[B2_CNT_PATH_SAMPLES], cnt.paths);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}

// ---------------- m = n: stored light vertex -> pinhole (the splat block of mis_sample_body's light walk) ----------------
template <bool COUNT>
__global__ void __launch_bounds__(MW_THREADS) k_mw_splat(RenderCtx c, CamFrame cf, MisWave w) {
    extern __shared__ uint2 s_stack[];
    uint2 *stack = s_stack + threadIdx.x;
    const DevScene &sc = c.sc;
    const int l_cap = c.light_depth - 1;
    const float inv_nl = 1.0f / (float) sc.n_lights, wh = (float) c.W * (float) c.H;
    const int n_rec = w.ints[MW_N_PAIRS];
    TravCounters tc{ 0u, 0u };
    MisCounters cnt{ 0u, 0u, 0u };
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_rec; i += gridDim.x * blockDim.x) {
        const uint2 rec = w.pairs[i];
        const int slot = (int) rec.x, n = (int) rec.y + 1;                    // the path has n surface vertices
        const MisVtx v = load_lv_vertex(w, n - 1, slot);
        uint32_t px = 0; float cos_c = 1.0f, d2c = 1.0f; float3 to_eye = mk3(0, 0, 1);
        if (!camera_project(c, cf, v.pos, px, cos_c, d2c, to_eye)) continue;   // (it did in k_mw_light: same arithmetic)
        const float cos_y = dot3(v.n, to_eye);
        float wgt = 0.0f;
        const float3 f = bsdf_evaluate(v.mtl, to_eye, v.wo, v.n);
        if (f.x > 0.0f || f.y > 0.0f || f.z > 0.0f) {
            float pL[B2_MIS_MAXV + 2], pE[B2_MIS_MAXV + 2];
            unsigned char dl[B2_MIS_MAXV + 2];
            light_densities(w, slot, n - 1, pL, pE, dl);
            pL[n] = v.pdf_fwd; dl[n] = (unsigned char) v.delta;
            pE[n] = camera_area_pdf(cf, cos_y, cos_c, d2c);
            if (n >= 2) {
                const float3 ppos = mk3(lvrow(w, n - 2, 0, slot)), pn = mk3(lvrow(w, n - 2, 1, slot));
                const float3 dv = ppos - v.pos;
                pE[n - 1] = area_pdf(bsdf_pdf(v.mtl, to_eye, v.wo, v.n), pn, v.wo, dot3(dv, dv));
            }
            wgt = mis_balance(pL, pE, dl, n, n, l_cap, inv_nl, wh);
        }
        if (wgt > 0.0f) {
            cnt.shadow++;
            if (!mis_occluded<COUNT>(sc, x3_add(v.pos, x3_muls(v.n, B2_EPSILON)), c.cam_eye, stack, blockDim.x, &tc)) {
                const float3 val = v.beta * f * (wgt * camera_area_pdf(cf, cos_y, cos_c, d2c) / wh);
                if (valid_color(val)) { film_add(c, px, val); cnt.paths++; }
            }
        }
    }
    warp_count_add(&c.counters[B2_CNT_SHADOW], cnt.shadow);
    warp_count_add(&c.counters[B2_CNT_PATH_SAMPLES], cnt.paths);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}

// ---------------- connection stage, cheap part: does (eye vertex, candidate) pass the geometric tests ----------------
// candidate j < n_lights: light j (m = 0, next-event estimation); else stored light vertex j - n_lights of the eye vertex's own sample
__device__ __forceinline__ bool mw_candidate_passes(const RenderCtx &c, const MisWave &w, int slot, float3 pos, float3 n, int nd_e, int cand) {
    const DevScene &sc = c.sc;
    if (cand < sc.n_lights) {
        const SpotLight L = spot_light(sc, cand);
        if (L.parallel) return false;
        const float3 d = pos - L.pos;
        const float d2 = dot3(d, d);
        if (d2 < 1e-12f) return false;
        const float3 om = d * fast_rsqrt(d2);
        if (dot3(om, L.dir) < L.cos_cut) return false;
        return dot3(n, om * -1.0f) > 0.0f;
    }
    const int k = cand - sc.n_lights;
    const uint32_t fl = f2u(lvrow(w, k, 2, slot).w);
    if ((fl & 1u) || (int) (fl >> 1) + nd_e > c.eye_depth) return false;
    const float3 ypos = mk3(lvrow(w, k, 0, slot)), yn = mk3(lvrow(w, k, 1, slot));
    const float3 dv = ypos - pos;
    const float d2 = dot3(dv, dv);
    if (d2 < 1e-10f) return false;
    const float3 wz = dv * fast_rsqrt(d2);
    return dot3(n, wz) > 0.0f && dot3(yn, wz * -1.0f) > 0.0f;
}

// ---------------- eye subpaths: one bounce (the eye walk of mis_sample_body without its candidate loop) ----------------
template <bool COUNT>
__global__ void __launch_bounds__(MW_THREADS) k_mw_eye(RenderCtx c, CamFrame cf, MisWave w, int n_items, int step, uint32_t pix0, int s0) {
    extern __shared__ uint2 s_stack[];
    uint2 *stack = s_stack + threadIdx.x;
    const int item = blockIdx.x * blockDim.x + threadIdx.x;
    const DevScene &sc = c.sc;
    const int n_cap = c.eye_depth;
    TravCounters tc{ 0u, 0u };
    MisCounters cnt{ 0u, 0u, 0u };
    bool alive = item < n_items;
    int slot = 0, T = 0, nd_e = 0, n_delta = 0;
    bool prev_delta = false;
    uint32_t pixel = 0u, sidx = 0u;
    float3 ro = mk3(0, 0, 0), rd = mk3(0, 0, 1), thr = mk3(1.0f, 1.0f, 1.0f), prev_pos = c.cam_eye, prev_n = mk3(0, 0, 1);
    float eta = 1.0f, last_pdf = 1.0f, cos_cam = 1.0f;
    if (alive) {
        slot = step == 0 ? item : w.live[item];
        sample_of_slot(c, pix0, s0, slot, pixel, sidx);
        if (step == 0) {
            alive = n_cap > 0;
            float jx, jy;
            rng_draw2(c.key, B2_RNG_EYE, pixel, sidx, jx, jy);
            uint32_t col;
            const uint32_t row = udiv_magic(pixel, (uint32_t) c.W, c.w_magic, col);
            const float3 pp = c.cUL + c.cdx * ((float) col + jx) + c.cdy * ((float) row + jy);
            rd = normalize3(pp - c.cam_eye); ro = c.cam_eye;
            cos_cam = dot3(rd, cf.fwd);
            prev_n = rd;
        } else {
            const float4 a = w.e_ro[slot], b = w.e_rd[slot], t4 = w.e_thr[slot], pp = w.e_ppos[slot], pn = w.e_pn[slot];
            const int4 i4 = w.e_int[slot];
            ro = mk3(a); eta = a.w; rd = mk3(b); last_pdf = b.w; thr = mk3(t4); prev_pos = mk3(pp); prev_n = mk3(pn);
            T = i4.x; nd_e = i4.y; n_delta = i4.z; prev_delta = i4.w != 0;
        }
    }
    HitRec h; h.prim = B2_PRIM_MISS; h.t = 0.0f;
    if (alive) { h = mis_closest<COUNT>(sc, ro, rd, stack, blockDim.x, &tc); cnt.closest++; }
    Surface s; s.is_light = false; s.pos = s.n = mk3(0, 0, 0); s.mtl.base = mk3(0, 0, 0); s.mtl.roughness = s.mtl.metallic = s.mtl.eta = 0.0f;
    if (alive && h.prim == B2_PRIM_MISS) alive = false;
    if (alive) {
        s = mis_surface(sc, ro, rd, h);
        if (s.is_light) { if (T == 0) { film_add(c, pixel, s.mtl.base); cnt.paths++; } alive = false; }   // the ball of a light is only ever seen directly
    }
    bool delta = false;
    if (alive) { delta = is_delta_material(s.mtl); if (!delta && nd_e + 1 > n_cap) alive = false; }
    const float3 wo = rd * -1.0f;
    float d2prev = 1.0f;
    int t = T;
    if (alive) {
        nd_e += delta ? 0 : 1;
        const float3 dprev = s.pos - prev_pos;
        d2prev = dot3(dprev, dprev);
        const float e_fwd = T == 0 ? camera_area_pdf(cf, fabsf(dot3(s.n, wo)), cos_cam, d2prev) : (prev_delta ? 1.0f : area_pdf(last_pdf, s.n, rd, d2prev));
        w.e_hist[(size_t) T * w.N + slot] = make_float4(e_fwd, 0.0f, delta ? 1.0f : 0.0f, 0.0f);
        t = ++T;                                                                             // this is eye vertex z_t
    }
    // ---- the record the connection stage reads (this vertex against every light and every light vertex of its sample) ----
    const bool connects = alive && !delta;
    const int q = warp_alloc(&w.ints[MW_N_CV], connects);
    if (connects) {
        const size_t N = (size_t) w.N;
        w.cv[0 * N + q] = mk4(s.pos, d2prev);
        w.cv[1 * N + q] = mk4(s.n, u2f((uint32_t) t | ((uint32_t) nd_e << 8)));
        w.cv[2 * N + q] = mk4(wo, u2f((uint32_t) slot));
        w.cv[3 * N + q] = mk4(thr, u2f(pixel));
        w.cv[4 * N + q] = mk4(s.mtl.base, s.mtl.roughness);
        w.cv[5 * N + q] = mk4(prev_n, s.mtl.metallic);
        w.cv[6 * N + q] = make_float4(s.mtl.eta, 0.0f, 0.0f, 0.0f);
    }
    // the geometric tests of this vertex against its candidates, here, where the vertex is in registers and neighbouring
    // threads hold neighbouring slots (a separate kernel over a (vertex, candidate) grid was 22 % of the render at 15 lanes
    // and 17 % of the issue slots: most of the grid is empty and its loads stride by the batch size)
    {
        const int n_cand = connects ? sc.n_lights + w.l_count[slot] : 0;
        for (int cand = 0; __any_sync(0xffffffffu, cand < n_cand); ++cand) {
            const bool have = cand < n_cand && mw_candidate_passes(c, w, slot, s.pos, s.n, nd_e, cand);
            const int pi = warp_alloc(&w.ints[MW_N_PAIRS], have);
            if (have) w.pairs[pi] = make_uint2((uint32_t) q, (uint32_t) cand);
        }
    }
    if (alive && ((!delta && nd_e == n_cap) || T >= B2_MIS_MAXV)) alive = false;
    BsdfSample bs; bs.wi = mk3(0, 0, 0); bs.f = mk3(0, 0, 0); bs.pdf = 0.0f; bs.new_eta = eta; bs.is_delta = false;
    if (alive) {
        float u_rr, u1, u2;
        rng_draw3(c.key, B2_RNG_EYE, pixel, sidx, (uint32_t) step, u_rr, u1, u2);
        mis_bsdf_sample(s.mtl, wo, s.n, u_rr, u1, u2, eta, bs);
        if (bs.pdf <= 0.0f) alive = false;
    }
    if (alive) {
        if (t >= 2) w.e_hist[(size_t) (t - 2) * w.N + slot].y = bs.is_delta ? 1.0f : area_pdf(bsdf_pdf(s.mtl, bs.wi, wo, s.n), prev_n, rd, d2prev);
        prev_pos = s.pos; prev_n = s.n;
        rd = bs.wi;
        if (bs.is_delta) {
            thr = thr * bs.f;
            eta = bs.new_eta;
            ro = s.pos + s.n * (dot3(bs.wi, s.n) < 0.0f ? -B2_EPSILON : B2_EPSILON);
            prev_delta = true;
            if (++n_delta >= c.max_delta) alive = false;
        } else {
            thr = thr * bs.f * fabsf(dot3(s.n, bs.wi)) / bs.pdf;
            if (!valid_color(thr) || (thr.x <= 0.0f && thr.y <= 0.0f && thr.z <= 0.0f)) alive = false;
            ro = s.pos + s.n * B2_EPSILON;
            last_pdf = bs.pdf; prev_delta = false;
        }
    }
    if (alive) {
        w.e_ro[slot] = mk4(ro, eta); w.e_rd[slot] = mk4(rd, last_pdf); w.e_thr[slot] = mk4(thr, 0.0f);
        w.e_ppos[slot] = mk4(prev_pos, 0.0f); w.e_pn[slot] = mk4(prev_n, 0.0f);
        w.e_int[slot] = make_int4(T, nd_e, n_delta, prev_delta ? 1 : 0);
    }
    const int li = warp_alloc(&w.ints[MW_N_LIVE_NEXT], alive);
    if (alive) w.live_next[li] = slot;
    warp_count_add(&c.counters[B2_CNT_CLOSEST], cnt.closest);
    warp_count_add(&c.counters[B2_CNT_PATH_SAMPLES], cnt.paths);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}

// ---------------- connection stage, expensive part: BSDFs, densities of the joined path, weight, shadow ray ----------------
template <bool COUNT>
__global__ void __launch_bounds__(MW_THREADS) k_mw_connect(RenderCtx c, MisWave w) {
    extern __shared__ uint2 s_stack[];
    uint2 *stack = s_stack + threadIdx.x;
    const DevScene &sc = c.sc;
    const int l_cap = c.light_depth - 1;
    const float inv_nl = 1.0f / (float) sc.n_lights, wh = (float) c.W * (float) c.H;
    const int n_pairs = w.ints[MW_N_PAIRS];
    const size_t N = (size_t) w.N;
    TravCounters tc{ 0u, 0u };
    MisCounters cnt{ 0u, 0u, 0u };
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_pairs; i += gridDim.x * blockDim.x) {
        const uint2 pr = w.pairs[i];
        const int q = (int) pr.x, cand = (int) pr.y;
        const float4 r0 = w.cv[0 * N + q], r1 = w.cv[1 * N + q], r2 = w.cv[2 * N + q], r3 = w.cv[3 * N + q], r4 = w.cv[4 * N + q], r5 = w.cv[5 * N + q], r6 = w.cv[6 * N + q];
        Surface s; s.is_light = false; s.pos = mk3(r0); s.n = mk3(r1);
        s.mtl.base = mk3(r4); s.mtl.roughness = r4.w; s.mtl.metallic = r5.w; s.mtl.eta = r6.x;
        const float d2prev = r0.w;
        const int t = (int) (f2u(r1.w) & 0xffu);
        const float3 wo = mk3(r2), rd = wo * -1.0f, thr = mk3(r3), prev_n = mk3(r5);
        const int slot = (int) f2u(r2.w);
        const uint32_t pixel = f2u(r3.w);
        // the candidate again (the eye bounce kept only the pair's indices)
        int m = 0; float d2 = 1.0f, cos_z = 0.0f, cos_y = 0.0f; float3 wz = mk3(0, 0, 1), target = mk3(0, 0, 0), scale = mk3(0, 0, 0);
        MisVtx y; y.pos = y.n = y.wo = y.beta = mk3(0, 0, 0); y.mtl = s.mtl; y.pdf_fwd = y.pdf_rev = 0.0f; y.delta = 0; y.nd = 0;
        if (cand < sc.n_lights) {
            const SpotLight L = spot_light(sc, cand);
            const float3 d = s.pos - L.pos;
            d2 = dot3(d, d);
            const float3 om = d * fast_rsqrt(d2);
            wz = om * -1.0f;
            cos_z = dot3(s.n, wz);
            m = 0; target = L.pos + om * L.r; scale = L.intensity;
            cos_y = 1.0f / (2.0f * B2_PI * (1.0f - L.cos_cut));                        // m = 0: the emission density rides in cos_y
        } else {
            y = load_lv_vertex(w, cand - sc.n_lights, slot);
            const float3 dv = y.pos - s.pos;
            d2 = dot3(dv, dv);
            wz = dv * fast_rsqrt(d2);
            cos_z = dot3(s.n, wz); cos_y = dot3(y.n, wz * -1.0f);
            m = cand - sc.n_lights + 1; target = x3_add(y.pos, x3_muls(y.n, B2_EPSILON)); scale = y.beta;
        }
        float wgt = 0.0f;
        float3 val = mk3(0, 0, 0);
        const float3 fz = bsdf_evaluate(s.mtl, wo, wz, s.n);
        const int n = m + t;
        if (fz.x > 0.0f || fz.y > 0.0f || fz.z > 0.0f) {
            float pL[2 * B2_MIS_MAXV + 2], pE[2 * B2_MIS_MAXV + 2];
            unsigned char dl[2 * B2_MIS_MAXV + 2];
            for (int i2 = m + 1; i2 <= n; ++i2) {                                                                        // x_i = z_{n + 1 - i}
                const float4 e = w.e_hist[(size_t) (n - i2) * N + slot];
                pE[i2] = e.x; pL[i2] = (n - i2 == t - 1) ? 0.0f : e.y; dl[i2] = e.z != 0.0f ? 1 : 0;
            }
            if (m + 2 <= n) pL[m + 2] = area_pdf(bsdf_pdf(s.mtl, wz, wo, s.n), prev_n, rd, d2prev);                      // z, lit from the light side, generates its eye-side neighbour
            if (m == 0) {
                pL[1] = area_pdf(cos_y, s.n, wz, d2);
                val = thr * fz * scale * (cos_z / d2);
                wgt = 1.0f;
            } else {
                const float3 wy = wz * -1.0f;
                const float3 fy = bsdf_evaluate(y.mtl, wy, y.wo, y.n);
                if (fy.x > 0.0f || fy.y > 0.0f || fy.z > 0.0f) {
                    light_densities(w, slot, m, pL, pE, dl);
                    pL[m + 1] = area_pdf(bsdf_pdf(y.mtl, y.wo, wy, y.n), s.n, wz, d2);                                    // y generates z
                    pE[m] = area_pdf(bsdf_pdf(s.mtl, wo, wz, s.n), y.n, wy, d2);                                          // z generates y
                    if (m >= 2) {
                        const float3 ppos = mk3(lvrow(w, m - 2, 0, slot)), pn = mk3(lvrow(w, m - 2, 1, slot));
                        const float3 dy = ppos - y.pos;
                        pE[m - 1] = area_pdf(bsdf_pdf(y.mtl, wy, y.wo, y.n), pn, y.wo, dot3(dy, dy));
                    }
                    val = thr * fz * fy * scale * (cos_z * cos_y / d2);
                    wgt = 1.0f;
                }
            }
            if (wgt > 0.0f) wgt = mis_balance(pL, pE, dl, n, m, l_cap, inv_nl, wh);
        }
        if (wgt > 0.0f) {
            cnt.shadow++;
            if (!mis_occluded<COUNT>(sc, x3_add(s.pos, x3_muls(s.n, B2_EPSILON)), target, stack, blockDim.x, &tc)) {
                val = val * wgt;
                if (valid_color(val)) { film_add(c, pixel, val); cnt.paths++; }
            }
        }
    }
    warp_count_add(&c.counters[B2_CNT_SHADOW], cnt.shadow);
    warp_count_add(&c.counters[B2_CNT_PATH_SAMPLES], cnt.paths);
    if (COUNT) { warp_count_add(&c.counters[B2_CNT_NODES], tc.nodes); warp_count_add(&c.counters[B2_CNT_TRIS], tc.tris); }
}

__global__ void __launch_bounds__(256) k_mw_film_out(const float *__restrict__ sum, float *__restrict__ out, size_t n, float inv, int accumulate) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = accumulate ? out[i] + sum[i] * inv : sum[i] * inv;
}

inline unsigned blocks_of(long long n, int threads) { return (unsigned) std::max<long long>(1, (n + threads - 1) / threads); }

}  // namespace

// float4 rows of wave state per slot: 3 + 1 (light walk), 24 x 6 (light vertices), 5 + 1 (eye walk), 24 (eye history), 7 (records)
static constexpr size_t MW_F4_PER_SLOT = 4 + (size_t) B2_MIS_MAXV * MW_LV_ROWS + 6 + B2_MIS_MAXV + MW_CV_ROWS;

int RenderWorkspace::ensure_mis_wave(int slots, int n_lights, size_t film_floats) {
    const size_t N = (size_t) std::max(slots, 1);
    B2_TRY(mw_f4.ensure(MW_F4_PER_SLOT * N));
    B2_TRY(mw_i32.ensure(3 * N + MW_N_INTS));
    B2_TRY(mw_pairs.ensure(N * (size_t) (n_lights + B2_MIS_MAXV)));
    B2_TRY(film.ensure(film_floats));
    if (!counters.p) B2_TRY(counters.alloc(B2_CNT_N));
    if (!h_ints) B2_CUDA(cudaMallocHost((void **) &h_ints, 8 * sizeof(int)));
    return B2_OK;
}

// B2_STRATEGIES_MIS, wavefront.  Called by b2_render_mis_impl (b2_integrators.cu) after its parameter checks.
int b2_render_mis_wave_impl(const DevScene &sc, int bvh_depth, RenderWorkspace &ws, const B2Camera &cam, const B2RenderParams &p,
                            float *d_film_out, cudaStream_t st, B2Stats *stats) {
    const int npix = p.width * p.height;
    const size_t film_floats = (size_t) 3 * npix;
    const int stride = p.sample_stride > 0 ? p.sample_stride : 1;
    int n_my = 0;
    for (long long s = p.sample_offset; s < p.spp; s += stride) n_my++;
    const long long total = (long long) npix * n_my;
    // batch: 2^22 samples (12.4 GB of wave state; one host read per bounce, so few large batches) unless the caller says
    // otherwise; halved until it fits half of the free memory
    long long batch = p.max_batch_paths > 0 ? p.max_batch_paths : (1ll << 22);
    batch = std::max<long long>(1, std::min(batch, std::max<long long>(total, 1)));
    {
        size_t free_b = 0, total_b = 0;
        B2_CUDA(cudaMemGetInfo(&free_b, &total_b));
        const size_t per_slot = MW_F4_PER_SLOT * sizeof(float4) + 3 * sizeof(int) + (size_t) (sc.n_lights + B2_MIS_MAXV) * sizeof(uint2);
        const size_t have = ws.mw_f4.n * sizeof(float4) + ws.mw_i32.n * sizeof(int) + ws.mw_pairs.n * sizeof(uint2);
        while (batch > 4096 && per_slot * (size_t) batch > (free_b + have) / 2) batch /= 2;
    }
    B2_TRY(ws.ensure_mis_wave((int) batch, sc.n_lights, film_floats));
    const size_t sb = stack_bytes(MW_THREADS, bvh_depth);
    if (sb > 200 * 1024) { b2_set_error("BVH depth %d needs more shared memory for traversal stacks than a CTA has", bvh_depth); return B2_ERR_STATE; }
    {
        static int configured[64];
        int dev = 0;
        B2_CUDA(cudaGetDevice(&dev));
        if (dev >= 64 || bvh_depth + 1 > configured[dev]) {
            B2_CUDA(cudaFuncSetAttribute(k_mw_light<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb));
            B2_CUDA(cudaFuncSetAttribute(k_mw_light<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb));
            B2_CUDA(cudaFuncSetAttribute(k_mw_eye<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb));
            B2_CUDA(cudaFuncSetAttribute(k_mw_eye<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb));
            B2_CUDA(cudaFuncSetAttribute(k_mw_splat<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb));
            B2_CUDA(cudaFuncSetAttribute(k_mw_splat<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb));
            B2_CUDA(cudaFuncSetAttribute(k_mw_connect<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb));
            B2_CUDA(cudaFuncSetAttribute(k_mw_connect<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sb));
            if (dev < 64) configured[dev] = bvh_depth + 1;
        }
    }
    RenderCtx c; b2_setup_ctx(c, sc, cam, p, ws, 1);
    const bool count = p.count_traversal != 0;
    const CamFrame cf = cam_frame(c.cdx, c.cdy);

    MisWave w;
    {
        const size_t N = (size_t) batch;
        float4 *f = ws.mw_f4.p;
        w.N = (int) batch;
        w.l_ro = f; f += N; w.l_rd = f; f += N; w.l_alpha = f; f += N; w.l_int = (int4 *) f; f += N;
        w.lv = f; f += (size_t) B2_MIS_MAXV * MW_LV_ROWS * N;
        w.e_ro = f; f += N; w.e_rd = f; f += N; w.e_thr = f; f += N; w.e_ppos = f; f += N; w.e_pn = f; f += N; w.e_int = (int4 *) f; f += N;
        w.e_hist = f; f += (size_t) B2_MIS_MAXV * N;
        w.cv = f; f += (size_t) MW_CV_ROWS * N;
        int *i = ws.mw_i32.p;
        w.ints = i; i += MW_N_INTS; w.l_count = i; i += N; w.live = i; i += N; w.live_next = i;
        w.pairs = ws.mw_pairs.p;
    }
    uint64_t n_launches = 0, n_batches = 0, n_waves = 0;
    CudaEvent ev0, ev1, ev_wave;
    B2_CUDA(ev0.create()); B2_CUDA(ev1.create()); B2_CUDA(ev_wave.create(cudaEventDisableTiming));
    B2_CUDA(cudaEventRecord(ev0, st));
    B2_CUDA(cudaMemsetAsync(ws.film.p, 0, sizeof(float) * film_floats, st));
    B2_CUDA(cudaMemsetAsync(ws.counters.p, 0, sizeof(unsigned long long) * B2_CNT_N, st));
    int *h = ws.h_ints;
    auto read_ints = [&]() -> int {                       // the wave's counters, through the pinned slot; waits for the wave
        B2_CUDA(cudaMemcpyAsync(h, w.ints, MW_N_INTS * sizeof(int), cudaMemcpyDeviceToHost, st));
        B2_CUDA(cudaEventRecord(ev_wave, st));
        B2_CUDA(cudaEventSynchronize(ev_wave));
        return B2_OK;
    };
    const int conn_grid = sc.n_sms * 16;
    for (long long w0 = 0; w0 < total && sc.n_lights > 0; w0 += batch) {
        const int n = (int) std::min<long long>(batch, total - w0);
        const uint32_t pix0 = (uint32_t) (w0 % npix); const int s0 = (int) (w0 / npix);
        n_batches++;
        B2_CUDA(cudaMemsetAsync(w.ints, 0, MW_N_INTS * sizeof(int), st));
        // ---- light subpaths, bounce by bounce (a stored vertex per bounce: at most B2_MIS_MAXV bounces) ----
        int n_live = n;
        for (int step = 0; n_live > 0 && step <= B2_MIS_MAXV; ++step) {
            if (step > 0) B2_CUDA(cudaMemsetAsync(&w.ints[MW_N_LIVE_NEXT], 0, sizeof(int), st));
            if (count) k_mw_light<true><<<blocks_of(n_live, MW_THREADS), MW_THREADS, sb, st>>>(c, cf, w, n_live, step, pix0, s0);
            else k_mw_light<false><<<blocks_of(n_live, MW_THREADS), MW_THREADS, sb, st>>>(c, cf, w, n_live, step, pix0, s0);
            n_launches++; n_waves++;
            B2_TRY(read_ints());
            n_live = h[MW_N_LIVE_NEXT];
            std::swap(w.live, w.live_next);
        }
        {   // the camera connections of every stored light vertex that lies inside the frustum (records gathered by the light bounces)
            const long long n_rec = h[MW_N_PAIRS];
            if (n_rec > 0) {
                if (count) k_mw_splat<true><<<(unsigned) std::min<long long>(blocks_of(n_rec, MW_THREADS), conn_grid), MW_THREADS, sb, st>>>(c, cf, w);
                else k_mw_splat<false><<<(unsigned) std::min<long long>(blocks_of(n_rec, MW_THREADS), conn_grid), MW_THREADS, sb, st>>>(c, cf, w);
                n_launches++;
            }
        }
        const int J = sc.n_lights + std::min(h[MW_S_MAX], B2_MIS_MAXV);
        // ---- eye subpaths: a bounce, then its vertices against the lights and the light vertices of their own samples ----
        n_live = n;
        for (int step = 0; n_live > 0 && step <= B2_MIS_MAXV; ++step) {
            B2_CUDA(cudaMemsetAsync(w.ints, 0, 3 * sizeof(int), st));       // n_live_next, n_cv, n_pairs
            if (count) k_mw_eye<true><<<blocks_of(n_live, MW_THREADS), MW_THREADS, sb, st>>>(c, cf, w, n_live, step, pix0, s0);
            else k_mw_eye<false><<<blocks_of(n_live, MW_THREADS), MW_THREADS, sb, st>>>(c, cf, w, n_live, step, pix0, s0);
            const long long cand = (long long) n_live * J;                   // n_pairs <= n_cv * J <= n_live * J
            if (count) k_mw_connect<true><<<(unsigned) std::min<long long>(blocks_of(cand, MW_THREADS), conn_grid), MW_THREADS, sb, st>>>(c, w);
            else k_mw_connect<false><<<(unsigned) std::min<long long>(blocks_of(cand, MW_THREADS), conn_grid), MW_THREADS, sb, st>>>(c, w);
            n_launches += 2; n_waves++;
            B2_TRY(read_ints());
            n_live = h[MW_N_LIVE_NEXT];
            std::swap(w.live, w.live_next);
        }
    }
    B2_CUDA(cudaGetLastError());
    const float inv = p.film_mode == B2_FILM_SUM ? 1.0f : 1.0f / (float) p.spp;
    k_mw_film_out<<<blocks_of((long long) film_floats, 256), 256, 0, st>>>(ws.film.p, d_film_out, film_floats, inv, p.film_mode == B2_FILM_ACCUMULATE);
    n_launches++;
    B2_CUDA(cudaGetLastError());
    B2_CUDA(cudaEventRecord(ev1, st));
    unsigned long long h_cnt[B2_CNT_N];
    B2_CUDA(cudaMemcpyAsync(h_cnt, ws.counters.p, sizeof(h_cnt), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev0, ev1);
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        stats->eye_samples = (unsigned long long) total;
        stats->closest_rays = h_cnt[B2_CNT_CLOSEST]; stats->shadow_rays = h_cnt[B2_CNT_SHADOW]; stats->path_samples = h_cnt[B2_CNT_PATH_SAMPLES];
        stats->delta_cap_hits = h_cnt[B2_CNT_DELTA_CAP];
        stats->nodes_visited = h_cnt[B2_CNT_NODES]; stats->tris_tested = h_cnt[B2_CNT_TRIS];
        stats->n_batches = n_batches; stats->n_waves = n_waves; stats->n_launches = n_launches;
        stats->ms_total = ms;
    }
    return B2_OK;
}
