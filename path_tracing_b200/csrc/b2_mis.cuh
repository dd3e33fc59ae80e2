// b2_mis.cuh -- B2_STRATEGIES_MIS: bidirectional path tracing with the full strategy set and real densities
// (SURVEY.md 8(f) rank 2).  What cpu_bdpt.cpp leaves out or gets wrong is done here the textbook way (Veach 1997, ch. 10):
//
//   * one light subpath PER EYE SAMPLE (the reference shares one pool of n_lights * spl subpaths between all pixels,
//     cpu_bdpt.cpp:218-325, so its pool noise never averages out);
//   * every strategy of a path x_0 (light) x_1 .. x_n (surfaces) x_{n+1} (pinhole): m = 0 next-event estimation,
//     0 < m < n connection of light vertex x_m with eye vertex x_{m+1}, m = n light tracing with a film splat (the
//     reference has neither m = 0 nor m = n);
//   * area densities of BOTH directions at every vertex when a connection is weighted (the reference weights with the
//     placeholder pdf_fwd = 0 of the vertex being connected, :385, and an unnormalised direction, :431);
//   * balance heuristic over exactly the strategies that can produce the path under the depth limits and delta surfaces.
//
// Same light model, BSDF (normalised GGX, transport order) and camera measurement as b2_strategies.cuh, so the image has the
// same expectation as B2_STRATEGIES_NEE at equal eye_depth -- that is the test (tests/test_gpu_strategies.py): there is
// no oracle for a mode the reference does not have.
//
// Path space of the mode: at most `eye_depth` non-delta surface vertices per path (delta vertices are free, capped by
// max_delta); a light subpath carries at most light_depth - 1 of them (vertex 0 of a reference light path is the light).
// Shadow rays are blocked by every object (the reference lets them through eta > 0 surfaces, cpu_bdpt.cpp:97; with real
// light subpaths that energy arrives through the refracting light paths instead).  Lights: one of the n_lights is drawn
// uniformly per light subpath; next-event estimation visits every light.  Parallel lights are ignored, as in
// b2_strategies.cuh.
//
// Weights.  Strategy m produces the path with density  q_m = c_m * prod_{i<=m} pL_i * prod_{i>m} pE_i ,  pL_i / pE_i = area
// density of x_i when it is generated from the light / eye side, c_m = (m ? 1/n_lights : 1) * (m == n ? W*H : 1): the
// W*H*spp light subpaths of a render may all splat into the pixel, its own spp eye samples carry the other strategies.
// A density through a delta surface is a Dirac factor that appears once in either direction and is replaced by 1 in both
// (any function of the PATH that is positive exactly on the feasible strategies keeps the estimator unbiased; q only has
// to be the same whichever strategy evaluates it).  w_k = q_k / sum_m q_m, accumulated as ratios from k outwards.
#pragma once
#include "b2_strategies.cuh"

#define B2_RNG_MIS_LIGHT 6u
#ifndef B2_MIS_MAXV
#define B2_MIS_MAXV 24                   /* stored vertices per subpath, delta ones included */
#endif

struct MisVtx { float3 pos, n, wo, beta; Mat mtl; float pdf_fwd, pdf_rev; int delta, nd; };   // nd = non-delta vertices up to and including this one

// w_k of a path of n surface vertices: pL / pE / dl indexed 1..n (index 0 and n + 1 unused).
B2_NOINLINE float mis_balance(const float *pL, const float *pE, const unsigned char *dl, int n, int k, int l_cap, float inv_nl, float wh) {
    const float ck = (k ? inv_nl : 1.0f) * (k == n ? wh : 1.0f);
    float sum = 1.0f, r = 1.0f;
    int nd = 0;
    for (int i = 1; i <= k; ++i) nd += dl[i] ? 0 : 1;
    for (int m = k + 1; m <= n; ++m) {                                 // more light vertices
        const float a = (m >= 2 && dl[m - 1]) ? 1.0f : pL[m], b = (m + 1 <= n && dl[m + 1]) ? 1.0f : pE[m];
        if (!(a > 0.0f)) break;
        if (!(b > 0.0f)) return 0.0f;
        r *= a / b;
        if (!dl[m] && ++nd > l_cap) break;
        if (!dl[m] && (m == n || !dl[m + 1])) sum += r * ((m == n ? inv_nl * wh : inv_nl) / ck);
    }
    r = 1.0f;
    for (int m = k - 1; m >= 0; --m) {                                 // more eye vertices
        const int i = m + 1;
        const float a = (i + 1 <= n && dl[i + 1]) ? 1.0f : pE[i], b = (i >= 2 && dl[i - 1]) ? 1.0f : pL[i];
        if (!(a > 0.0f)) break;
        if (!(b > 0.0f)) return 0.0f;
        r *= a / b;
        if ((m == 0 || !dl[m]) && !dl[m + 1]) sum += r * ((m ? inv_nl : 1.0f) / ck);
    }
    const float w = 1.0f / sum;
    return (w > 0.0f && w <= 1.0f) ? w : 0.0f;                          // NaN / Inf / overflowed ratios drop the sample
}

B2_HD float area_pdf(float pdf_omega, float3 n_to, float3 dir, float d2) { return pdf_omega * fabsf(dot3(n_to, dir)) / fmaxf(d2, 1e-12f); }
B2_HD void film_add(const RenderCtx &c, uint32_t pixel, float3 v) {
    if (v.x != 0.0f) atomic_add_f32(&c.film[3 * pixel], v.x);
    if (v.y != 0.0f) atomic_add_f32(&c.film[3 * pixel + 1], v.y);
    if (v.z != 0.0f) atomic_add_f32(&c.film[3 * pixel + 2], v.z);
}
// camera density of a surface point seen from the pinhole, per unit area: the pixel footprint (b2_strategies.cuh / DESIGN 9)
B2_HD float camera_area_pdf(const CamFrame &cf, float cos_y, float cos_c, float d2) { return cos_y / fmaxf(d2 * cos_c * cos_c * cos_c * cf.pix_area, 1e-20f); }

struct MisCounters { uint32_t closest, shadow, paths; };

// Out-of-line copies of the big pieces: the body needs each of them at two places (light walk, eye walk), and inlined the
// kernel was 9 900 SASS instructions that did not fit the instruction cache (ncu: stall_no_instruction 18.3 per issue).
template <bool COUNT>
B2_NOINLINE HitRec mis_closest(const DevScene &sc, float3 ro, float3 rd, uint2 *stack, int sstride, TravCounters *tc) {
    return closest_hit_bvh<COUNT>(sc, ro, rd, stack, sstride, tc);
}
template <bool COUNT>
B2_NOINLINE bool mis_occluded(const DevScene &sc, float3 p1, float3 p2, uint2 *stack, int sstride, TravCounters *tc) {
    const ShadowRay sr = make_shadow_ray(p1, p2);
    return occluded_bvh<COUNT, true>(sc, sr.o, sr.d, sr.tmax, stack, sstride, tc);
}
B2_NOINLINE void mis_bsdf_sample(const Mat &m, float3 wo, float3 n, float u_rr, float u1, float u2, float eta, BsdfSample &bs) { bsdf_sample(m, wo, n, u_rr, u1, u2, eta, bs); }
B2_NOINLINE Surface mis_surface(const DevScene &sc, float3 ro, float3 rd, HitRec h) { return surface_at(sc, ro, rd, h); }

// The body is written warp-synchronously: every lane of a warp runs every loop until no lane has work left (B2_WARP_ANY),
// lanes without work are predicated off, and the expensive parts (closest hit, BSDF + weight, shadow ray) sit at one place of
// the loop each so that the lanes that need them run them together.  A first version with per-lane `break` / `continue`
// through nested loops ran with 3.7 of 32 lanes active (ncu: smsp__thread_inst_executed_per_inst_executed, profiles/
// r01_mis_summary.md).  On the host (tests/hostsim) a "warp" is one lane.
#if defined(__CUDA_ARCH__)
#define B2_WARP_ANY(p) (__any_sync(0xffffffffu, (p)) != 0)
#else
#define B2_WARP_ANY(p) (p)
#endif

// One eye sample (pixel, sidx): its light subpath with the m = n splats, its eye subpath with the m = 0 .. n - 1 strategies.
// `valid` = false for the lanes past the end of a batch: they only take part in the warp votes.
template <bool COUNT>
B2_HD void mis_sample_body(const RenderCtx &c, const CamFrame &cf, bool valid, uint32_t pixel, uint32_t sidx, MisVtx *lv, uint2 *stack, int sstride,
                           TravCounters *tc, MisCounters &cnt) {
    const DevScene &sc = c.sc;
    const int n_cap = c.eye_depth, l_cap = c.light_depth - 1;
    const float inv_nl = 1.0f / (float) sc.n_lights, wh = (float) c.W * (float) c.H;
    float pL[2 * B2_MIS_MAXV + 2], pE[2 * B2_MIS_MAXV + 2];
    unsigned char dl[2 * B2_MIS_MAXV + 2];
    BsdfSample bs; bs.wi = mk3(0, 0, 0); bs.f = mk3(0, 0, 0); bs.pdf = 0.0f;
    float3 ro = mk3(0, 0, 0), rd = mk3(0, 0, 1);

    // ---------------- light subpath ----------------
    int S = 0;
    {
        bool alive = valid && l_cap > 0 && n_cap > 0;
        float3 alpha = mk3(0, 0, 0);
        float eta = 1.0f, last_pdf = 0.0f, first_off = 0.0f;
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
            alpha = mk3(sc.lights[4 * li + 2]) * (float) sc.n_lights;                 // I / (pdf_omega = 1 / Omega) / P(light) = illum * n_lights
            last_pdf = alive ? 1.0f / omega : 0.0f; first_off = L.r;
        }
        bool prev_delta = false;
        int nd = 0, n_delta = 0;
        for (int step = 0; B2_WARP_ANY(alive); ++step) {
            HitRec h; h.prim = B2_PRIM_MISS; h.t = 0.0f;
            if (alive) { h = mis_closest<COUNT>(sc, ro, rd, stack, sstride, tc); cnt.closest++; }
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
                lv[S++] = v;
                if (!v.delta && nd <= n_cap && camera_project(c, cf, v.pos, px, cos_c, d2c, to_eye)) { cos_y = dot3(v.n, to_eye); splat = cos_y > 0.0f; }
            }
            float w = 0.0f;
            float3 f = mk3(0, 0, 0);
            if (splat) {
                f = bsdf_evaluate(v.mtl, to_eye, v.wo, v.n);
                if (f.x > 0.0f || f.y > 0.0f || f.z > 0.0f) {
                    const int n = S;
                    for (int i = 1; i <= n; ++i) { pL[i] = lv[i - 1].pdf_fwd; pE[i] = lv[i - 1].pdf_rev; dl[i] = (unsigned char) lv[i - 1].delta; }
                    pE[n] = camera_area_pdf(cf, cos_y, cos_c, d2c);
                    if (n >= 2) { const float3 dv = lv[n - 2].pos - v.pos; pE[n - 1] = area_pdf(bsdf_pdf(v.mtl, to_eye, v.wo, v.n), lv[n - 2].n, v.wo, dot3(dv, dv)); }
                    w = mis_balance(pL, pE, dl, n, n, l_cap, inv_nl, wh);
                }
            }
            if (w > 0.0f) {
                cnt.shadow++;
                if (!mis_occluded<COUNT>(sc, x3_add(v.pos, x3_muls(v.n, B2_EPSILON)), c.cam_eye, stack, sstride, tc)) {
                    const float3 val = alpha * f * (w * camera_area_pdf(cf, cos_y, cos_c, d2c) / wh);
                    if (valid_color(val)) { film_add(c, px, val); cnt.paths++; }
                }
            }
            if (alive && ((!v.delta && nd == l_cap) || S >= B2_MIS_MAXV)) alive = false;   // nothing behind the last allowed vertex can connect
            if (alive) {
                float u_rr, u1, u2;
                rng_draw3(c.key, B2_RNG_MIS_LIGHT, pixel, sidx, (uint32_t) step, u_rr, u1, u2);
                mis_bsdf_sample(s.mtl, v.wo, s.n, u_rr, u1, u2, eta, bs);
                if (bs.pdf <= 0.0f) alive = false;
            }
            if (alive) {
                if (S >= 2) {                                                           // density of the previous vertex seen from this one
                    const float3 dv = lv[S - 2].pos - v.pos;
                    lv[S - 2].pdf_rev = bs.is_delta ? 1.0f : area_pdf(bsdf_pdf(s.mtl, bs.wi, v.wo, s.n), lv[S - 2].n, v.wo, dot3(dv, dv));
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
        }
    }

    // ---------------- eye subpath ----------------
    float e_fwd[B2_MIS_MAXV], e_rev[B2_MIS_MAXV];
    unsigned char e_dl[B2_MIS_MAXV];
    float3 own = mk3(0, 0, 0);
    bool alive = valid && n_cap > 0;
    float cos_cam = 1.0f;
    if (alive) {
        float jx, jy;
        rng_draw2(c.key, B2_RNG_EYE, pixel, sidx, jx, jy);
        uint32_t col;
        const uint32_t row = udiv_magic(pixel, (uint32_t) c.W, c.w_magic, col);
        const float3 pp = c.cUL + c.cdx * ((float) col + jx) + c.cdy * ((float) row + jy);
        rd = normalize3(pp - c.cam_eye); ro = c.cam_eye;
        cos_cam = dot3(rd, cf.fwd);
    }
    float3 thr = mk3(1.0f, 1.0f, 1.0f), prev_pos = c.cam_eye, prev_n = rd;
    float eta = 1.0f, last_pdf = 1.0f;
    bool prev_delta = false;
    int T = 0, nd_e = 0, n_delta = 0;
    const int J = sc.n_lights + S;                                                          // candidates of an eye vertex: the lights, then the light vertices
    for (int step = 0; B2_WARP_ANY(alive); ++step) {
        HitRec h; h.prim = B2_PRIM_MISS; h.t = 0.0f;
        if (alive) { h = mis_closest<COUNT>(sc, ro, rd, stack, sstride, tc); cnt.closest++; }
        Surface s; s.is_light = false; s.pos = s.n = mk3(0, 0, 0); s.mtl.base = mk3(0, 0, 0); s.mtl.roughness = s.mtl.metallic = s.mtl.eta = 0.0f;
        if (alive && h.prim == B2_PRIM_MISS) alive = false;
        if (alive) {
            s = mis_surface(sc, ro, rd, h);
            if (s.is_light) { if (T == 0) { own = own + s.mtl.base; cnt.paths++; } alive = false; }   // the ball of a light is only ever seen directly
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
            e_fwd[T] = T == 0 ? camera_area_pdf(cf, fabsf(dot3(s.n, wo)), cos_cam, d2prev) : (prev_delta ? 1.0f : area_pdf(last_pdf, s.n, rd, d2prev));
            e_rev[T] = 0.0f; e_dl[T] = delta ? 1 : 0;
            t = ++T;                                                                         // this is eye vertex z_t
        }
        // ---- this vertex against every light (m = 0) and every light vertex (0 < m < n) ----
        const bool connects = alive && !delta;
        int j = 0;
        for (;;) {
            // cheap part: the next candidate that passes the geometric tests
            bool have = false;
            int m = 0; float d2 = 1.0f, cos_z = 0.0f, cos_y = 0.0f; float3 wz = mk3(0, 0, 1), target = mk3(0, 0, 0), scale = mk3(0, 0, 0);
            while (connects && j < J && !have) {
                const int cand = j++;
                if (cand < sc.n_lights) {
                    const SpotLight L = spot_light(sc, cand);
                    if (L.parallel) continue;
                    const float3 d = s.pos - L.pos;
                    d2 = dot3(d, d);
                    if (d2 < 1e-12f) continue;
                    const float3 om = d * fast_rsqrt(d2);
                    if (dot3(om, L.dir) < L.cos_cut) continue;
                    wz = om * -1.0f;
                    cos_z = dot3(s.n, wz);
                    if (cos_z <= 0.0f) continue;
                    m = 0; target = L.pos + om * L.r; scale = L.intensity;
                    cos_y = 1.0f / (2.0f * B2_PI * (1.0f - L.cos_cut));                        // m = 0: the emission density rides in cos_y
                    have = true;
                } else {
                    const MisVtx &y = lv[cand - sc.n_lights];
                    if (y.delta || y.nd + nd_e > n_cap) continue;
                    const float3 dv = y.pos - s.pos;
                    d2 = dot3(dv, dv);
                    if (d2 < 1e-10f) continue;
                    wz = dv * fast_rsqrt(d2);
                    cos_z = dot3(s.n, wz); cos_y = dot3(y.n, wz * -1.0f);
                    if (cos_z <= 0.0f || cos_y <= 0.0f) continue;
                    m = cand - sc.n_lights + 1; target = x3_add(y.pos, x3_muls(y.n, B2_EPSILON)); scale = y.beta;
                    have = true;
                }
            }
            if (!B2_WARP_ANY(have)) break;
            // expensive part, lanes with a candidate together: BSDFs, weight, shadow ray
            float w = 0.0f;
            float3 val = mk3(0, 0, 0);
            if (have) {
                const float3 fz = bsdf_evaluate(s.mtl, wo, wz, s.n);
                const int n = m + t;
                if (fz.x > 0.0f || fz.y > 0.0f || fz.z > 0.0f) {
                    for (int i = m + 1; i <= n; ++i) { pE[i] = e_fwd[n - i]; pL[i] = e_rev[n - i]; dl[i] = e_dl[n - i]; }    // x_i = z_{n + 1 - i}
                    if (m + 2 <= n) pL[m + 2] = area_pdf(bsdf_pdf(s.mtl, wz, wo, s.n), prev_n, rd, d2prev);                  // z, lit from the light side, generates its eye-side neighbour
                    if (m == 0) {
                        pL[1] = area_pdf(cos_y, s.n, wz, d2);
                        val = thr * fz * scale * (cos_z / d2);
                        w = 1.0f;
                    } else {
                        const MisVtx &y = lv[m - 1];
                        const float3 wy = wz * -1.0f;
                        const float3 fy = bsdf_evaluate(y.mtl, wy, y.wo, y.n);
                        if (fy.x > 0.0f || fy.y > 0.0f || fy.z > 0.0f) {
                            for (int i = 1; i <= m; ++i) { pL[i] = lv[i - 1].pdf_fwd; pE[i] = lv[i - 1].pdf_rev; dl[i] = (unsigned char) lv[i - 1].delta; }
                            pL[m + 1] = area_pdf(bsdf_pdf(y.mtl, y.wo, wy, y.n), s.n, wz, d2);                                // y generates z
                            pE[m] = area_pdf(bsdf_pdf(s.mtl, wo, wz, s.n), y.n, wy, d2);                                      // z generates y
                            if (m >= 2) { const float3 dy = lv[m - 2].pos - y.pos; pE[m - 1] = area_pdf(bsdf_pdf(y.mtl, wy, y.wo, y.n), lv[m - 2].n, y.wo, dot3(dy, dy)); }
                            val = thr * fz * fy * scale * (cos_z * cos_y / d2);
                            w = 1.0f;
                        }
                    }
                    if (w > 0.0f) w = mis_balance(pL, pE, dl, n, m, l_cap, inv_nl, wh);
                }
            }
            if (w > 0.0f) {
                cnt.shadow++;
                if (!mis_occluded<COUNT>(sc, x3_add(s.pos, x3_muls(s.n, B2_EPSILON)), target, stack, sstride, tc)) {
                    val = val * w;
                    if (valid_color(val)) { own = own + val; cnt.paths++; }
                }
            }
        }
        if (alive && ((!delta && nd_e == n_cap) || T >= B2_MIS_MAXV)) alive = false;
        if (alive) {
            float u_rr, u1, u2;
            rng_draw3(c.key, B2_RNG_EYE, pixel, sidx, (uint32_t) step, u_rr, u1, u2);
            mis_bsdf_sample(s.mtl, wo, s.n, u_rr, u1, u2, eta, bs);
            if (bs.pdf <= 0.0f) alive = false;
        }
        if (alive) {
            if (t >= 2) e_rev[t - 2] = bs.is_delta ? 1.0f : area_pdf(bsdf_pdf(s.mtl, bs.wi, wo, s.n), prev_n, rd, d2prev);
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
    }
    if (valid) film_add(c, pixel, own);
}
