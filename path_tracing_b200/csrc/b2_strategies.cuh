// b2_strategies.cuh -- sampling strategies that are NOT part of the reference estimator (B2_STRATEGIES_* > 0).
//
// cpu_bdpt.cpp connects every eye vertex with a shared, one-shot pool of light vertices and has no t = 1 strategy
// (SURVEY.md hard part 1); the north_star asks for "atomic splatting of t=1 light-tracing contributions into the film".
// That strategy is built here as its own, separately validated estimator, off in parity mode:
//
//   light tracing (t = 1) : W*H*spp light subpaths per render; every non-delta vertex is connected to the pinhole and
//                           its contribution is splatted into the pixel it projects to with atomicAdd (k_light_trace)
//   next-event estimation : the eye subpaths of the wavefront (k_extend, unchanged) with every eye vertex connected to
//        (s = 1)            every light itself, weight 1 (k_nee) -- the independent estimator the splats are validated
//                           against: both are unbiased for the same integral, so their images must agree
//   B2_STRATEGIES_FULL    : the two combined with constant weights 1/2 + 1/2 (a valid, if not optimal, MIS combination:
//                           each path of >= 1 surface vertex can be sampled by both), plus the depth-0 emitter hits
//
// Light model of these modes (the reference's connection code treats the sampled emission direction as the light's
// "normal", which is not a physical model, so one is fixed here): light k is a point source at `pos` with total power
// `illum`, uniform intensity I = illum / Omega inside the cone of half-angle min(cutoff, pi) around `dir`,
// Omega = 2 pi (1 - cos(cutoff)); rays leave from the ball surface pos + r * omega as in cpu_bdpt.cpp:253.  Parallel lights
// are ignored by these modes.  The BSDF is the reference's (b2_bsdf.cuh) with ONE correction, the normalised GGX D
// (mat_mark_fix: with the reference's D, bsdf_pdf is not the density of bsdf_sample and every multi-bounce estimator is
// biased in its own way; the direct-light-only comparison of the two strategies passes either way).  The BSDF is not
// reciprocal (Schlick on wo only), so the light subpaths evaluate it in transport order, f(wo = towards the eye side,
// wi = towards the light side).
#pragma once
#include "b2_bdpt.cuh"

#define B2_RNG_LTRACE 3u

struct SpotLight { float3 pos, dir, intensity; float r, cos_cut; bool parallel; };
B2_HD SpotLight spot_light(const DevScene &sc, int li) {
    const float4 l0 = sc.lights[4 * li], l1 = sc.lights[4 * li + 1], l2 = sc.lights[4 * li + 2], l3 = sc.lights[4 * li + 3];
    SpotLight L;
    L.pos = mk3(l0); L.r = l0.w; L.dir = mk3(l3); L.parallel = l2.w != 0.0f;
    L.cos_cut = l1.w >= B2_PI ? -1.0f : cosf(l1.w);
    const float omega = 2.0f * B2_PI * (1.0f - L.cos_cut);
    L.intensity = omega > 0.0f ? mk3(l2) * (1.0f / omega) : mk3(0, 0, 0);
    return L;
}
B2_HD bool is_delta_material(const Mat &m) {                         // the two delta branches of bsdf_sample (b2_bsdf.cuh)
    return (m.eta > 0.0f && mat_rough(m) < 0.001f && m.metallic < 0.01f) || (m.metallic > 0.99f && mat_rough(m) < 0.001f);
}

// Pinhole projection: pixel of world point p, cosine at the lens, squared distance, direction p -> eye.
struct CamFrame { float3 fwd; float inv_dx2, inv_dy2, pix_area; };
inline CamFrame cam_frame(float3 cdx, float3 cdy) {
    CamFrame f;
    const float3 c = make_float3(cdx.y * cdy.z - cdx.z * cdy.y, cdx.z * cdy.x - cdx.x * cdy.z, cdx.x * cdy.y - cdx.y * cdy.x);
    const float lc = sqrtf(c.x * c.x + c.y * c.y + c.z * c.z);
    f.fwd = make_float3(c.x / lc, c.y / lc, c.z / lc);             // cdx x cdy = -|dx||dy| cw: the viewing direction
    const float dx2 = cdx.x * cdx.x + cdx.y * cdx.y + cdx.z * cdx.z, dy2 = cdy.x * cdy.x + cdy.y * cdy.y + cdy.z * cdy.z;
    f.inv_dx2 = 1.0f / dx2; f.inv_dy2 = 1.0f / dy2; f.pix_area = sqrtf(dx2) * sqrtf(dy2);
    return f;
}
B2_HD bool camera_project(const RenderCtx &c, const CamFrame &cf, float3 p, uint32_t &pixel, float &cos_c, float &d2, float3 &to_eye) {
    const float3 v = p - c.cam_eye;
    const float z = dot3(v, cf.fwd);
    if (!(z > 1e-6f)) return false;
    d2 = dot3(v, v);
    const float inv_d = fast_rsqrt(d2);
    cos_c = z * inv_d;
    to_eye = v * -inv_d;
    const float3 rel = c.cam_eye + v * fast_rcp(z) - c.cUL;         // on the image plane at distance 1, relative to its upper-left corner
    const float px = dot3(rel, c.cdx) * cf.inv_dx2, py = dot3(rel, c.cdy) * cf.inv_dy2;
    if (!(px >= 0.0f && py >= 0.0f && px < (float) c.W && py < (float) c.H)) return false;
    pixel = (uint32_t) py * (uint32_t) c.W + (uint32_t) px;
    return true;
}

// One light subpath with t = 1 splats.  `path` is the global index of the subpath (stream id and light choice), `alpha0_scale`
// = n_lights / (W*H) * weight: with N_L = W*H*spp subpaths and a film that is divided by spp at resolve, every subpath of
// light k starts with illum_k * alpha0_scale (derivation in DESIGN.md section 9).
template <bool COUNT>
B2_HD void light_trace_body(const RenderCtx &c, const CamFrame &cf, unsigned long long path, float alpha0_scale, uint2 *stack, int sstride,
                            TravCounters *tc, uint32_t &n_closest, uint32_t &n_shadow, uint32_t &n_splat) {
    const DevScene &sc = c.sc;
    const int li = (int) (path % (unsigned long long) sc.n_lights);
    const SpotLight L = spot_light(sc, li);
    if (L.parallel) return;
    const uint32_t id0 = (uint32_t) path, id1 = (uint32_t) (path >> 32);
    float r1, r2;
    rng_draw2(c.key, B2_RNG_LTRACE, id0, id1, r1, r2);
    float3 u_vec = (fabsf(L.dir.x) > 0.9f) ? mk3(0, 1, 0) : mk3(1, 0, 0);
    const float3 v_vec = normalize3(cross3(L.dir, u_vec));
    u_vec = normalize3(cross3(v_vec, L.dir));
    const float ct = 1.0f - r1 * (1.0f - L.cos_cut), st = fast_sqrt(fmaxf(0.0f, 1.0f - ct * ct));
    float sphi, cphi;
    sincos_2pi(r2, sphi, cphi);
    float3 rd = normalize3(u_vec * (st * cphi) + v_vec * (st * sphi) + L.dir * ct);
    float3 ro = L.pos + rd * L.r;
    float3 alpha = mk3(sc.lights[4 * li + 2]) * alpha0_scale;
    float eta = 1.0f;
    int n_delta = 0;
    BsdfSample bs; bs.wi = mk3(0, 0, 0); bs.f = mk3(0, 0, 0); bs.pdf = 0.0f;
    for (int depth = 1; depth < c.light_depth; ++depth) {
        const HitRec h = closest_hit_bvh<COUNT>(sc, ro, rd, stack, sstride, tc);
        n_closest++;
        if (h.prim == B2_PRIM_MISS) break;
        const Surface s = surface_at(sc, ro, rd, h);
        if (s.is_light) break;
        const float3 to_light = rd * -1.0f;
        if (!is_delta_material(s.mtl)) {                                    // ---- t = 1: connect to the pinhole and splat ----
            uint32_t pixel; float cos_c, d2; float3 to_eye;
            if (camera_project(c, cf, s.pos, pixel, cos_c, d2, to_eye)) {
                const float cos_y = dot3(s.n, to_eye);
                if (cos_y > 0.0f) {
                    const float3 f = bsdf_evaluate(s.mtl, to_eye, to_light, s.n);
                    if (f.x > 0.0f || f.y > 0.0f || f.z > 0.0f) {
                        const ShadowRay sr = make_shadow_ray(x3_add(s.pos, x3_muls(s.n, B2_EPSILON)), c.cam_eye);
                        n_shadow++;
                        if (!occluded_bvh<COUNT>(sc, sr.o, sr.d, sr.tmax, stack, sstride, tc)) {
                            const float g = fast_div(cos_y, d2 * cos_c * cos_c * cos_c * cf.pix_area);
                            const float3 v = alpha * f * g;
                            if (valid_color(v)) {
                                atomic_add_f32(&c.film[3 * pixel], v.x); atomic_add_f32(&c.film[3 * pixel + 1], v.y); atomic_add_f32(&c.film[3 * pixel + 2], v.z);
                                n_splat++;
                            }
                        }
                    }
                }
            }
        }
        float u_rr, u1, u2;
        rng_draw3(c.key, B2_RNG_LTRACE, id0, id1, (uint32_t) (depth - 1 + n_delta), u_rr, u1, u2);
        bsdf_sample(s.mtl, to_light, s.n, u_rr, u1, u2, eta, bs);
        if (bs.pdf <= 0.0f && !bs.is_delta) break;
        if (bs.is_delta) {
            alpha = alpha * bs.f;
            rd = bs.wi; eta = bs.new_eta;
            ro = s.pos + s.n * (dot3(bs.wi, s.n) < 0.0f ? -B2_EPSILON : B2_EPSILON);
            if (++n_delta >= c.max_delta) break;
            depth--;
            continue;
        }
        const float3 f_adj = bsdf_evaluate(s.mtl, bs.wi, to_light, s.n);   // transport order: wo = towards the eye side
        alpha = alpha * f_adj * fabsf(dot3(s.n, bs.wi)) / bs.pdf;
        if (!valid_color(alpha) || (alpha.x <= 0.0f && alpha.y <= 0.0f && alpha.z <= 0.0f)) break;
        rd = bs.wi;
        ro = s.pos + s.n * B2_EPSILON;
    }
}

// s = 1: one eye vertex against every (non-parallel) light, weight `w`.  Returns the sum of the contributions.
template <bool COUNT>
B2_HD float3 nee_body(const RenderCtx &c, const EyeVtx &ev, float w, uint2 *stack, int sstride, TravCounters *tc, uint32_t &n_shadow, uint32_t &n_path) {
    float3 sum = mk3(0, 0, 0);
    if (is_delta_material(ev.mtl)) return sum;
    for (int li = 0; li < c.sc.n_lights; ++li) {
        const SpotLight L = spot_light(c.sc, li);
        if (L.parallel) continue;
        const float3 d = ev.pos - L.pos;
        const float d2 = dot3(d, d);
        if (d2 < 1e-12f) continue;
        const float3 om = d * fast_rsqrt(d2);                               // light -> vertex
        if (dot3(om, L.dir) < L.cos_cut) continue;
        const float3 wi = om * -1.0f;
        const float cos_x = dot3(ev.n, wi);
        if (cos_x <= 0.0f) continue;
        const float3 f = bsdf_evaluate(ev.mtl, ev.wo, wi, ev.n);
        if (f.x <= 0.0f && f.y <= 0.0f && f.z <= 0.0f) continue;
        const ShadowRay sr = make_shadow_ray(x3_add(ev.pos, x3_muls(ev.n, B2_EPSILON)), L.pos + om * L.r);
        n_shadow++;
        if (occluded_bvh<COUNT>(c.sc, sr.o, sr.d, sr.tmax, stack, sstride, tc)) continue;
        const float3 v = ev.thr * f * L.intensity * (w * fast_div(cos_x, d2));
        if (valid_color(v)) { sum = sum + v; n_path++; }
    }
    return sum;
}
