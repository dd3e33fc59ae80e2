// b2_pt.cuh -- per-item bodies of the two other integrators of the reference's command line:
//
//   pt   cuda_path_trace_kernel, src/pt_cu.cu:20-252  -- unidirectional path tracing with next-event estimation
//   ppm  ppm_eye_trace / ppm_photon_trace / ppm_resolve_image, src/ppm_cu.cu:64-322 -- one photon-mapping pass
//
// Both reuse the scene tables, the wide BVH, the BSDF and the Philox streams of the BDPT path.  What differs from
// cpu_bdpt.cpp is restated from the reference's GPU-side helpers (include/geometric.cuh:240-390, selected with
// DevScene::gpu_sem): ordinary spheres are hit from the inside too, and shadow rays are blocked by every object.  The
// reference has no CPU implementation of these integrators and seeds curand from the clock (pt_cu.cu:279,
// ppm_cu.cu:357), so parity is statistical: against oracle/bdpt_oracle.c's restatement on identical Philox streams, and
// against the reference's own kernels run on the same GPU (oracle/_ref/libref_cuda.so, tests/test_gpu_pt_ppm.py).
// Quirks kept (each cited below): the BSDF-hits-light MIS branch is a stub that never adds (pt_cu.cu:105-119), emitted
// radiance = illum / (ball area * cone fraction), the photon flux carries a factor n_lights (ppm_cu.cu:209), hit points
// are created on the first non-delta surface whatever the BSDF sample says, spp is ignored by ppm.
// Deviations (documented in DESIGN.md 10): the NEE point on a light ball is drawn with two uniforms instead of rejection
// sampling (same distribution); a delta-bounce cap (max_delta) bounds the reference's unbounded `depth--` loops; a delta
// sample with pdf 0 (total internal reflection) ends the path instead of continuing with stale outputs.
#pragma once
#include "b2_bdpt.cuh"

#define B2_RNG_PT_NEE 4u      // ctr = (pixel, sample, 1 + depth of the vertex, 4): light choice, two uniforms for the point on the ball
#define B2_RNG_PHOTON 5u      // ctr = (photon index, 0, block, 5): block 0 = emission, 1 + k = k-th bsdf_sample call
#define B2_PPM_RADIUS 0.05f   // PPM_RADIUS, include/ppm_cu.cuh:5

B2_HD void rng_draw4(uint2 key, uint32_t kind, uint32_t id0, uint32_t id1, uint32_t block, float &a, float &b, float &c, float &d) {
    const uint4 o = philox4x32_10(make_uint4(id0, id1, block, kind), key);
    a = u32_to_unit(o.x); b = u32_to_unit(o.y); c = u32_to_unit(o.z); d = u32_to_unit(o.w);
}
// the materials that take direct light / photons: not transparent, not a perfect mirror (pt_cu.cu:126, ppm_cu.cu:221)
B2_HD bool gathers_light(const Mat &m) { return m.eta <= 0.0f && (m.metallic < 0.99f || mat_rough(m) > 0.01f); }

// ---- path tracing ----
// Register state of a path between two bounces; in HBM: ray_o {o, eta}, ray_d {d, -}, thr {thr, depth | n_delta << 8 |
// last_is_delta << 31}, last_pos {pixel, sample, -, -} (bits).
struct PtState { float3 ro, rd, thr; float eta; uint32_t pixel, sidx; int depth, n_delta; bool last_delta; };
B2_HD PtState pt_load(const RenderCtx &c, int p) {
    const float4 o = c.ray_o[p], d = c.ray_d[p], t = c.thr[p], x = c.last_pos[p];
    PtState s;
    s.ro = mk3(o); s.eta = o.w; s.rd = mk3(d); s.thr = mk3(t);
    const uint32_t w = f2u(t.w);
    s.depth = (int) (w & 0xffu); s.n_delta = (int) ((w >> 8) & 0x7fffffu); s.last_delta = (w >> 31) != 0u;
    s.pixel = f2u(x.x); s.sidx = f2u(x.y);
    return s;
}
B2_HD void pt_store(const RenderCtx &c, int p, const PtState &s) {
    c.ray_o[p] = mk4(s.ro, s.eta); c.ray_d[p] = mk4(s.rd, 0.0f);
    c.thr[p] = mk4(s.thr, u2f((uint32_t) s.depth | ((uint32_t) s.n_delta << 8) | (s.last_delta ? 0x80000000u : 0u)));
    c.last_pos[p] = make_float4(u2f(s.pixel), u2f(s.sidx), 0.0f, 0.0f);
}
B2_HD PtState pt_primary(const RenderCtx &c, uint32_t pix0, int s0, uint32_t q) {
    const PathState b = primary_path(c, pix0, s0, q);                     // pt_cu.cu:37-44: the same jittered pinhole ray
    PtState s;
    s.ro = b.ro; s.rd = b.rd; s.thr = b.thr; s.eta = 1.0f; s.pixel = b.pixel; s.sidx = b.sidx; s.depth = 0; s.n_delta = 0; s.last_delta = true;
    return s;
}

// Radiance of a light ball seen at `pos` (pt_cu.cu:60-96): the first light whose ball surface is within 1e-2 of the hit;
// illum / (4 pi r^2 * cone fraction), the cone fraction forced to 1 at depth 0 and 0 behind the shade.
B2_HD float3 pt_emission(const DevScene &sc, float3 pos, float3 illum, int depth) {
    for (int i = 0; i < sc.n_lights; ++i) {
        const float4 l0 = sc.lights[4 * i], l1 = sc.lights[4 * i + 1], l2 = sc.lights[4 * i + 2], l3 = sc.lights[4 * i + 3];
        const float3 cth = pos - mk3(l0);
        if (!(fabsf(length3(cth) - l0.w) < 1e-2f)) continue;
        const float area = 4.0f * B2_PI * l0.w * l0.w;
        float cone = 1.0f;
        if (l1.w > 0.0f && l2.w == 0.0f) {
            cone = (1.0f - l3.w) * 0.5f;
            if (depth == 0) cone = 1.0f;
            else if (dot3(mk3(l3), normalize3(cth)) < l3.w) cone = 0.0f;
        }
        if (!(cone > 0.0f)) return mk3(0, 0, 0);
        return illum / (area * cone);
    }
    return mk3(0, 0, 0);
}

// One bounce of a path (pt_cu.cu:48-249 for one value of `depth`) in two parts, so that the wavefront can run each as its
// own kernel (b2_integrators.cu): pt_shade = light hit / emission, the BSDF sample and the next ray; pt_nee = the
// next-event estimation of the same vertex from an eye-vertex record (EyeVtx with .path = sample index).
//   pt_shade returns true when the path continues (ps = next ray); `add` = the emitted radiance this bounce adds to the
//   pixel; want_nee: `ev` is a vertex that takes direct light (:126).
B2_HD bool pt_shade(const RenderCtx &c, PtState &ps, HitRec h, float3 &add, EyeVtx &ev, bool &want_nee, uint32_t &n_path, uint32_t &n_cap) {
    const DevScene &sc = c.sc;
    add = mk3(0, 0, 0);
    want_nee = false;
    if (h.prim == B2_PRIM_MISS) return false;
    const Surface s = surface_at(sc, ps.ro, ps.rd, h);
    const float3 wo = ps.rd * -1.0f;
    if (s.is_light) {                                                     // :58-122; the !last_is_delta branch never adds (:105-119)
        if (ps.last_delta) {
            const float3 contrib = ps.thr * pt_emission(sc, s.pos, s.mtl.base, ps.depth);
            if (valid_color(contrib) && (contrib.x > 0.0f || contrib.y > 0.0f || contrib.z > 0.0f)) { add = clamp_radiance(contrib, 15.0f); n_path++; }
        }
        return false;
    }
    if (gathers_light(s.mtl) && sc.n_lights > 0) {
        want_nee = true;
        ev.pos = s.pos; ev.n = s.n; ev.wo = wo; ev.thr = ps.thr; ev.mtl = s.mtl; ev.depth = ps.depth; ev.pixel = ps.pixel; ev.path = (int) ps.sidx;
    }
    float u_rr, ua, ub;
    rng_draw3(c.key, B2_RNG_EYE, ps.pixel, ps.sidx, (uint32_t) (ps.depth + ps.n_delta), u_rr, ua, ub);
    BsdfSample bs; bs.wi = mk3(0, 0, 0); bs.f = mk3(0, 0, 0); bs.pdf = 0.0f;
    bsdf_sample(s.mtl, wo, s.n, u_rr, ua, ub, ps.eta, bs);
    if (bs.pdf <= 0.0f) return false;                                     // :222 (and the pdf-0 delta case, see the header)
    if (bs.is_delta) {                                                    // :224-239
        ps.thr = ps.thr * bs.f;
        ps.rd = bs.wi; ps.eta = bs.new_eta;
        ps.ro = s.pos + s.n * (dot3(bs.wi, s.n) < 0.0f ? -B2_EPSILON : B2_EPSILON);
        ps.last_delta = true;
        if (!valid_color(ps.thr)) return false;
        if (++ps.n_delta >= c.max_delta) { n_cap++; return false; }
        return true;
    }
    ps.thr = ps.thr * bs.f * fabsf(dot3(s.n, bs.wi)) / bs.pdf;            // :241-249
    if (!valid_color(ps.thr)) return false;
    ps.rd = bs.wi; ps.ro = s.pos + s.n * B2_EPSILON;
    ps.last_delta = false;
    return ++ps.depth < c.eye_depth;
}
// Next-event estimation of one vertex (:125-208): one light chosen uniformly; a parallel light is a direction, any other a
// point on its ball with the power heuristic against the BSDF density.  The draw is keyed by the vertex's depth: a
// vertex that takes direct light is never a delta bounce, so (pixel, sample, depth) names it uniquely.
// Three steps, so that the wavefront kernel can trace the shadow rays of many vertices together (k_pt_nee):
//   pt_nee_prepare : light sample, cosine / spot-cone tests -> shadow ray + what pt_nee_finish needs (false: no contribution)
//   (visibility)   : any-hit, every object opaque
//   pt_nee_finish  : BSDF value and density, power heuristic, validity + clamp
struct NeeSample { float3 wi; float dist2, cos_s, cos_l; int light; };          // light < 0: parallel light -(light + 1)
B2_HD bool pt_nee_prepare(const RenderCtx &c, const EyeVtx &ev, ShadowRay &sr, NeeSample &ns) {
    const DevScene &sc = c.sc;
    float ul, u1, u2, unused;
    rng_draw4(c.key, B2_RNG_PT_NEE, ev.pixel, (uint32_t) ev.path, 1u + (uint32_t) ev.depth, ul, u1, u2, unused);
    int li = (int) (ul * (float) sc.n_lights);
    if (li > sc.n_lights - 1) li = sc.n_lights - 1;
    const float4 l0 = sc.lights[4 * li], l1 = sc.lights[4 * li + 1], l2 = sc.lights[4 * li + 2], l3 = sc.lights[4 * li + 3];
    const float3 p1 = x3_add(ev.pos, x3_muls(ev.n, B2_EPSILON));
    if (l2.w != 0.0f) {                                                   // parallel light, :131-151
        ns.wi = mk3(l3) * -1.0f; ns.light = -(li + 1); ns.dist2 = 0.0f; ns.cos_l = 1.0f;
        ns.cos_s = fmaxf(0.0f, dot3(ev.n, ns.wi));
        if (!(ns.cos_s > 0.0f)) return false;
        sr = make_shadow_ray(p1, ev.pos + ns.wi * 1e4f);
        return true;
    }
    const float z = 1.0f - 2.0f * u1, r = fast_sqrt(fmaxf(0.0f, 1.0f - z * z));      // point on the light ball, :152-206
    float sphi, cphi;
    sincos_2pi(u2, sphi, cphi);
    const float3 dl = mk3(r * cphi, r * sphi, z);
    const float3 lpos = mk3(l0) + dl * l0.w;
    float3 wi = lpos - ev.pos;
    ns.dist2 = dot3(wi, wi);
    wi = div_sqrt3(wi, ns.dist2);
    ns.wi = wi; ns.light = li;
    ns.cos_s = fmaxf(0.0f, dot3(ev.n, wi)); ns.cos_l = fmaxf(0.0f, -dot3(dl, wi));
    if (!(ns.cos_s > 0.0f && ns.cos_l > 0.0f) || (l1.w > 0.0f && -dot3(mk3(l3), wi) < l3.w)) return false;
    sr = make_shadow_ray(p1, x3_add(lpos, x3_muls(dl, B2_EPSILON)));
    return true;
}
B2_HD float3 pt_nee_finish(const RenderCtx &c, const EyeVtx &ev, const NeeSample &ns, uint32_t &n_path) {
    const DevScene &sc = c.sc;
    float3 contrib;
    if (ns.light < 0) {
        const float4 l2 = sc.lights[4 * (-ns.light - 1) + 2];
        contrib = ev.thr * bsdf_evaluate(ev.mtl, ev.wo, ns.wi, ev.n) * mk3(l2) * (ns.cos_s * (float) sc.n_lights);
    } else {
        const float4 l0 = sc.lights[4 * ns.light], l2 = sc.lights[4 * ns.light + 2];
        const float area = 4.0f * B2_PI * l0.w * l0.w;
        const float pdf_l = fast_div(ns.dist2, (float) sc.n_lights * area * fmaxf(ns.cos_l, 1e-6f));
        const float pdf_b = bsdf_pdf(ev.mtl, ev.wo, ns.wi, ev.n);
        const float pl2 = pdf_l * pdf_l, pb2 = pdf_b * pdf_b;
        const float w = fast_div(pl2, fmaxf(pl2 + pb2, 1e-8f));          // power heuristic, :192-194
        contrib = ev.thr * bsdf_evaluate(ev.mtl, ev.wo, ns.wi, ev.n) * mk3(l2) * (fast_div(ns.cos_s, pdf_l) * w);
    }
    if (!valid_color(contrib)) return mk3(0, 0, 0);
    n_path++;
    return clamp_radiance(contrib, 15.0f);
}
template <bool COUNT>
B2_HD float3 pt_nee(const RenderCtx &c, const EyeVtx &ev, uint2 *stack, int sstride, TravCounters *tc, uint32_t &n_shadow, uint32_t &n_path) {
    ShadowRay sr; NeeSample ns;
    if (!pt_nee_prepare(c, ev, sr, ns)) return mk3(0, 0, 0);
    n_shadow++;
    if (occluded_bvh<COUNT, true>(c.sc, sr.o, sr.d, sr.tmax, stack, sstride, tc)) return mk3(0, 0, 0);
    return pt_nee_finish(c, ev, ns, n_path);
}
// Both parts for one path (the sequential host driver of tests/hostsim).
template <bool COUNT>
B2_HD bool pt_bounce(const RenderCtx &c, PtState &ps, HitRec h, uint2 *stack, int sstride, TravCounters *tc,
                     float3 &add, uint32_t &n_shadow, uint32_t &n_path, uint32_t &n_cap) {
    EyeVtx ev; bool want_nee;
    const bool alive = pt_shade(c, ps, h, add, ev, want_nee, n_path, n_cap);
    if (want_nee) add = add + pt_nee<COUNT>(c, ev, stack, sstride, tc, n_shadow, n_path);
    return alive;
}

// ---- photon mapping ----
// Hit point of a pixel (CudaHitPoint, geometric.cuh:53-65, without its dead fields): 5 x float4
//   h0 = {pos, cell key low bits}  h1 = {normal, pixel}  h2 = {wo, roughness}  h3 = {throughput, metallic}  h4 = {base, eta}
// kept in CELL order (sorted by the 63-bit key of their grid cell) so that a photon reads the hit points of a cell as one
// contiguous run; flux = {accum_flux, photon_count} per hit point, same order.
#define B2_HP_F4 5
struct PpmCtx {
    float4 *hp; float4 *flux; float *direct;      // n_hp records; W*H*3 direct term (light seen through a delta chain)
    unsigned long long *keys; uint32_t *order;    // cell key per hit point / sort permutation
    ulonglong2 *table; uint32_t table_mask;       // open-addressing cell table: {key + 1, start | count << 32}; 0 = empty
    int *n_hp;
    float radius, inv_cell;
};
B2_HD unsigned long long ppm_cell_key(int gx, int gy, int gz) {           // 21 bits per axis, biased: cells of get_grid_index (ppm_cu.cu:32-38)
    return ((unsigned long long) ((uint32_t) (gx + (1 << 20)) & 0x1fffffu) << 42) | ((unsigned long long) ((uint32_t) (gy + (1 << 20)) & 0x1fffffu) << 21) |
           (unsigned long long) ((uint32_t) (gz + (1 << 20)) & 0x1fffffu);
}
B2_HD void ppm_cell_of(const DevScene &sc, float inv_cell, float3 p, int &gx, int &gy, int &gz) {
    gx = (int) floorf((p.x - sc.scene_min.x) * inv_cell); gy = (int) floorf((p.y - sc.scene_min.y) * inv_cell); gz = (int) floorf((p.z - sc.scene_min.z) * inv_cell);
}
B2_HD uint32_t ppm_hash(unsigned long long k) { k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33; return (uint32_t) k; }

// Eye pass of one pixel (ppm_cu.cu:64-152): follow delta bounces until the first other surface.  Returns true when a hit
// point was created; `direct` = the light seen through the delta chain (:112-118), zero otherwise.
struct HitPoint { float3 pos, n, wo, thr; Mat mtl; uint32_t pixel; };
template <bool COUNT>
B2_HD bool ppm_eye_body(const RenderCtx &c, uint32_t pixel, uint32_t sidx, uint2 *stack, int sstride, TravCounters *tc,
                        HitPoint &hp, float3 &direct, uint32_t &n_closest, uint32_t &n_cap) {
    direct = mk3(0, 0, 0);
    if (c.eye_depth <= 0) return false;
    PathState b;
    {   // the jittered primary ray of pixel `pixel`, sample `sidx`
        float jx, jy;
        rng_draw2(c.key, B2_RNG_EYE, pixel, sidx, jx, jy);
        uint32_t col;
        const uint32_t row = udiv_magic(pixel, (uint32_t) c.W, c.w_magic, col);
        const float3 pp = x3_add(x3_add(c.cUL, x3_muls(c.cdx, x_add((float) col, jx))), x3_muls(c.cdy, x_add((float) row, jy)));
        b.rd = f3_normalize(x3_sub(pp, c.cam_eye)); b.ro = c.cam_eye;
    }
    float3 thr = mk3(1.0f, 1.0f, 1.0f);
    float eta = 1.0f;
    for (int n_delta = 0;; ++n_delta) {
        const HitRec h = closest_hit_bvh<COUNT>(c.sc, b.ro, b.rd, stack, sstride, tc);
        n_closest++;
        if (h.prim == B2_PRIM_MISS) return false;
        const Surface s = surface_at(c.sc, b.ro, b.rd, h);
        const float3 wo = b.rd * -1.0f;
        if (s.is_light) {
            const float3 contrib = thr * s.mtl.base;
            if (valid_color(contrib)) direct = clamp_radiance(contrib, 15.0f);
            return false;
        }
        float u_rr, ua, ub;
        rng_draw3(c.key, B2_RNG_EYE, pixel, sidx, (uint32_t) n_delta, u_rr, ua, ub);
        BsdfSample bs; bs.wi = mk3(0, 0, 0); bs.f = mk3(0, 0, 0); bs.pdf = 0.0f;
        bsdf_sample(s.mtl, wo, s.n, u_rr, ua, ub, eta, bs);
        if (!bs.is_delta) {                                               // :141-150
            hp.pos = s.pos; hp.n = s.n; hp.wo = wo; hp.thr = thr; hp.mtl = s.mtl; hp.pixel = pixel;
            return true;
        }
        if (bs.pdf <= 0.0f) return false;                                 // :130-131
        thr = thr * bs.f;
        b.rd = bs.wi; eta = bs.new_eta;
        b.ro = s.pos + s.n * (dot3(bs.wi, s.n) < 0.0f ? -B2_EPSILON : B2_EPSILON);
        if (!valid_color(thr)) return false;
        if (n_delta + 1 >= c.max_delta) { n_cap++; return false; }
    }
}

// One photon (ppm_cu.cu:158-287): emission as the BDPT light subpaths, flux = illum * n_lights / spl, at every surface
// that gathers light the flux goes to the hit points within `radius` whose normal agrees (:221-260), then the photon
// scatters.  flux_add(hit point index, rgb) is the accumulation (an atomic on the device).
template <bool COUNT, class FluxAdd>
B2_HD void ppm_photon_body(const RenderCtx &c, const PpmCtx &pc, int idx, uint2 *stack, int sstride, TravCounters *tc, FluxAdd &flux_add,
                           uint32_t &n_closest, uint32_t &n_gather, uint32_t &n_cap) {
    const DevScene &sc = c.sc;
    const int li = idx % sc.n_lights;
    const float4 l0 = sc.lights[4 * li], l2 = sc.lights[4 * li + 2], l3 = sc.lights[4 * li + 3];
    float r1, r2;
    rng_draw2(c.key, B2_RNG_PHOTON, (uint32_t) idx, 0u, r1, r2);
    const float3 w = mk3(l3);
    float3 u_vec = (fabsf(w.x) > 0.9f) ? mk3(0, 1, 0) : mk3(1, 0, 0);
    const float3 v_vec = normalize3(cross3(w, u_vec));
    u_vec = normalize3(cross3(v_vec, w));
    float3 ro, rd;
    if (l2.w != 0.0f) {                                                   // :171-188
        rd = w;
        const float3 center = (sc.scene_min + sc.scene_max) * 0.5f;
        const float radius = length3(sc.scene_max - sc.scene_min) * 0.5f;
        ro = center - rd * (radius * 2.0f) + u_vec * ((r1 - 0.5f) * radius * 2.0f) + v_vec * ((r2 - 0.5f) * radius * 2.0f);
    } else {                                                              // :189-204
        const float theta = acosf(1.0f - r1 * (1.0f - l3.w));
        const float phi = 2.0f * B2_PI * r2;
        const float st = sinf(theta);
        rd = normalize3(u_vec * (st * cosf(phi)) + v_vec * (st * sinf(phi)) + w * cosf(theta));
        ro = mk3(l0) + rd * l0.w;
    }
    float3 flux = mk3(l2) * ((float) sc.n_lights / fmaxf((float) c.spl, 1.0f));   // :209
    float eta = 1.0f;
    int n_delta = 0;
    const float r2max = pc.radius * pc.radius;
    for (int depth = 0; depth < c.light_depth; ++depth) {
        const HitRec h = closest_hit_bvh<COUNT>(sc, ro, rd, stack, sstride, tc);
        n_closest++;
        if (h.prim == B2_PRIM_MISS) break;
        const Surface s = surface_at(sc, ro, rd, h);
        if (s.is_light) break;
        const float3 wi_l = rd * -1.0f;
        if (gathers_light(s.mtl)) {
            int gx, gy, gz;
            ppm_cell_of(sc, pc.inv_cell, s.pos, gx, gy, gz);
            for (int dz = -1; dz <= 1; ++dz) for (int dy = -1; dy <= 1; ++dy) for (int dx = -1; dx <= 1; ++dx) {
                const unsigned long long key = ppm_cell_key(gx + dx, gy + dy, gz + dz) + 1ull;
                uint32_t slot = ppm_hash(key) & pc.table_mask;
                ulonglong2 e = pc.table[slot];
                while (e.x != 0ull && e.x != key) { slot = (slot + 1u) & pc.table_mask; e = pc.table[slot]; }
                if (e.x == 0ull) continue;
                const uint32_t start = (uint32_t) e.y, count = (uint32_t) (e.y >> 32);
                for (uint32_t k = start; k < start + count; ++k) {
                    const float4 h0 = pc.hp[B2_HP_F4 * (size_t) k], h1 = pc.hp[B2_HP_F4 * (size_t) k + 1];
                    if (!(dot3(mk3(h1), s.n) > 0.01f)) continue;
                    const float3 dd = mk3(h0) - s.pos;
                    if (!(dot3(dd, dd) < r2max)) continue;
                    const float4 h2 = pc.hp[B2_HP_F4 * (size_t) k + 2], h3 = pc.hp[B2_HP_F4 * (size_t) k + 3], h4 = pc.hp[B2_HP_F4 * (size_t) k + 4];
                    Mat m; m.base = mk3(h4); m.roughness = h2.w; m.metallic = h3.w; m.eta = h4.w;
                    const float3 f = bsdf_evaluate(m, mk3(h2), wi_l, mk3(h1));
                    if (valid_color(f)) { flux_add(k, flux * f * mk3(h3)); n_gather++; }
                }
            }
        }
        float u_rr, ua, ub;
        rng_draw3(c.key, B2_RNG_PHOTON, (uint32_t) idx, 0u, (uint32_t) (depth + n_delta), u_rr, ua, ub);
        BsdfSample bs; bs.wi = mk3(0, 0, 0); bs.f = mk3(0, 0, 0); bs.pdf = 0.0f;
        bsdf_sample(s.mtl, wi_l, s.n, u_rr, ua, ub, eta, bs);
        if (bs.pdf <= 0.0f) break;                                        // :272
        if (bs.is_delta) {
            flux = flux * bs.f;
            if (++n_delta >= c.max_delta) { n_cap++; break; }
            depth--;
        } else flux = flux * bs.f * fabsf(dot3(s.n, bs.wi)) / bs.pdf;
        if (!valid_color(flux)) break;
        rd = bs.wi; eta = bs.new_eta;
        ro = s.pos + s.n * (dot3(bs.wi, s.n) < 0.0f ? -B2_EPSILON : B2_EPSILON);
    }
}

// ppm_resolve_image, ppm_cu.cu:293-315: direct term + clamp(flux / max(pi r^2, 1e-6), 15)
B2_HD float3 ppm_resolve(float3 direct, float3 flux, bool valid, float radius) {
    if (!valid) return direct;
    const float3 rad = flux / fmaxf(B2_PI * radius * radius, 1e-6f);
    return valid_color(rad) ? direct + clamp_radiance(rad, 15.0f) : direct;
}
