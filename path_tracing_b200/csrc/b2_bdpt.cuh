// b2_bdpt.cuh -- per-item bodies of the wavefront BDPT stages.  The estimator is the reference's CPU one
// (src/cpu_bdpt.cpp:173-488), quirk for quirk: one shared light-vertex pool per render,
// connections from every eye vertex (delta ones included) to every pool vertex, placeholder pdf_fwd = 0 at
// connection time, per-connection clamp at 15, no Russian roulette, delta bounces that do not advance depth.
// The __global__ wrappers and the host wave loop are in b2_render.cu; tests/hostsim runs the same bodies
// sequentially on the host.
//
// Stages of one wave (all paths of a batch advance one bounce):
//   closest  : ray -> (prim, t)                              [closest_hit_bvh]
//   shade    : hit -> eye vertex record (CV) + BSDF sample + next ray / termination
//   connect  : CV x pool vertex -> weighted contribution + shadow segment, appended to the shadow queue
//   shadow   : any-hit test of each queued segment; visible ones are splatted into the film (atomicAdd)
//   compact  : surviving path ids -> next wave's queue
#pragma once
#include "b2_bvh.cuh"
#include "b2_rng.cuh"

// ---- light-vertex pool record: 5 x float4 (80 B), CudaLightVertex of bdpt_cu.cuh:6-17 without the dead fields ----
//   f0 = {pos, flags}   f1 = {normal, source_cutoff}   f2 = {throughput, pdf_fwd}   f3 = {base_color, pdf_rev}
//   f4 = {roughness, metallic, eta, 0}
#define B2_LV_F4 5
#define B2_LV_SRC 1u
#define B2_LV_PARALLEL 2u
struct LightVtx { float3 pos, n, thr; Mat mtl; uint32_t flags; float cutoff, pdf_fwd, pdf_rev; };
B2_HD LightVtx load_lv(const float4 *pool, int i) {
    const float4 a = pool[B2_LV_F4 * i], b = pool[B2_LV_F4 * i + 1], c = pool[B2_LV_F4 * i + 2], d = pool[B2_LV_F4 * i + 3], e = pool[B2_LV_F4 * i + 4];
    LightVtx v;
    v.pos = mk3(a); v.flags = f2u(a.w); v.n = mk3(b); v.cutoff = b.w; v.thr = mk3(c); v.pdf_fwd = c.w;
    v.mtl.base = mk3(d); v.pdf_rev = d.w; v.mtl.roughness = e.x; v.mtl.metallic = e.y; v.mtl.eta = e.z;
    return v;
}
B2_HD void store_lv(float4 *pool, int i, const LightVtx &v) {
    pool[B2_LV_F4 * i] = mk4(v.pos, u2f(v.flags));
    pool[B2_LV_F4 * i + 1] = mk4(v.n, v.cutoff);
    pool[B2_LV_F4 * i + 2] = mk4(v.thr, v.pdf_fwd);
    pool[B2_LV_F4 * i + 3] = mk4(v.mtl.base, v.pdf_rev);
    pool[B2_LV_F4 * i + 4] = make_float4(v.mtl.roughness, v.mtl.metallic, v.mtl.eta, 0.0f);
}

// ---- eye-vertex record handed from shade to connect: 5 x float4 (80 B) ----
//   f0 = {pos, depth}  f1 = {normal, path}  f2 = {wo, roughness}  f3 = {throughput, metallic}  f4 = {base_color, eta}
#define B2_CV_F4 5
struct EyeVtx { float3 pos, n, wo, thr; Mat mtl; int depth, path; };
B2_HD EyeVtx load_cv(const float4 *cv, int q) {
    const float4 a = cv[B2_CV_F4 * q], b = cv[B2_CV_F4 * q + 1], c = cv[B2_CV_F4 * q + 2], d = cv[B2_CV_F4 * q + 3], e = cv[B2_CV_F4 * q + 4];
    EyeVtx v;
    v.pos = mk3(a); v.depth = (int) f2u(a.w); v.n = mk3(b); v.path = (int) f2u(b.w); v.wo = mk3(c); v.mtl.roughness = c.w;
    v.thr = mk3(d); v.mtl.metallic = d.w; v.mtl.base = mk3(e); v.mtl.eta = e.w;
    return v;
}
B2_HD void store_cv(float4 *cv, int q, const EyeVtx &v) {
    cv[B2_CV_F4 * q] = mk4(v.pos, u2f((uint32_t) v.depth));
    cv[B2_CV_F4 * q + 1] = mk4(v.n, u2f((uint32_t) v.path));
    cv[B2_CV_F4 * q + 2] = mk4(v.wo, v.mtl.roughness);
    cv[B2_CV_F4 * q + 3] = mk4(v.thr, v.mtl.metallic);
    cv[B2_CV_F4 * q + 4] = mk4(v.mtl.base, v.mtl.eta);
}
#define B2_CV_NONE 0xffffffffu     // depth field of a queue slot that produced no eye vertex this wave

// ---- shadow-queue entry: 3 x float4 (48 B) = {origin, tmax} {dir, pixel} {contribution, 0} ----
#define B2_SQ_F4 3

enum { B2_CNT_EYE_SAMPLES = 0, B2_CNT_CLOSEST, B2_CNT_SHADOW, B2_CNT_PATH_SAMPLES, B2_CNT_DELTA_CAP, B2_CNT_NODES, B2_CNT_TRIS, B2_CNT_SQ_OVERFLOW, B2_CNT_N };

struct RenderCtx {
    DevScene sc;
    float3 cam_eye, cUL, cdx, cdy;          // cpu_bdpt.cpp:190-200
    int W, H, spl, eye_depth, light_depth, max_delta;
    int sample_offset, sample_stride;
    uint2 key;
    // light-vertex pool
    float4 *pool; int n_pool;
    int *pool_active; int *n_pool_active;
    // per-path state, capacity max_paths
    int max_paths;
    float4 *ray_o;      // {origin, current eta}
    float4 *ray_d;      // {direction, last_pdf_omega}
    float4 *thr;        // {throughput, -}
    float4 *last_pos;   // {last_pos, -}
    float4 *last_n;     // {last_normal, -}
    uint32_t *pix, *sidx;
    uint32_t *pstate;   // depth | n_delta << 8
    int2 *hit;          // {prim, float bits of t}, indexed by queue slot
    float4 *hist_pe;    // [eye_depth][max_paths] {pos, eta} of stored (non-delta) eye vertices
    float2 *hist_pdf;   // [eye_depth][max_paths] {pdf_fwd, pdf_rev}
    int *live;          // queue of live path ids for this wave
    int *live_next; int *n_live_next;
    float4 *cv;         // eye-vertex records, indexed by queue slot
    float4 *sq; uint32_t *n_sq; uint32_t sq_cap;
    float *film;        // W*H*3 running sums
    unsigned long long *counters;
};

// camera basis exactly as run_cpu_bdpt derives it (glm arithmetic, cpu_bdpt.cpp:190-199); host-side helper
inline void camera_basis(const float eye[3], const float look_at[3], const float view_up[3], float fov, int W, int H,
                         float3 &cam_eye, float3 &cUL, float3 &cdx, float3 &cdy) {
    float3 e = make_float3(eye[0], eye[1], eye[2]), la = make_float3(look_at[0], look_at[1], look_at[2]), vu = make_float3(view_up[0], view_up[1], view_up[2]);
    float aspect = float(W) / float(H);
    float theta = fov * B2_PI / 180.0f;
    float half_height = tanf(theta / 2.0f);
    float half_width = aspect * half_height;
    float3 cw = g3_normalize(x3_sub(e, la));
    float3 cu = g3_normalize(x3_cross(vu, cw));
    float3 cv = x3_cross(cw, cu);
    cUL = x3_sub(x3_add(x3_sub(e, x3_muls(cu, half_width)), x3_muls(cv, half_height)), cw);
    cdx = x3_divs(x3_muls(cu, 2.0f * half_width), float(W));
    cdy = x3_divs(x3_muls(cv, -2.0f * half_height), float(H));
    cam_eye = e;
}

// Surface point of a hit: position, geometric normal flipped against the ray, material (cpu_bdpt.cpp:49-58, 67-76)
struct Surface { float3 pos, n; Mat mtl; bool is_light; };
B2_HD Surface surface_at(const DevScene &sc, float3 ro, float3 rd, HitRec h) {
    Surface s;
    s.pos = x3_add(ro, x3_muls(rd, h.t));
    const int kind = prim_kind(h.prim), i = prim_index(h.prim);
    if (kind == B2_KIND_TRI) {
        const float4 b = sc.tri[3 * i + 1], c = sc.tri[3 * i + 2];
        float3 n = g3_normalize(x3_cross(mk3(b), mk3(c)));                 // Triangle::geom_normal, object.cpp:68-70
        if (x3_dot(n, rd) > 0.0f) n = x3_neg(n);
        s.n = n; s.mtl = load_mat(sc, (int) (f2u(b.w) & ~B2_MAT_TRANSPARENT)); s.is_light = false;
    } else if (kind == B2_KIND_SPHERE) {
        const float4 a = sc.sph[2 * i], b = sc.sph[2 * i + 1];
        float3 n = g3_normalize(x3_sub(s.pos, mk3(a)));                    // Sphere::normal_at, object.cpp:58-66 (scale 1)
        if (x3_dot(n, rd) > 0.0f) n = x3_neg(n);
        s.n = n; s.mtl = load_mat(sc, (int) (f2u(b.y) & ~B2_MAT_TRANSPARENT)); s.is_light = false;
    } else {
        const float4 a = sc.lights[4 * i], il = sc.lights[4 * i + 2];
        float3 n = f3_normalize(x3_sub(s.pos, mk3(a)));                    // cpu_bdpt.cpp:74-76
        if (x3_dot(n, rd) > 0.0f) n = x3_muls(n, -1.0f);
        s.n = n; s.mtl.base = mk3(il); s.mtl.eta = 0.0f; s.mtl.roughness = 1.0f; s.mtl.metallic = 0.0f; s.is_light = true;
    }
    return s;
}

// =====================================================================================================
// Light subpaths -> pool (cpu_bdpt.cpp:218-325).  One item = one light subpath idx in [0, n_lights*spl).
// The pool is zero-filled beforehand (the reference value-initialises it, :203).
// =====================================================================================================
template <bool COUNT>
B2_HD void light_path_body(const RenderCtx &c, int idx, uint2 *stack, int sstride, TravCounters *tc, uint32_t &n_closest, uint32_t &n_cap) {
    const DevScene &sc = c.sc;
    const int light_idx = idx % sc.n_lights;
    const float4 l0 = sc.lights[4 * light_idx], l1 = sc.lights[4 * light_idx + 1], l2 = sc.lights[4 * light_idx + 2], l3 = sc.lights[4 * light_idx + 3];
    const bool is_parallel = l2.w != 0.0f;
    const int base = idx * c.light_depth;
    float3 ro, rd;
    float eta = 1.0f;
    float r1, r2;
    rng_draw2(c.key, B2_RNG_LIGHT, (uint32_t) idx, 0u, r1, r2);
    const float3 w = mk3(l3);                                              // normalize(light.dir)
    float3 u_vec = (fabsf(w.x) > 0.9f) ? mk3(0, 1, 0) : mk3(1, 0, 0);
    const float3 v_vec = normalize3(cross3(w, u_vec));
    u_vec = normalize3(cross3(v_vec, w));
    if (is_parallel) {                                                     // :227-240
        rd = w;
        const float3 center = (sc.scene_min + sc.scene_max) * 0.5f;
        const float radius = length3(sc.scene_max - sc.scene_min) * 0.5f;
        const float ou = (r1 - 0.5f) * radius * 2.0f, ov = (r2 - 0.5f) * radius * 2.0f;
        ro = center - rd * (radius * 2.0f) + u_vec * ou + v_vec * ov;
    } else {                                                               // :241-254
        const float theta = acosf(1.0f - r1 * (1.0f - l3.w));
        const float phi = 2.0f * B2_PI * r2;
        const float st = sinf(theta);
        const float3 ld = mk3(st * cosf(phi), st * sinf(phi), cosf(theta));
        rd = normalize3(u_vec * ld.x + v_vec * ld.y + w * ld.z);
        ro = mk3(l0) + rd * l0.w;
    }
    float3 thr = mk3(l2) / fmaxf((float) c.spl, 1.0f);                     // :256
    {
        LightVtx v0; v0.pos = ro; v0.n = rd; v0.thr = thr; v0.flags = B2_LV_SRC | (is_parallel ? B2_LV_PARALLEL : 0u);
        v0.cutoff = l1.w; v0.pdf_fwd = 0.0f; v0.pdf_rev = 0.0f; v0.mtl.base = mk3(0, 0, 0); v0.mtl.roughness = v0.mtl.metallic = v0.mtl.eta = 0.0f;
        store_lv(c.pool, base, v0);
    }
    float3 last_n = rd, last_pos = ro;
    float last_pdf = 1.0f / B2_PI;
    int n_delta = 0;
    BsdfSample bs; bs.wi = mk3(0, 0, 0); bs.f = mk3(0, 0, 0); bs.pdf = 0.0f;
    for (int depth = 1; depth < c.light_depth; ++depth) {                  // :271-324
        const HitRec h = closest_hit_bvh<COUNT>(sc, ro, rd, stack, sstride, tc);
        n_closest++;
        if (h.prim == B2_PRIM_MISS) break;
        const Surface s = surface_at(sc, ro, rd, h);
        LightVtx v;
        if (s.is_light) {                                                  // pseudo-source vertex, :277-282
            v.pos = s.pos; v.n = s.n; v.thr = thr; v.mtl = s.mtl; v.flags = B2_LV_SRC; v.cutoff = 0.0f; v.pdf_fwd = 0.0f; v.pdf_rev = 0.0f;
            store_lv(c.pool, base + depth, v);
            break;
        }
        if (length3(thr) < 1e-4f) break;
        const float3 dp = s.pos - last_pos;
        const float dist2 = dot3(dp, dp);
        if (dist2 < 1e-6f) break;
        const float3 wo = rd * -1.0f;
        const float cos_hit = fabsf(dot3(s.n, wo));
        const float cos_prev = fabsf(dot3(last_n, rd));
        const float pdf_fwd = last_pdf * cos_hit / dist2;
        float u_rr, u1, u2;
        rng_draw3(c.key, B2_RNG_LIGHT, (uint32_t) idx, 0u, (uint32_t) (depth - 1 + n_delta), u_rr, u1, u2);
        bsdf_sample(s.mtl, wo, s.n, u_rr, u1, u2, eta, bs);
        if (bs.pdf <= 0.0f && !bs.is_delta) break;
        if (bs.is_delta) {                                                 // :302-308 (last_* NOT updated)
            thr = thr * bs.f;
            rd = bs.wi; eta = bs.new_eta;
            ro = s.pos + s.n * (dot3(bs.wi, s.n) < 0.0f ? -B2_EPSILON : B2_EPSILON);
            if (++n_delta >= c.max_delta) { n_cap++; break; }
            depth--;
            continue;
        }
        v.pos = s.pos; v.n = s.n; v.thr = thr; v.mtl = s.mtl; v.flags = 0u; v.cutoff = 0.0f;
        v.pdf_fwd = pdf_fwd;
        v.pdf_rev = bsdf_pdf(s.mtl, bs.wi, wo, s.n) * cos_prev / dist2;
        store_lv(c.pool, base + depth, v);
        thr = thr * bs.f * fabsf(dot3(s.n, bs.wi)) / bs.pdf;
        if (!valid_color(thr)) break;
        rd = bs.wi;
        ro = s.pos + s.n * B2_EPSILON;
        last_pdf = bs.pdf; last_n = s.n; last_pos = s.pos;
    }
}
// A pool vertex takes part in connections iff length(throughput) >= 1e-6 (cpu_bdpt.cpp:391)
B2_HD bool pool_vertex_active(const float4 *pool, int i) { return !(length3(mk3(pool[B2_LV_F4 * i + 2])) < 1e-6f); }

// =====================================================================================================
// Eye subpaths
// =====================================================================================================
// Path i of a batch covers work item w = w0 + i: pixel = w % npix, local sample = w / npix (cpu_bdpt.cpp:344-361)
B2_HD void eye_gen_body(const RenderCtx &c, int i, long long w0) {
    const long long w = w0 + i;
    const int npix = c.W * c.H;
    const uint32_t pixel = (uint32_t) (w % npix);
    const uint32_t s = (uint32_t) (c.sample_offset + (int) (w / npix) * c.sample_stride);
    float jx, jy;
    rng_draw2(c.key, B2_RNG_EYE, pixel, s, jx, jy);
    const float px = x_add((float) (pixel % (uint32_t) c.W), jx), py = x_add((float) (pixel / (uint32_t) c.W), jy);
    const float3 pp = x3_add(x3_add(c.cUL, x3_muls(c.cdx, px)), x3_muls(c.cdy, py));
    const float3 d = f3_normalize(x3_sub(pp, c.cam_eye));
    c.ray_o[i] = mk4(c.cam_eye, 1.0f);
    c.ray_d[i] = mk4(d, 1.0f);
    c.thr[i] = make_float4(1.0f, 1.0f, 1.0f, 0.0f);
    c.last_pos[i] = mk4(c.cam_eye, 0.0f);
    c.last_n[i] = mk4(d, 0.0f);
    c.pix[i] = pixel; c.sidx[i] = s; c.pstate[i] = 0u;
    c.live[i] = i;
}

// One bounce of queue slot q (cpu_bdpt.cpp:365-472 minus the connection loop, which runs from the CV record).
// Returns true if the path stays alive.
B2_HD bool shade_body(const RenderCtx &c, int q, uint32_t &n_emit_hits, uint32_t &n_cap) {
    const DevScene &sc = c.sc;
    const int path = c.live[q];
    const int2 hr = c.hit[q];
    c.cv[B2_CV_F4 * q].w = u2f(B2_CV_NONE);
    if (hr.x == B2_PRIM_MISS) return false;
    const float4 o4 = c.ray_o[path], d4 = c.ray_d[path];
    const float3 ro = mk3(o4), rd = mk3(d4);
    HitRec h; h.prim = hr.x; h.t = u2f((uint32_t) hr.y);
    const Surface s = surface_at(sc, ro, rd, h);
    const uint32_t ps = c.pstate[path];
    int depth = (int) (ps & 0xffu), n_delta = (int) (ps >> 8);
    const uint32_t pixel = c.pix[path];
    if (s.is_light && depth == 0) {                                        // :372-375
        if (valid_color(s.mtl.base)) {
            atomic_add_f32(&c.film[3 * pixel], s.mtl.base.x); atomic_add_f32(&c.film[3 * pixel + 1], s.mtl.base.y); atomic_add_f32(&c.film[3 * pixel + 2], s.mtl.base.z);
        }
        n_emit_hits++;
        return false;
    }
    const float3 last_pos = mk3(c.last_pos[path]), last_n = mk3(c.last_n[path]);
    const float3 wo = rd * -1.0f;
    float pdf_fwd = 1.0f;
    const float3 dp = s.pos - last_pos;
    const float dist2 = dot3(dp, dp);
    if (depth > 0) pdf_fwd = d4.w * fabsf(dot3(s.n, wo)) / fmaxf(dist2, 1e-6f);   // :378-382
    float3 thr = mk3(c.thr[path]);
    EyeVtx ev; ev.pos = s.pos; ev.n = s.n; ev.wo = wo; ev.thr = thr; ev.mtl = s.mtl; ev.depth = depth; ev.path = path;
    store_cv(c.cv, q, ev);
    float u_rr, u1, u2;
    rng_draw3(c.key, B2_RNG_EYE, pixel, c.sidx[path], (uint32_t) (depth + n_delta), u_rr, u1, u2);
    BsdfSample bs; bs.wi = mk3(0, 0, 0); bs.f = mk3(0, 0, 0); bs.pdf = 0.0f;
    bsdf_sample(s.mtl, wo, s.n, u_rr, u1, u2, o4.w, bs);
    if (bs.pdf <= 0.0f && !bs.is_delta) return false;
    if (bs.is_delta) {                                                     // :453-459
        thr = thr * bs.f;
        c.thr[path] = mk4(thr, 0.0f);
        c.ray_d[path] = mk4(bs.wi, 1.0f);
        c.ray_o[path] = mk4(s.pos + s.n * (dot3(bs.wi, s.n) < 0.0f ? -B2_EPSILON : B2_EPSILON), bs.new_eta);
        c.last_pos[path] = mk4(s.pos, 0.0f); c.last_n[path] = mk4(s.n, 0.0f);
        n_delta++;
        c.pstate[path] = (uint32_t) depth | ((uint32_t) n_delta << 8);
        if (n_delta >= c.max_delta) { n_cap++; return false; }
        return true;
    }
    const float pdf_rev = bsdf_pdf(s.mtl, bs.wi, wo, s.n) * fabsf(dot3(last_n, rd)) / fmaxf(dist2, 1e-6f);   // :461-465
    c.hist_pe[(size_t) depth * c.max_paths + path] = mk4(s.pos, s.mtl.eta);
    c.hist_pdf[(size_t) depth * c.max_paths + path] = make_float2(pdf_fwd, pdf_rev);
    thr = thr * bs.f * fabsf(dot3(s.n, bs.wi)) / bs.pdf;
    if (!valid_color(thr)) return false;
    depth++;
    if (depth >= c.eye_depth) return false;
    c.thr[path] = mk4(thr, 0.0f);
    c.ray_d[path] = mk4(bs.wi, bs.pdf);
    c.ray_o[path] = mk4(s.pos + s.n * B2_EPSILON, o4.w);
    c.last_pos[path] = mk4(s.pos, 0.0f); c.last_n[path] = mk4(s.n, 0.0f);
    c.pstate[path] = (uint32_t) depth | ((uint32_t) n_delta << 8);
    return true;
}

// =====================================================================================================
// MIS weight (cpu_calculate_mis_weight, cpu_bdpt.cpp:112-167) for eye vertex `ev` (s_idx = ev.depth) and pool
// vertex lv_index (t_idx = lv_index % light_depth).  d_vec is passed UNNORMALISED, as the reference does (:431).
// At connection time the current eye vertex still carries the placeholders pdf_fwd = 0, pdf_rev = 1 (:385).
// =====================================================================================================
B2_HD float mis_weight_body(const RenderCtx &c, const EyeVtx &ev, const LightVtx &lv, int lv_index, int t_idx, float3 d_vec, float dist2) {
    const int s_idx = ev.depth;
    const float3 ns = normalize3(ev.n), nt = normalize3(lv.n);
    const float3 md = d_vec * -1.0f;
    const float cos_s = fmaxf(0.0f, dot3(ns, d_vec));
    const float cos_t = fmaxf(0.0f, dot3(nt, md));
    if (cos_s <= 0.0f || cos_t <= 0.0f || dist2 < 1e-6f) return 0.0f;
    const float3 prev_e = (s_idx == 0) ? c.cam_eye : mk3(c.hist_pe[(size_t) (s_idx - 1) * c.max_paths + ev.path]);
    const float3 wo_s = normalize3(prev_e - ev.pos);
    const float3 wo_t = (t_idx == 0) ? normalize3(lv.n) : normalize3(mk3(c.pool[B2_LV_F4 * (lv_index - 1)]) - lv.pos);
    const float pdf_omega_s = fmaxf(bsdf_pdf(ev.mtl, wo_s, d_vec, ns), 1e-6f);
    const float pdf_omega_t = fmaxf(bsdf_pdf(lv.mtl, wo_t, md, nt), 1e-6f);
    const float pdf_s_to_t = pdf_omega_s * cos_t / dist2;
    const float pdf_t_to_s = pdf_omega_t * cos_s / dist2;
    float sum = 1.0f, ratio = 1.0f, prev_rev = pdf_t_to_s;
    for (int i = s_idx; i > 0; --i) {                                      // :143-148
        float eta_i, fwd_i, rev_i;
        if (i == s_idx) { eta_i = ev.mtl.eta; fwd_i = 0.0f; rev_i = 1.0f; }
        else {
            eta_i = c.hist_pe[(size_t) i * c.max_paths + ev.path].w;
            const float2 p = c.hist_pdf[(size_t) i * c.max_paths + ev.path];
            fwd_i = p.x; rev_i = p.y;
        }
        if (eta_i > 0.0f) break;
        ratio *= prev_rev / fmaxf(fwd_i, 1e-8f);
        sum += ratio;
        prev_rev = rev_i;
    }
    ratio = 1.0f; prev_rev = pdf_s_to_t;
    const int base = lv_index - t_idx;
    for (int i = t_idx; i > 0; --i) {                                      // :153-163
        const float4 f0 = c.pool[B2_LV_F4 * (base + i)], f2 = c.pool[B2_LV_F4 * (base + i) + 2];
        if (f2u(f0.w) & B2_LV_SRC) { ratio *= prev_rev / fmaxf(f2.w, 1e-8f); sum += ratio; break; }
        if (c.pool[B2_LV_F4 * (base + i) + 4].z > 0.0f) break;
        ratio *= prev_rev / fmaxf(f2.w, 1e-8f);
        sum += ratio;
        prev_rev = c.pool[B2_LV_F4 * (base + i) + 3].w;
    }
    if (isnan(sum) || isinf(sum) || sum <= 0.0f) return 0.0f;
    return 1.0f / sum;
}

// =====================================================================================================
// Connection of one eye vertex with one pool vertex (cpu_bdpt.cpp:389-439 up to the visibility query).
// Returns true when a shadow segment has to be traced; `e` then holds the queue entry (contribution already
// MIS-weighted and clamped; the shadow stage multiplies it by the 0/1 transmittance).
// =====================================================================================================
struct ShadowEntry { float3 o; float tmax; float3 d; uint32_t pixel; float3 contrib; };
B2_HD bool connect_pair(const RenderCtx &c, const EyeVtx &ev, uint32_t pixel, int lv_index, ShadowEntry &e) {
    const LightVtx lv = load_lv(c.pool, lv_index);
    const float3 d_vec = lv.pos - ev.pos;
    const float dist2 = dot3(d_vec, d_vec);
    if (dist2 < 1e-6f) return false;
    const float dist = sqrtf(dist2);
    const float3 wi = d_vec / dist;
    const float3 mwi = wi * -1.0f;
    const float cosE = fmaxf(0.0f, dot3(ev.n, wi));
    const float cosL = fmaxf(0.0f, dot3(lv.n, mwi));
    if (cosE <= 0.0f || cosL <= 0.0f) return false;
    const int t_idx = lv_index % c.light_depth;
    const bool is_src = lv.flags & B2_LV_SRC;
    if (is_src && lv.cutoff > 0.0f && !(lv.flags & B2_LV_PARALLEL)) {      // spot-cone gate, :405-409
        const int li = (lv_index / c.light_depth) % c.sc.n_lights;
        const float4 l3 = c.sc.lights[4 * li + 3];
        if (dot3(mk3(l3), mwi) < l3.w) return false;            // l3.w = cosf(light.cutoff) == cosf(lv.source_cutoff)
    }
    const float3 fE = bsdf_evaluate(ev.mtl, ev.wo, wi, ev.n);
    float3 fL = mk3(1.0f, 1.0f, 1.0f);
    if (!is_src && t_idx > 0) {
        const float3 wo_l = normalize3(mk3(c.pool[B2_LV_F4 * (lv_index - 1)]) - lv.pos);
        fL = bsdf_evaluate(lv.mtl, wo_l, mwi, lv.n);
    }
    if ((fE.x <= 0.0f && fE.y <= 0.0f && fE.z <= 0.0f) || (fL.x <= 0.0f && fL.y <= 0.0f && fL.z <= 0.0f)) return false;
    const float G = (cosE * cosL) / fmaxf(dist2, 1e-4f);
    const float w = mis_weight_body(c, ev, lv, lv_index, t_idx, d_vec, dist2);
    float3 contrib = ev.thr * fE * G * fL * lv.thr * w;
    contrib = valid_color(contrib) ? clamp_radiance(contrib, 15.0f) : mk3(0.0f, 0.0f, 0.0f);
    const ShadowRay sr = make_shadow_ray(x3_add(ev.pos, x3_muls(ev.n, B2_EPSILON)), x3_add(lv.pos, x3_muls(lv.n, B2_EPSILON)));   // :424
    e.o = sr.o; e.tmax = sr.tmax; e.d = sr.d; e.pixel = pixel; e.contrib = contrib;
    return true;
}
B2_HD void store_sq(float4 *sq, size_t slot, const ShadowEntry &e) {
    sq[B2_SQ_F4 * slot] = mk4(e.o, e.tmax);
    sq[B2_SQ_F4 * slot + 1] = mk4(e.d, u2f(e.pixel));
    sq[B2_SQ_F4 * slot + 2] = mk4(e.contrib, 0.0f);
}
// Shadow stage for queue slot k: returns true if the segment is unoccluded (and splats its contribution).
template <bool COUNT>
B2_HD bool shadow_body(const RenderCtx &c, size_t k, uint2 *stack, int sstride, TravCounters *tc) {
    const float4 a = c.sq[B2_SQ_F4 * k], b = c.sq[B2_SQ_F4 * k + 1];
    if (occluded_bvh<COUNT>(c.sc, mk3(a), mk3(b), a.w, stack, sstride, tc)) return false;
    const float4 cc = c.sq[B2_SQ_F4 * k + 2];
    const uint32_t pixel = f2u(b.w);
    if (cc.x != 0.0f) atomic_add_f32(&c.film[3 * pixel], cc.x);
    if (cc.y != 0.0f) atomic_add_f32(&c.film[3 * pixel + 1], cc.y);
    if (cc.z != 0.0f) atomic_add_f32(&c.film[3 * pixel + 2], cc.z);
    return true;
}
