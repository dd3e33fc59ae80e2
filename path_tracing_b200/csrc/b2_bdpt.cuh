// b2_bdpt.cuh -- per-item bodies of the wavefront BDPT stages.  The estimator is the reference's CPU one
// (src/cpu_bdpt.cpp:173-488), quirk for quirk: one shared light-vertex pool per render,
// connections from every eye vertex (delta ones included) to every pool vertex, placeholder pdf_fwd = 0 at
// connection time, per-connection clamp at 15, no Russian roulette, delta bounces that do not advance depth.
// The __global__ wrappers and the host wave loop are in b2_render.cu; tests/hostsim runs the same bodies
// sequentially on the host.
//
// Stages of one wave (all paths of a batch advance one bounce):
//   closest  : ray -> (prim, t)                              [closest_hit_bvh]
//   shade    : hit -> eye vertex record (CV) + BSDF sample + next ray / termination; survivors are
//              compacted into the next wave's queue (warp ballot + prefix)
//   connect  : CV x pool vertex: cheap pre-tests (connect_pretest) for every pair; surviving pairs are
//              compacted in shared memory and finished at full lane occupancy (connect_finish: BSDFs,
//              shadow segment any-hit, MIS weight) into the path's radiance accumulator
//   flush    : at the end of a batch every path adds its accumulator to the film (atomicAdd)
#pragma once
#include "b2_bvh.cuh"
#include "b2_rng.cuh"

// ---- light-vertex pool record: 5 x float4 (80 B), CudaLightVertex of bdpt_cu.cuh:6-17 without the dead fields ----
//   f0 = {pos, flags}   f1 = {normal, source_cutoff}   f2 = {throughput, pdf_fwd}   f3 = {base_color, pdf_rev}
//   f4 = {roughness, metallic, eta, 0}
#define B2_LV_F4 5
#define B2_LV_SRC 1u
#define B2_LV_PARALLEL 2u
struct LightVtx { float3 pos, n, thr; Mat mtl; uint32_t flags; float cutoff, pdf_fwd, pdf_rev, walk; };   // walk: mis_light_sum, written by k_pool_active
B2_HD LightVtx load_lv(const float4 *pool, int i) {
    const float4 a = pool[B2_LV_F4 * i], b = pool[B2_LV_F4 * i + 1], c = pool[B2_LV_F4 * i + 2], d = pool[B2_LV_F4 * i + 3], e = pool[B2_LV_F4 * i + 4];
    LightVtx v;
    v.pos = mk3(a); v.flags = f2u(a.w); v.n = mk3(b); v.cutoff = b.w; v.thr = mk3(c); v.pdf_fwd = c.w;
    v.mtl.base = mk3(d); v.pdf_rev = d.w; v.mtl.roughness = e.x; v.mtl.metallic = e.y; v.mtl.eta = e.z; v.walk = e.w;
    return v;
}
B2_HD void store_lv(float4 *pool, int i, const LightVtx &v) {
    pool[B2_LV_F4 * i] = mk4(v.pos, u2f(v.flags));
    pool[B2_LV_F4 * i + 1] = mk4(v.n, v.cutoff);
    pool[B2_LV_F4 * i + 2] = mk4(v.thr, v.pdf_fwd);
    pool[B2_LV_F4 * i + 3] = mk4(v.mtl.base, v.pdf_rev);
    pool[B2_LV_F4 * i + 4] = make_float4(v.mtl.roughness, v.mtl.metallic, v.mtl.eta, 0.0f);
}

// ---- eye-vertex record handed from the extend stage to the connect stage: 5 x float4 (80 B) ----
//   f0 = {pos, depth | pixel << 8}  f1 = {normal, path}  f2 = {wo, roughness}  f3 = {throughput, metallic}  f4 = {base_color, eta}
// Records are appended compactly (one per eye vertex actually created this wave), so the connect stage has no idle threads.
#define B2_CV_F4 5
struct EyeVtx { float3 pos, n, wo, thr; Mat mtl; int depth, path; uint32_t pixel; };
B2_HD EyeVtx load_cv(const float4 *cv, int q) {
    const float4 a = cv[B2_CV_F4 * q], b = cv[B2_CV_F4 * q + 1], c = cv[B2_CV_F4 * q + 2], d = cv[B2_CV_F4 * q + 3], e = cv[B2_CV_F4 * q + 4];
    EyeVtx v;
    v.pos = mk3(a); v.depth = (int) (f2u(a.w) & 0xffu); v.pixel = f2u(a.w) >> 8; v.n = mk3(b); v.path = (int) f2u(b.w); v.wo = mk3(c); v.mtl.roughness = c.w;
    v.thr = mk3(d); v.mtl.metallic = d.w; v.mtl.base = mk3(e); v.mtl.eta = e.w;
    return v;
}
B2_HD void store_cv(float4 *cv, int q, const EyeVtx &v) {
    cv[B2_CV_F4 * q] = mk4(v.pos, u2f((uint32_t) v.depth | (v.pixel << 8)));
    cv[B2_CV_F4 * q + 1] = mk4(v.n, u2f((uint32_t) v.path));
    cv[B2_CV_F4 * q + 2] = mk4(v.wo, v.mtl.roughness);
    cv[B2_CV_F4 * q + 3] = mk4(v.thr, v.mtl.metallic);
    cv[B2_CV_F4 * q + 4] = mk4(v.mtl.base, v.mtl.eta);
}

// ---- connection pre-test record of an ACTIVE pool vertex: 3 x float4 (48 B), built once per render ----
//   r0 = {pos, cos(cutoff) of the spot-cone gate or B2_NO_GATE}  r1 = {normal, pool index}  r2 = {normalize(light.dir), 0}
#define B2_PRE_F4 3
#define B2_NO_GATE (-2.0f)

enum { B2_CNT_EYE_SAMPLES = 0, B2_CNT_CLOSEST, B2_CNT_SHADOW, B2_CNT_PATH_SAMPLES, B2_CNT_DELTA_CAP, B2_CNT_NODES, B2_CNT_TRIS, B2_CNT_PAIRS, B2_CNT_N };

struct RenderCtx {
    DevScene sc;
    float3 cam_eye, cUL, cdx, cdy;          // cpu_bdpt.cpp:190-200
    int W, H, spl, eye_depth, light_depth, max_delta;
    uint32_t w_magic, npix_magic;           // udiv_magic_of(W), udiv_magic_of(W * H): see udiv_magic()
    int sample_offset, sample_stride;
    uint2 key;
    // light-vertex pool
    float4 *pool; int n_pool;
    float4 *pre; int *n_pool_active;        // pre-test records of the active pool vertices, ascending pool index
    // per-path state, capacity max_paths; a slot is allocated when a path creates its first eye vertex
    int max_paths;
    float4 *ray_o;      // {origin, current eta}
    float4 *ray_d;      // {direction, last_pdf_omega}
    float4 *thr;        // {throughput, depth | n_delta << 8}
    float4 *last_pos;   // {last_pos, pixel}
    float4 *last_n;     // {last_normal, sample index}
    float4 *hist_pe;    // [eye_depth][max_paths] {pos, eta} of stored (non-delta) eye vertices
    float2 *hist_pdf;   // [eye_depth][max_paths] {pdf_fwd, pdf_rev}
    int *live;          // queue of live path slots for this wave
    int *live_next; int *n_live_next;
    int *n_paths;       // slots handed out in this batch
    float4 *cv; int *n_cv;   // eye-vertex records created this wave
    unsigned int *conn_cursor;   // next unclaimed work item of the connection stage (zeroed with n_live_next / n_cv every wave)
    float *film;        // W*H*3 running sums
    unsigned long long *counters;
};

// Unsigned division by a per-render constant d without the ~20-instruction (32-bit) / ~70-instruction (64-bit) division
// sequences: with m = floor(2^32 / d), q = mulhi(n, m) is floor(n / d) or one less, for every 32-bit n.
inline uint32_t udiv_magic_of(uint32_t d) { return d <= 1u ? 0xffffffffu : (uint32_t) (0x100000000ull / d); }
B2_HD uint32_t udiv_magic(uint32_t n, uint32_t d, uint32_t m, uint32_t &rem) {
    uint32_t q = mulhi32(n, m), r = n - q * d;
    if (r >= d) { q++; r -= d; }
    rem = r;
    return q;
}

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
        if (sc.ggx_fix) mat_mark_fix(s.mtl);
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
        const float pdf_fwd = fast_div(last_pdf * cos_hit, dist2);
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
        v.pdf_rev = fast_div(bsdf_pdf(s.mtl, bs.wi, wo, s.n) * cos_prev, dist2);
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
// Register-resident state of one eye subpath between two bounces.
struct PathState {
    float3 ro, rd; float eta, last_pdf;     // next ray, current medium eta, solid-angle pdf of the last sampled direction
    float3 thr, last_pos, last_n;
    uint32_t pixel, sidx; int depth, n_delta;
};
B2_HD PathState load_path(const RenderCtx &c, int p) {
    const float4 o = c.ray_o[p], d = c.ray_d[p], t = c.thr[p], lp = c.last_pos[p], ln = c.last_n[p];
    PathState s;
    s.ro = mk3(o); s.eta = o.w; s.rd = mk3(d); s.last_pdf = d.w; s.thr = mk3(t); s.depth = (int) (f2u(t.w) & 0xffu); s.n_delta = (int) (f2u(t.w) >> 8);
    s.last_pos = mk3(lp); s.pixel = f2u(lp.w); s.last_n = mk3(ln); s.sidx = f2u(ln.w);
    return s;
}
B2_HD void store_path(const RenderCtx &c, int p, const PathState &s) {
    c.ray_o[p] = mk4(s.ro, s.eta); c.ray_d[p] = mk4(s.rd, s.last_pdf);
    c.thr[p] = mk4(s.thr, u2f((uint32_t) s.depth | ((uint32_t) s.n_delta << 8)));
    c.last_pos[p] = mk4(s.last_pos, u2f(s.pixel)); c.last_n[p] = mk4(s.last_n, u2f(s.sidx));
}
// Work item w of a render: pixel = w % npix, local sample = w / npix (cpu_bdpt.cpp:344-361): jittered primary ray.
// A batch starts at w0 = s0 * npix + pix0 (split on the host); item q of the batch is w0 + q, pix0 + q < 2^32.
B2_HD PathState primary_path(const RenderCtx &c, uint32_t pix0, int s0, uint32_t q) {
    PathState s;
    const uint32_t ds = udiv_magic(pix0 + q, (uint32_t) (c.W * c.H), c.npix_magic, s.pixel);
    s.sidx = (uint32_t) (c.sample_offset + (s0 + (int) ds) * c.sample_stride);
    float jx, jy;
    rng_draw2(c.key, B2_RNG_EYE, s.pixel, s.sidx, jx, jy);
    uint32_t col;
    const uint32_t row = udiv_magic(s.pixel, (uint32_t) c.W, c.w_magic, col);
    const float px = x_add((float) col, jx), py = x_add((float) row, jy);
    const float3 pp = x3_add(x3_add(c.cUL, x3_muls(c.cdx, px)), x3_muls(c.cdy, py));
    s.rd = f3_normalize(x3_sub(pp, c.cam_eye));
    s.ro = c.cam_eye; s.eta = 1.0f; s.last_pdf = 1.0f;
    s.thr = mk3(1.0f, 1.0f, 1.0f); s.last_pos = c.cam_eye; s.last_n = s.rd;
    s.depth = 0; s.n_delta = 0;
    return s;
}

// One bounce (cpu_bdpt.cpp:365-472 minus the connection loop, which runs later from the eye-vertex record).
struct ShadeOut {
    bool has_cv;        // an eye vertex was created: `ev` is valid (its .path is filled in by the caller)
    bool alive;         // the path continues: `ps` holds the next ray
    bool store_hist;    // non-delta vertex of a continuing path: hist[ev.depth] = {pos, eta}, {pdf_fwd, pdf_rev}
    bool emitter_hit;   // depth-0 light hit: `emitted` goes to the film (:372-375)
    bool delta_cap;
    EyeVtx ev; float pdf_fwd, pdf_rev; float3 emitted;
};
B2_HD void shade_step(const RenderCtx &c, PathState &ps, HitRec h, ShadeOut &o) {
    o.has_cv = o.alive = o.store_hist = o.emitter_hit = o.delta_cap = false;
    if (h.prim == B2_PRIM_MISS) return;
    const Surface s = surface_at(c.sc, ps.ro, ps.rd, h);
    if (s.is_light && ps.depth == 0) {                                     // :372-375
        o.emitter_hit = true; o.emitted = s.mtl.base;
        return;
    }
    if (s.is_light && c.sc.ggx_fix) return;                                // outside parity mode a light ball is only ever seen directly (b2_strategies.cuh)
    const float3 wo = ps.rd * -1.0f;
    float pdf_fwd = 1.0f;
    const float3 dp = s.pos - ps.last_pos;
    const float dist2 = dot3(dp, dp);
    if (ps.depth > 0) pdf_fwd = fast_div(ps.last_pdf * fabsf(dot3(s.n, wo)), fmaxf(dist2, 1e-6f));   // :378-382
    o.has_cv = true;
    o.ev.pos = s.pos; o.ev.n = s.n; o.ev.wo = wo; o.ev.thr = ps.thr; o.ev.mtl = s.mtl; o.ev.depth = ps.depth; o.ev.pixel = ps.pixel; o.ev.path = -1;
    float u_rr, u1, u2;
    rng_draw3(c.key, B2_RNG_EYE, ps.pixel, ps.sidx, (uint32_t) (ps.depth + ps.n_delta), u_rr, u1, u2);
    BsdfSample bs; bs.wi = mk3(0, 0, 0); bs.f = mk3(0, 0, 0); bs.pdf = 0.0f;
    bsdf_sample(s.mtl, wo, s.n, u_rr, u1, u2, ps.eta, bs);
    if (bs.pdf <= 0.0f && !bs.is_delta) return;
    if (bs.is_delta) {                                                     // :453-459
        ps.thr = ps.thr * bs.f;
        ps.rd = bs.wi; ps.last_pdf = 1.0f;
        ps.ro = s.pos + s.n * (dot3(bs.wi, s.n) < 0.0f ? -B2_EPSILON : B2_EPSILON); ps.eta = bs.new_eta;
        ps.last_pos = s.pos; ps.last_n = s.n;
        ps.n_delta++;
        if (ps.n_delta >= c.max_delta) { o.delta_cap = true; return; }
        o.alive = true;
        return;
    }
    o.pdf_fwd = pdf_fwd;
    o.pdf_rev = fast_div(bsdf_pdf(s.mtl, bs.wi, wo, s.n) * fabsf(dot3(ps.last_n, ps.rd)), fmaxf(dist2, 1e-6f));   // :461-465
    ps.thr = ps.thr * bs.f * fabsf(dot3(s.n, bs.wi)) / bs.pdf;
    if (!valid_color(ps.thr)) return;
    ps.depth++;
    if (ps.depth >= c.eye_depth) return;
    ps.rd = bs.wi; ps.last_pdf = bs.pdf;
    ps.ro = s.pos + s.n * B2_EPSILON;
    ps.last_pos = s.pos; ps.last_n = s.n;
    o.alive = true; o.store_hist = true;
}

// =====================================================================================================
// MIS weight (cpu_calculate_mis_weight, cpu_bdpt.cpp:112-167), written once over vertex accessors so the
// render (per-path history arrays + pool records) and the table kernel behind b2_mis_weight (flat arrays)
// share it.  d_vec is passed UNNORMALISED, as the reference does (:431).
//   EA: pos(i) eta(i) pdf_fwd(i) pdf_rev(i) of eye vertex i; cur_n / cur_mtl of vertex s_idx
//   LA: pos(i) eta(i) pdf_fwd(i) pdf_rev(i) is_src(i) of light vertex i of the SAME light subpath; cur_n / cur_mtl of t_idx
// =====================================================================================================
// The eye-side walk (:143-148) is LINEAR in pdf_t_to_s, the only term of it that depends on the light vertex: every ratio
// carries that factor exactly once.  mis_eye_sum is the walk for pdf_t_to_s = 1; the connection stage computes it once per eye
// vertex and work item (k_connect) instead of once per visible pair -- at depth d a pair read 2 (d - 1) scattered history
// rows, which made the deep waves of C3 fetch up to 10 GB from HBM per launch (profiles/r02_connect_summary.md, section 5).
// Same terms as the sequential walk, summed before instead of after the multiplication by pdf_t_to_s (fp32 rounding only).
template <class EA>
B2_HD float mis_eye_sum(const EA &E, int s_idx) {
    float sum = 0.0f, ratio = 1.0f, prev_rev = 1.0f;
    for (int i = s_idx; i > 0; --i) {
        if (E.eta(i) > 0.0f) break;
        ratio *= fast_div(prev_rev, fmaxf(E.pdf_fwd(i), 1e-8f));
        sum += ratio;
        prev_rev = E.pdf_rev(i);
    }
    return sum;
}
// The light-side walk (:153-163) is linear in pdf_s_to_t in the same way; its value for pdf_s_to_t = 1 is a constant of the pool
// vertex, stored in the spare slot of its record when the pool is compacted (k_pool_active).
template <class LA>
B2_HD float mis_light_sum(const LA &L, int t_idx) {
    float sum = 0.0f, ratio = 1.0f, prev_rev = 1.0f;
    for (int i = t_idx; i > 0; --i) {
        if (L.is_src(i)) { ratio *= fast_div(prev_rev, fmaxf(L.pdf_fwd(i), 1e-8f)); sum += ratio; break; }
        if (L.eta(i) > 0.0f) break;
        ratio *= fast_div(prev_rev, fmaxf(L.pdf_fwd(i), 1e-8f));
        sum += ratio;
        prev_rev = L.pdf_rev(i);
    }
    return sum;
}
// eye_sum / light_sum < 0: walk that subpath here (the table kernel, the one-thread-per-sample strategies); >= 0: mis_eye_sum(E, s_idx)
template <class EA, class LA>
B2_HD float mis_weight_core(const EA &E, int s_idx, const LA &L, int t_idx, float3 d_vec, float dist2, float3 cam_pos, float eye_sum = -1.0f, float light_sum = -1.0f) {
    if (s_idx < 0 || t_idx < 0) return 0.0f;
    const float3 ns = normalize3(E.cur_n()), nt = normalize3(L.cur_n());
    const float3 md = d_vec * -1.0f;
    const float cos_s = fmaxf(0.0f, dot3(ns, d_vec));
    const float cos_t = fmaxf(0.0f, dot3(nt, md));
    if (cos_s <= 0.0f || cos_t <= 0.0f || dist2 < 1e-6f) return 0.0f;
    const float3 ev_pos = E.pos(s_idx), lv_pos = L.pos(t_idx);
    const float3 wo_s = normalize3(((s_idx == 0) ? cam_pos : E.pos(s_idx - 1)) - ev_pos);
    const float3 wo_t = (t_idx == 0) ? normalize3(L.cur_n()) : normalize3(L.pos(t_idx - 1) - lv_pos);
#if B2_MAT_LOCAL_COPY
    const Mat m_s = E.cur_mtl(), m_t = L.cur_mtl();     // the noinline callee takes an address: of these 24 B, not of the whole vertex record
#else
    const Mat &m_s = E.cur_mtl(), &m_t = L.cur_mtl();
#endif
    const float pdf_omega_s = fmaxf(bsdf_pdf(m_s, wo_s, d_vec, ns), 1e-6f);
    const float pdf_omega_t = fmaxf(bsdf_pdf(m_t, wo_t, md, nt), 1e-6f);
    const float inv_d2 = fast_rcp(dist2);
    const float pdf_s_to_t = pdf_omega_s * cos_t * inv_d2;
    const float pdf_t_to_s = pdf_omega_t * cos_s * inv_d2;
    float sum = 1.0f, ratio = 1.0f, prev_rev = pdf_t_to_s;
    if (eye_sum >= 0.0f) sum += pdf_t_to_s * eye_sum;
    else for (int i = s_idx; i > 0; --i) {                                 // :143-148
        if (E.eta(i) > 0.0f) break;
        ratio *= fast_div(prev_rev, fmaxf(E.pdf_fwd(i), 1e-8f));
        sum += ratio;
        prev_rev = E.pdf_rev(i);
    }
    ratio = 1.0f; prev_rev = pdf_s_to_t;
    if (light_sum >= 0.0f) sum += pdf_s_to_t * light_sum;
    else for (int i = t_idx; i > 0; --i) {                                 // :153-163
        if (L.is_src(i)) { ratio *= fast_div(prev_rev, fmaxf(L.pdf_fwd(i), 1e-8f)); sum += ratio; break; }
        if (L.eta(i) > 0.0f) break;
        ratio *= fast_div(prev_rev, fmaxf(L.pdf_fwd(i), 1e-8f));
        sum += ratio;
        prev_rev = L.pdf_rev(i);
    }
    if (isnan(sum) || isinf(sum) || sum <= 0.0f) return 0.0f;
    return fast_rcp(sum);
}

// Render-side accessors.  At connection time the current eye vertex still carries the placeholders
// pdf_fwd = 0, pdf_rev = 1 (cpu_bdpt.cpp:385); earlier vertices come from the path's history arrays.
struct RenderEyeAcc {
    const RenderCtx &c; const EyeVtx &ev;
    B2_HD float3 pos(int i) const { return i == ev.depth ? ev.pos : mk3(c.hist_pe[(size_t) i * c.max_paths + ev.path]); }
    B2_HD float eta(int i) const { return i == ev.depth ? ev.mtl.eta : c.hist_pe[(size_t) i * c.max_paths + ev.path].w; }
    B2_HD float pdf_fwd(int i) const { return i == ev.depth ? 0.0f : c.hist_pdf[(size_t) i * c.max_paths + ev.path].x; }
    B2_HD float pdf_rev(int i) const { return i == ev.depth ? 1.0f : c.hist_pdf[(size_t) i * c.max_paths + ev.path].y; }
    B2_HD float3 cur_n() const { return ev.n; }
    B2_HD const Mat &cur_mtl() const { return ev.mtl; }
};
struct RenderLightAcc {
    const RenderCtx &c; const LightVtx &lv; int base;        // base = pool index of the subpath's vertex 0
    B2_HD float3 pos(int i) const { return mk3(c.pool[B2_LV_F4 * (base + i)]); }
    B2_HD float eta(int i) const { return c.pool[B2_LV_F4 * (base + i) + 4].z; }
    B2_HD float pdf_fwd(int i) const { return c.pool[B2_LV_F4 * (base + i) + 2].w; }
    B2_HD float pdf_rev(int i) const { return c.pool[B2_LV_F4 * (base + i) + 3].w; }
    B2_HD bool is_src(int i) const { return (f2u(c.pool[B2_LV_F4 * (base + i)].w) & B2_LV_SRC) != 0u; }
    B2_HD float3 cur_n() const { return lv.n; }
    B2_HD const Mat &cur_mtl() const { return lv.mtl; }
};

// =====================================================================================================
// Connection of one eye vertex with one pool vertex (cpu_bdpt.cpp:389-439), split in two:
//   connect_pretest : the geometric rejects up to and including the spot-cone gate (:391-409); needs only
//                     pos/normal of both ends, so it is what runs for EVERY (eye vertex, pool vertex) pair
//   connect_finish  : BSDFs, black skip, visibility, G, MIS, clamp (:411-438) for the surviving pairs
// Pool vertices with |throughput| < 1e-6 (:391) never reach either: they are not in the active list.
// =====================================================================================================
B2_HD bool connect_pretest(const RenderCtx &c, float3 ev_pos, float3 ev_n, int lv_index) {
    const float4 f0 = c.pool[B2_LV_F4 * lv_index], f1 = c.pool[B2_LV_F4 * lv_index + 1];
    const float3 d_vec = mk3(f0) - ev_pos;
    const float dist2 = dot3(d_vec, d_vec);
    if (dist2 < 1e-6f) return false;
    const float dist = sqrtf(dist2);
    const float3 wi = d_vec / dist;
    const float3 mwi = wi * -1.0f;
    const float cosE = fmaxf(0.0f, dot3(ev_n, wi));
    const float cosL = fmaxf(0.0f, dot3(mk3(f1), mwi));
    if (cosE <= 0.0f || cosL <= 0.0f) return false;
    const uint32_t flags = f2u(f0.w);
    if ((flags & B2_LV_SRC) && f1.w > 0.0f && !(flags & B2_LV_PARALLEL)) {  // spot-cone gate, :405-409
        const int li = (lv_index / c.light_depth) % c.sc.n_lights;
        const float4 l3 = c.sc.lights[4 * li + 3];              // {normalize(light.dir), cosf(light.cutoff)}
        if (dot3(mk3(l3), mwi) < l3.w) return false;
    }
    return true;
}
// The same tests from a pre-test record, division-free: cos > 0 <=> dot(n, d) > 0, and the cone gate
// dot(l, -d/|d|) < cos(cutoff) <=> -dot(l, d) < cos(cutoff) * |d|.  (Sign-equivalent up to fp32 rounding exactly at the
// boundary; the cosines are recomputed as the reference does in connect_finish.)
B2_HD void make_pretest_rec(const RenderCtx &c, int lv_index, float4 *rec) {
    const float4 f0 = c.pool[B2_LV_F4 * lv_index], f1 = c.pool[B2_LV_F4 * lv_index + 1];
    const uint32_t flags = f2u(f0.w);
    float gate = B2_NO_GATE; float3 ldir = mk3(0, 0, 0);
    if ((flags & B2_LV_SRC) && f1.w > 0.0f && !(flags & B2_LV_PARALLEL)) {
        const float4 l3 = c.sc.lights[4 * ((lv_index / c.light_depth) % c.sc.n_lights) + 3];
        gate = l3.w; ldir = mk3(l3);
    }
    rec[0] = mk4(mk3(f0), gate); rec[1] = mk4(mk3(f1), u2f((uint32_t) lv_index)); rec[2] = mk4(ldir, 0.0f);
}
B2_HD bool pretest_rec(float3 ev_pos, float3 ev_n, float4 r0, float4 r1, float4 r2) {
    const float3 d = mk3(r0) - ev_pos;
    const float dist2 = dot3(d, d);
    bool pass = !(dist2 < 1e-6f) && dot3(ev_n, d) > 0.0f && dot3(mk3(r1), d) < 0.0f;
    if (r0.w > -1.5f) pass = pass && !(-dot3(mk3(r2), d) < r0.w * fast_sqrt(dist2));
    return pass;
}
// connect_finish in three steps, so that the connection kernel can run each at full lane occupancy with the shadow
// rays of a whole warp traced in between (k_connect):
//   connect_prepare : BSDFs, black skip (:411-422), shadow ray (:424), pc = everything of :433 but the MIS weight
//   (visibility)    : occluded_bvh / any_step on the shadow ray
//   connect_weight  : MIS weight (:431), validity + clamp (:435-437)
B2_HD bool connect_prepare(const RenderCtx &c, const EyeVtx &ev, int lv_index, ShadowRay &sr, float3 &pc) {
    const LightVtx lv = load_lv(c.pool, lv_index);
    const float3 d_vec = lv.pos - ev.pos;
    const float dist2 = dot3(d_vec, d_vec);
    const float3 wi = div_sqrt3(d_vec, dist2);
    const float3 mwi = wi * -1.0f;
    const float cosE = fmaxf(0.0f, dot3(ev.n, wi));
    const float cosL = fmaxf(0.0f, dot3(lv.n, mwi));
    const int t_idx = lv_index % c.light_depth;
    const bool is_src = (lv.flags & B2_LV_SRC) != 0u;
#if B2_MAT_LOCAL_COPY
    const Mat m_e = ev.mtl, m_l = lv.mtl;
#else
    const Mat &m_e = ev.mtl, &m_l = lv.mtl;
#endif
    const float3 fE = bsdf_evaluate(m_e, ev.wo, wi, ev.n);                    // :411-412
    float3 fL = mk3(1.0f, 1.0f, 1.0f);
    if (!is_src && t_idx > 0) {                                               // :413-419 (previous STORED vertex)
        const float3 wo_l = normalize3(mk3(c.pool[B2_LV_F4 * (lv_index - 1)]) - lv.pos);
        fL = bsdf_evaluate(m_l, wo_l, mwi, lv.n);
    }
    if ((fE.x <= 0.0f && fE.y <= 0.0f && fE.z <= 0.0f) || (fL.x <= 0.0f && fL.y <= 0.0f && fL.z <= 0.0f)) return false;
    sr = make_shadow_ray(x3_add(ev.pos, x3_muls(ev.n, B2_EPSILON)), x3_add(lv.pos, x3_muls(lv.n, B2_EPSILON)));   // :424
    const float G = fast_div(cosE * cosL, fmaxf(dist2, 1e-4f));              // :427
    pc = ev.thr * fE * G * fL * lv.thr;                                       // :433 without w (transmittance == 1 here)
    return true;
}
// eye_sum: mis_eye_sum of ev, or < 0; pool_walk: the pool records carry mis_light_sum (the render's pool; not the host simulator's)
B2_HD float3 connect_weight(const RenderCtx &c, const EyeVtx &ev, int lv_index, float3 pc, float eye_sum = -1.0f, bool pool_walk = false) {
    const LightVtx lv = load_lv(c.pool, lv_index);
    const float3 d_vec = lv.pos - ev.pos;
    const float dist2 = dot3(d_vec, d_vec);
    const int t_idx = lv_index % c.light_depth;
    const RenderEyeAcc EA{ c, ev };
    const RenderLightAcc LA{ c, lv, lv_index - t_idx };
    const float w = mis_weight_core(EA, ev.depth, LA, t_idx, d_vec, dist2, c.cam_eye, eye_sum, pool_walk ? lv.walk : -1.0f);   // :431
    const float3 contrib = pc * w;
    return valid_color(contrib) ? clamp_radiance(contrib, 15.0f) : mk3(0.0f, 0.0f, 0.0f);   // :435-437
}
// Returns the (validated, clamped) contribution, zero if the pair contributes nothing.  n_shadow / n_path count
// visibility queries (:424) and connections that reach the MIS evaluation (:431).
template <bool COUNT>
B2_HD float3 connect_finish(const RenderCtx &c, const EyeVtx &ev, int lv_index, uint2 *stack, int sstride, TravCounters *tc,
                            uint32_t &n_shadow, uint32_t &n_path) {
    const float3 zero = mk3(0.0f, 0.0f, 0.0f);
    ShadowRay sr; float3 pc;
    if (!connect_prepare(c, ev, lv_index, sr, pc)) return zero;
    n_shadow++;
    if (occluded_bvh<COUNT>(c.sc, sr.o, sr.d, sr.tmax, stack, sstride, tc)) return zero;
    n_path++;
    return connect_weight(c, ev, lv_index, pc);
}
