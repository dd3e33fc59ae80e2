// b2_isect.cuh -- primitive tests (bit-exact restatement of the CPU intersector's arithmetic on the
// device) and the two traversal flavours: closest-hit (cpu_find_closest_hit, cpu_bdpt.cpp:30-80) and
// any-hit (cpu_check_visibility, cpu_bdpt.cpp:82-107), each as exhaustive and wide-BVH variants.
#pragma once
#include "b2_scene.cuh"

// Triangle::check_intersect, object.cpp:72-95 (Moeller-Trumbore, no culling, t == tMax accepted).
B2_HD bool tri_test(float3 ro, float3 rd, float3 v0, float3 e1, float3 e2, float tMin, float tMax, float &t) {
    float3 pvec = x3_cross(rd, e2);
    float det = x3_dot(e1, pvec);
    if (fabsf(det) <= 1e-6f) return false;        // (double)fabs(det) < 1e-6  <=>  |det| <= float(1e-6)
    float inv = x_div(1.0f, det);
    float3 tvec = x3_sub(ro, v0);
    float u = x_mul(x3_dot(tvec, pvec), inv);
    if (u < 0.0f || u > 1.0f) return false;
    float3 qvec = x3_cross(tvec, e1);
    float v = x_mul(x3_dot(rd, qvec), inv);
    if (v < 0.0f || x_add(u, v) > 1.0f) return false;
    float tt = x_mul(x3_dot(e2, qvec), inv);
    if (tt < tMin || tt > tMax) return false;
    t = tt;
    return true;
}
// The same test with contracted multiply-adds and one approximate reciprocal (~35 instead of ~75 instructions).  Only the
// render's shadow rays use it (k_connect's trace phase): their end points come out of shading arithmetic that already
// differs from the CPU's in the last bits, so an any-hit answer that can only differ for a ray grazing an edge within
// rounding is within the render's statistical parity; closest hits, b2_intersect and every parity test keep tri_test.
B2_HD bool tri_test_fast(float3 ro, float3 rd, float3 v0, float3 e1, float3 e2, float tMin, float tMax) {
    const float3 pvec = cross3(rd, e2);
    const float det = dot3(e1, pvec);
    if (fabsf(det) <= 1e-6f) return false;
    const float inv = fast_rcp(det);
    const float3 tvec = ro - v0;
    const float u = dot3(tvec, pvec) * inv;
    if (u < 0.0f || u > 1.0f) return false;
    const float3 qvec = cross3(tvec, e1);
    const float v = dot3(rd, qvec) * inv;
    if (v < 0.0f || u + v > 1.0f) return false;
    const float tt = dot3(e2, qvec) * inv;
    return !(tt < tMin || tt > tMax);
}
// Sphere::check_intersect, object.cpp:16-56 with scale (1,1,1): refuses origins on/inside the sphere.
// Conservative pre-reject of the two quadratic tests below: the discriminant with contracted arithmetic and
// no square root; "certainly negative" = below minus 1e-4 of the magnitude of its terms, two orders of magnitude more than the
// rounding of either form (a dozen fp32 operations, |rd| = 1 within 1e-6), so a rejected ray is one the exact test rejects too
// and hit ids / t stay bit-equal.  A camera ray of C2 ran the exact test (IEEE sqrt, divisions, re-normalisation) against every
// sphere and every light ball; almost all of them miss by far.
B2_HD bool quad_certainly_misses(float3 ro, float3 rd, float3 c, float r) {
    const float3 oc = ro - c;
    const float b = dot3(oc, rd), l2 = dot3(oc, oc), d2 = dot3(rd, rd), r2 = r * r;
    // exact forms: h = b^2 - (|oc|^2 - r^2) with |rd| = 1 (light balls); disc = 4 (b'^2 - a (|oc|^2 - r^2)), b' = b / |rd|, a = 1 (spheres)
    const float h = b * b - d2 * (l2 - r2);
    return h < -1e-4f * (b * b + d2 * (l2 + r2));
}
B2_HD bool sphere_test(float3 ro, float3 rd, float3 c, float r, float tMin, float tMax, float &t) {
    if (quad_certainly_misses(ro, rd, c, r)) return false;
    float3 D = g3_normalize(rd);
    float3 OC = x3_sub(ro, c);
    float a = x_add(x_add(x_mul(D.x, D.x), x_mul(D.y, D.y)), x_mul(D.z, D.z));
    float b = x_mul(2.0f, x_add(x_add(x_mul(D.x, OC.x), x_mul(D.y, OC.y)), x_mul(D.z, OC.z)));
    float cc = x_sub(x_add(x_add(x_mul(OC.x, OC.x), x_mul(OC.y, OC.y)), x_mul(OC.z, OC.z)), x_mul(r, r));
    if (cc <= 1e-6f) return false;
    float disc = x_sub(x_mul(b, b), x_mul(x_mul(4.0f, a), cc));
    if (disc < 0.0f) return false;
    float sd = x_sqrt(0.0f < disc ? disc : 0.0f);
    float two_a = x_mul(2.0f, a);
    float t0 = x_div(x_sub(-b, sd), two_a);
    float t1 = x_div(x_add(-b, sd), two_a);
    if (t0 > t1) { float tmp = t0; t0 = t1; t1 = tmp; }
    float tc = (t0 >= tMin) ? t0 : t1;
    if (tc < tMin || tc > tMax) return false;
    t = tc;
    return true;
}
// intersect_sphere (light balls), geometric.cuh:240-259: strict EPSILON < t < max_dist.
B2_HD bool light_ball_test(float3 ro, float3 rd, float3 c, float r, float max_dist, float &t) {
    if (quad_certainly_misses(ro, rd, c, r)) return false;
    float3 oc = x3_sub(ro, c);
    float b = x3_dot(oc, rd);
    float cc = x_sub(x3_dot(oc, oc), x_mul(r, r));
    float h = x_sub(x_mul(b, b), cc);
    if (h < 0.0f) return false;
    h = x_sqrt(h);
    float th = x_sub(-b, h);
    if (th > B2_EPSILON && th < max_dist) { t = th; return true; }
    th = x_add(-b, h);
    if (th > B2_EPSILON && th < max_dist) { t = th; return true; }
    return false;
}
// AABB::intersectAABB, object.cpp:104-121 (group pre-cull), with the ray's fixed [tmin,tmax].
B2_HD bool group_slab(float3 ro, float3 rd, float3 mn, float3 mx, float tmin, float tmax) {
    const float p[3] = { ro.x, ro.y, ro.z }, d[3] = { rd.x, rd.y, rd.z };
    const float lo[3] = { mn.x, mn.y, mn.z }, hi[3] = { mx.x, mx.y, mx.z };
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        float invD = x_div(1.0f, d[a]);
        float t0 = x_mul(x_sub(lo[a], p[a]), invD);
        float t1 = x_mul(x_sub(hi[a], p[a]), invD);
        if (invD < 0) { float tmp = t0; t0 = t1; t1 = tmp; }
        tmin = t0 > tmin ? t0 : tmin;
        tmax = t1 < tmax ? t1 : tmax;
        if (tmax <= tmin) return false;
    }
    return true;
}
B2_HD bool group_pass(const DevScene &sc, int g, float3 ro, float3 rd, float tmin, float tmax) {
    return group_slab(ro, rd, mk3(sc.groups[2 * g]), mk3(sc.groups[2 * g + 1]), tmin, tmax);
}

// The reference visits groups in key order and culls group g with intersectAABB(ray, 1e-4, best.t) (cpu_bdpt.cpp:43-44),
// best.t being the closest hit of the EARLIER groups.  For an order-free traversal that only matters on exact ties: a hit
// at t in group g loses against an equal-t hit of an earlier group iff that cull rejects g with tMax = t, i.e. iff the hit
// lies exactly on the near face of g's box (min(exit, t) <= max(entry, 1e-4)).  Class of a candidate hit:
//   0 = its group's box is missed altogether (tMax = infinity): no hit
//   1 = "pass": on a tie it behaves like the reference's t <= best.t acceptance -- the later object wins
//   2 = "fail": its group would be culled behind an equal-t hit of an earlier group -- it loses ties against every pass
//       hit, and among fail hits the EARLIEST wins (the first tied group is never culled, the later ones are)
B2_HD int group_tie_class(const DevScene &sc, int g, float3 ro, float3 rd, float tmin, float t) {
    const float4 mn4 = sc.groups[2 * g], mx4 = sc.groups[2 * g + 1];
    const float p[3] = { ro.x, ro.y, ro.z }, d[3] = { rd.x, rd.y, rd.z };
    const float lo[3] = { mn4.x, mn4.y, mn4.z }, hi[3] = { mx4.x, mx4.y, mx4.z };
    float tmax_inf = B2_T_INF, tmax_t = t;
    bool dead_inf = false, dead_t = false;
#pragma unroll
    for (int a = 0; a < 3; ++a) {                       // the arithmetic of group_slab (AABB::intersectAABB, object.cpp:104-121)
        float invD = x_div(1.0f, d[a]);
        float t0 = x_mul(x_sub(lo[a], p[a]), invD);
        float t1 = x_mul(x_sub(hi[a], p[a]), invD);
        if (invD < 0) { float tmp = t0; t0 = t1; t1 = tmp; }
        tmin = t0 > tmin ? t0 : tmin;
        tmax_inf = t1 < tmax_inf ? t1 : tmax_inf;
        tmax_t = t1 < tmax_t ? t1 : tmax_t;
        if (tmax_inf <= tmin) dead_inf = true;
        if (tmax_t <= tmin) dead_t = true;
    }
    return dead_inf ? 0 : (dead_t ? 2 : 1);
}

struct HitRec { int prim; float t; };
struct TravCounters { uint32_t nodes, tris; };

// Candidate acceptance in an order-free traversal: the reference visits objects in `order` and accepts
// t <= best (object.cpp:92), so the winner is min t, ties -> highest order.
B2_HD bool closer(float t, uint32_t order, int cls, float best_t, uint32_t best_order, int best_cls) {
    if (t != best_t) return t < best_t;
    if (cls != best_cls) return cls == 1;                // a pass hit beats a fail hit (group_tie_class)
    return cls == 1 ? order > best_order : order < best_order;
}
// group of a stored hit (-1: light ball, or no group cull in the pt / ppm semantics)
B2_HD int group_of_prim(const DevScene &sc, int prim) {
    if (prim == B2_PRIM_MISS || sc.gpu_sem) return -1;
    const int k = prim_kind(prim), i = prim_index(prim);
    if (k == B2_KIND_TRI) return (int) f2u(sc.tri[3 * i + 2].w);
    if (k == B2_KIND_SPHERE) return (int) f2u(sc.sph[2 * i + 1].z);
    return -1;
}
// A primitive test reported a hit at t <= best.t for object (order, group g): does it replace `best`?  Strictly closer hits
// (the common case) only need the reference's group cull with tMax = infinity; exact ties compare tie classes, which are pure
// functions of (group, ray, t) and are computed only then -- nothing is carried along in the traversal state.
B2_HD bool accept_hit(const DevScene &sc, float3 ro, float3 rd, float t, uint32_t order, int g, const HitRec &best, uint32_t best_order) {
    if (best.prim == B2_PRIM_MISS || t < best.t) return group_pass(sc, g, ro, rd, 1e-4f, B2_T_INF);
    if (t > best.t) return false;
    const int cls = group_tie_class(sc, g, ro, rd, 1e-4f, t);
    if (cls == 0) return false;
    const int bg = group_of_prim(sc, best.prim);
    const int best_cls = bg >= 0 ? group_tie_class(sc, bg, ro, rd, 1e-4f, best.t) : 1;
    return closer(t, order, cls, best.t, best_order, best_cls);
}

// ---- spheres + light balls (short tables, tested after the triangles) ----
// sc.gpu_sem (the pt / ppm integrators): ordinary spheres go through intersect_sphere like the light balls -- both roots,
// so a ray that starts inside a sphere hits its far side (find_closest_hit, geometric.cuh:337-351) -- and there is no group cull.
B2_HD void closest_spheres(const DevScene &sc, float3 ro, float3 rd, HitRec &best, uint32_t &best_order) {
    for (int i = 0; i < sc.n_sph; ++i) {
        float4 a = sc.sph[2 * i], b = sc.sph[2 * i + 1];
        float t;
        if (sc.gpu_sem) {
            if (light_ball_test(ro, rd, mk3(a), a.w, best.t, t)) { best.t = t; best.prim = prim_make(B2_KIND_SPHERE, i); best_order = f2u(b.x); }
            continue;
        }
        if (sphere_test(ro, rd, mk3(a), a.w, 1e-4f, best.t, t)) {
            uint32_t order = f2u(b.x);
            if (accept_hit(sc, ro, rd, t, order, (int) f2u(b.z), best, best_order)) { best.t = t; best.prim = prim_make(B2_KIND_SPHERE, i); best_order = order; }
        }
    }
}
B2_HD void closest_lights(const DevScene &sc, float3 ro, float3 rd, HitRec &best) {
    for (int k = 0; k < sc.n_lights; ++k) {                      // cpu_bdpt.cpp:64-78
        float4 a = sc.lights[4 * k];
        float t;
        if (light_ball_test(ro, rd, mk3(a), a.w, best.t, t)) { best.t = t; best.prim = prim_make(B2_KIND_LIGHT, k); }
    }
}
// sc.gpu_sem: check_visibility of geometric.cuh:293-325 -- intersect_sphere within (EPSILON, tMax) and t > tMin, every
// object opaque (its transparency test reads mtl_old, which the pt / ppm helpers never fill: taken as zero-initialised).
B2_HD bool occluded_spheres(const DevScene &sc, float3 ro, float3 rd, float tMin, float tMax) {
    for (int i = 0; i < sc.n_sph; ++i) {
        float4 a = sc.sph[2 * i], b = sc.sph[2 * i + 1];
        float t;
        if (sc.gpu_sem) {
            if (light_ball_test(ro, rd, mk3(a), a.w, tMax, t) && t > tMin) return true;
            continue;
        }
        if (f2u(b.y) & B2_MAT_TRANSPARENT) continue;
        if (sphere_test(ro, rd, mk3(a), a.w, tMin, tMax, t) && group_pass(sc, (int) f2u(b.z), ro, rd, tMin, tMax)) return true;
    }
    return false;
}

// Sphere::check_intersect with contracted multiply-adds, one approximate square root and reciprocal, and WITHOUT the
// re-normalisation of a direction that is already normalised within rounding (object.cpp:17: t differs by |rd| = 1 +- 1e-7):
// ~15 instructions for a ray that misses the sphere instead of ~80.  Same class as tri_test_fast: only the render's shadow
// rays use it (k_connect), where an answer can only change for a segment that grazes the sphere within rounding.
B2_HD bool sphere_test_fast(float3 ro, float3 rd, float3 c, float r, float tMin, float tMax) {
    const float3 OC = ro - c;
    const float a = dot3(rd, rd), hb = dot3(rd, OC), cc = dot3(OC, OC) - r * r;
    if (cc <= 1e-6f) return false;
    const float disc = hb * hb - a * cc;
    if (disc < 0.0f) return false;
    const float sd = fast_sqrt(disc), inv_a = fast_rcp(a);
    const float t0 = (-hb - sd) * inv_a, t1 = (-hb + sd) * inv_a;
    const float tc = (t0 >= tMin) ? t0 : t1;
    return !(tc < tMin || tc > tMax);
}

// ---- exhaustive reference kernels' bodies (parity cross-check of the BVH path) ----
B2_HD HitRec closest_hit_brute(const DevScene &sc, float3 ro, float3 rd) {
    HitRec best; best.prim = B2_PRIM_MISS; best.t = B2_T_INF;
    uint32_t best_order = 0;
    for (int i = 0; i < sc.n_tris; ++i) {
        float4 a = sc.tri[3 * i], b = sc.tri[3 * i + 1], c = sc.tri[3 * i + 2];
        float t;
        if (tri_test(ro, rd, mk3(a), mk3(b), mk3(c), 1e-4f, best.t, t)) {
            uint32_t order = f2u(a.w);
            if (accept_hit(sc, ro, rd, t, order, (int) f2u(c.w), best, best_order)) { best.t = t; best.prim = prim_make(B2_KIND_TRI, i); best_order = order; }
        }
    }
    closest_spheres(sc, ro, rd, best, best_order);
    closest_lights(sc, ro, rd, best);
    return best;
}
// Segment p1 -> p2 exactly as cpu_check_visibility sets it up (cpu_bdpt.cpp:83-91).
struct ShadowRay { float3 o, d; float tmax; };
B2_HD ShadowRay make_shadow_ray(float3 p1, float3 p2) {
    ShadowRay r;
    float3 diff = x3_sub(p2, p1);
    float dist = x3_length(diff);
    r.o = p1; r.d = x3_divs(diff, dist); r.tmax = x_sub(dist, 1e-3f);
    return r;
}
B2_HD bool occluded_brute(const DevScene &sc, float3 ro, float3 rd, float tMax) {
    const float tMin = 1e-3f;
    for (int i = 0; i < sc.n_tris; ++i) {
        float4 a = sc.tri[3 * i], b = sc.tri[3 * i + 1], c = sc.tri[3 * i + 2];
        if (f2u(b.w) & B2_MAT_TRANSPARENT) continue;
        float t;
        if (tri_test(ro, rd, mk3(a), mk3(b), mk3(c), tMin, tMax, t) && group_pass(sc, (int) f2u(c.w), ro, rd, tMin, tMax)) return true;
    }
    return occluded_spheres(sc, ro, rd, tMin, tMax);
}
