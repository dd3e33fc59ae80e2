// tests/hostsim/hostsim.cpp -- TEST-ONLY host compilation of the device headers (b2_*.cuh).
// Purpose: debug the BVH build, the traversal and the wavefront stage logic on the CPU-only dev box, and
// compare them with the oracle in `-m "not gpu"` tests.  This library is NOT part of the product: libb2pt.so
// contains no host execution path, and nothing under path_tracing_b200/ loads this file.
#define B2_TRACK_STACK_DEPTH 1      // csrc/b2_bvh.cuh records the deepest traversal stack (hs_max_stack_depth)
#include "hostsim_bvh.hpp"
#include <cstdio>

struct HsScene {
    HostScene hs;
    HostBvh bvh;
    DevScene view;
};

static void hs_make_view(HsScene *s) {
    DevScene &v = s->view;
    v.tri = s->bvh.tri.data(); v.tri_obj = s->bvh.tri_obj.data(); v.nodes = s->bvh.nodes.data();
    v.sph = s->hs.sph.data(); v.lights = s->hs.lights.data(); v.mats = s->hs.mats.data(); v.groups = s->hs.groups.data();
    v.n_tris = (int) s->hs.tris.size(); v.n_nodes = (int) (s->bvh.nodes.size() / B2_NODE_UINT4);
    v.n_sph = s->hs.n_sph; v.n_lights = s->hs.n_lights; v.n_mats = s->hs.n_mats; v.n_groups = s->hs.n_groups;
    v.scene_min = s->hs.scene_min; v.scene_max = s->hs.scene_max;
    v.ggx_fix = 0; v.gpu_sem = 0;                       // parity mode
}

extern "C" {

void *hs_scene_create(const B2Tri *tris, size_t n_tris, const B2Sphere *sph, size_t n_sph, const B2Material *mats, size_t n_mats,
                      const B2Light *lights, size_t n_lights, int max_leaf, float pad_rel) {
    HsScene *s = new HsScene;
    const char *err = nullptr;
    if (!flatten_scene(tris, n_tris, sph, n_sph, mats, n_mats, lights, n_lights, s->hs, &err)) { fprintf(stderr, "hostsim: %s\n", err); delete s; return nullptr; }
    // max_leaf: 1..3; + 16 selects the radix-tree builder (B2_BUILDER_LBVH) instead of PLOC
    const int builder = (max_leaf & 16) ? B2_BUILDER_LBVH : B2_BUILDER_PLOC;
    const int radius = (max_leaf >> 8) & 0xff;                     // bits 8..15: PLOC search radius (0 = default 2)
    max_leaf &= 15;
    host_build_bvh(s->hs, max_leaf > 0 ? max_leaf : 3, pad_rel > 0 ? pad_rel : 3.8146973e-6f, s->bvh, builder, radius > 0 ? radius : 2);
    hs_make_view(s);
    return s;
}
void hs_scene_destroy(void *h) { delete (HsScene *) h; }
void hs_bvh_info(void *h, int *n_nodes, int *depth) { HsScene *s = (HsScene *) h; *n_nodes = s->view.n_nodes; *depth = s->bvh.depth; }

static int public_id(const DevScene &sc, int prim) {
    if (prim == B2_PRIM_MISS) return -1;
    int k = prim_kind(prim), i = prim_index(prim);
    if (k == B2_KIND_TRI) return sc.tri_obj[i];
    if (k == B2_KIND_SPHERE) return (int) f2u(sc.sph[2 * i + 1].w);
    return -2 - i;
}
// mode: 0 closest BVH, 1 any BVH, 16 closest brute, 17 any brute.  counters: [nodes, tris] totals (may be NULL)
// deepest traversal stack (entries) any walk of this library reached since the last reset
int hs_max_stack_depth(int reset) { const int d = b2_max_stack_depth; if (reset) b2_max_stack_depth = 0; return d; }

void hs_intersect(void *h, const float *org, const float *dir, size_t n, int mode, int32_t *prim_id, float *t, uint64_t *counters) {
    HsScene *s = (HsScene *) h;
    const DevScene &sc = s->view;
    uint2 stack[B2_STACK_CAP];
    TravCounters c{ 0, 0 };
    uint64_t cn = 0, ct = 0;
    for (size_t i = 0; i < n; ++i) {
        float3 a = make_float3(org[3 * i], org[3 * i + 1], org[3 * i + 2]), b = make_float3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]);
        c.nodes = c.tris = 0;
        if ((mode & 1) == 0) {
            HitRec r = (mode & 16) ? closest_hit_brute(sc, a, b) : closest_hit_bvh<true>(sc, a, b, stack, 1, &c);
            prim_id[i] = public_id(sc, r.prim);
            if (t) t[i] = r.t;
        } else {
            ShadowRay sr = make_shadow_ray(a, b);
            bool occ = (mode & 16) ? occluded_brute(sc, sr.o, sr.d, sr.tmax) : occluded_bvh<true>(sc, sr.o, sr.d, sr.tmax, stack, 1, &c);
            prim_id[i] = occ ? 0 : 1;
        }
        cn += c.nodes; ct += c.tris;
    }
    if (counters) { counters[0] = cn; counters[1] = ct; }
}

}  // extern "C"

#include "hostsim_render.hpp"
#include "hostsim_pt.hpp"
#include "hostsim_strategies.hpp"
extern "C" {
// integrator 1 = pt, 2 = ppm (B2RenderParams.integrator)
void hs_render_other(void *h, const B2Camera *cam, const B2RenderParams *p, float *film, unsigned long long *counters) {
    HsScene *s = (HsScene *) h;
    HsRenderOut out;
    if (p->integrator == B2_INTEGRATOR_PPM) hs_render_ppm(s->view, *cam, *p, film, out); else hs_render_pt(s->view, *cam, *p, film, out);
    if (counters) memcpy(counters, out.counters, sizeof(out.counters));
}
// B2RenderParams.strategies 1..4 (b2_strategies.cuh, b2_mis.cuh)
void hs_render_strategies_c(void *h, const B2Camera *cam, const B2RenderParams *p, float *film, unsigned long long *counters) {
    HsScene *s = (HsScene *) h;
    HsRenderOut out;
    hs_render_strategies(s->view, *cam, *p, film, out);
    if (counters) memcpy(counters, out.counters, sizeof(out.counters));
}
void hs_render_bdpt(void *h, const B2Camera *cam, const B2RenderParams *p, float *film, unsigned long long *counters) {
    HsScene *s = (HsScene *) h;
    HsRenderOut out;
    hs_render(s->view, *cam, *p, film, out);
    if (counters) { memcpy(counters, out.counters, sizeof(out.counters)); counters[B2_CNT_N] = out.pretest_mismatch; }
}
// the same render with the MIS walks hoisted out of the pair loop (k_pool_active + k_connect's form of the weight)
void hs_render_bdpt_factored(void *h, const B2Camera *cam, const B2RenderParams *p, float *film, unsigned long long *counters) {
    HsScene *s = (HsScene *) h;
    HsRenderOut out;
    hs_render(s->view, *cam, *p, film, out, true);
    if (counters) { memcpy(counters, out.counters, sizeof(out.counters)); counters[B2_CNT_N] = out.pretest_mismatch; }
}
// mis_weight_core over flat vertex arrays (the layout of b2_mis_weight: eye 17 floats, light 20 floats per vertex, V vertices per
// chain); factored != 0: both walks through mis_eye_sum / mis_light_sum, as the connection stage evaluates the weight
namespace {
struct HsFlatEye {
    const float *v; int s;
    float3 pos(int i) const { return make_float3(v[17 * i], v[17 * i + 1], v[17 * i + 2]); }
    float eta(int i) const { return v[17 * i + 14]; }
    float pdf_fwd(int i) const { return v[17 * i + 15]; }
    float pdf_rev(int i) const { return v[17 * i + 16]; }
    float3 cur_n() const { return make_float3(v[17 * s + 3], v[17 * s + 4], v[17 * s + 5]); }
    Mat cur_mtl() const { const float *m = v + 17 * s + 9; Mat r; r.base = make_float3(m[0], m[1], m[2]); r.roughness = m[3]; r.metallic = m[4]; r.eta = m[5]; return r; }
};
struct HsFlatLight {
    const float *v; int t;
    float3 pos(int i) const { return make_float3(v[20 * i], v[20 * i + 1], v[20 * i + 2]); }
    float eta(int i) const { return v[20 * i + 14]; }
    bool is_src(int i) const { return v[20 * i + 15] != 0.0f; }
    float pdf_fwd(int i) const { return v[20 * i + 18]; }
    float pdf_rev(int i) const { return v[20 * i + 19]; }
    float3 cur_n() const { return make_float3(v[20 * t + 3], v[20 * t + 4], v[20 * t + 5]); }
    Mat cur_mtl() const { const float *m = v + 20 * t + 9; Mat r; r.base = make_float3(m[0], m[1], m[2]); r.roughness = m[3]; r.metallic = m[4]; r.eta = m[5]; return r; }
};
}
void hs_mis_weight(size_t n, int V, const float *eye17, const int32_t *s_idx, const float *lit20, const int32_t *t_idx, const float *d_vec,
                   const float *dist2, const float *cam_pos, int factored, float *out) {
    for (size_t i = 0; i < n; ++i) {
        const int s = s_idx[i], t = t_idx[i];
        if (s < 0 || t < 0 || s >= V || t >= V) { out[i] = 0.0f; continue; }
        const HsFlatEye E{ eye17 + (size_t) 17 * V * i, s };
        const HsFlatLight L{ lit20 + (size_t) 20 * V * i, t };
        const float3 d = make_float3(d_vec[3 * i], d_vec[3 * i + 1], d_vec[3 * i + 2]), cp = make_float3(cam_pos[3 * i], cam_pos[3 * i + 1], cam_pos[3 * i + 2]);
        out[i] = factored ? mis_weight_core(E, s, L, t, d, dist2[i], cp, mis_eye_sum(E, s), mis_light_sum(L, t))
                          : mis_weight_core(E, s, L, t, d, dist2[i], cp);
    }
}
// pool20 layout as b2_light_pool / or_light_pool
void hs_light_pool(void *h, const B2RenderParams *p, float *pool20) {
    HsScene *s = (HsScene *) h;
    B2Camera cam; memset(&cam, 0, sizeof(cam)); cam.look_at[2] = 1; cam.view_up[1] = 1; cam.fov = 50;
    B2RenderParams pp = *p; if (pp.width <= 0) pp.width = pp.height = 1;
    RenderCtx c; hs_setup_ctx(c, s->view, cam, pp);
    std::vector<float4> pool, pre; unsigned long long cnt[B2_CNT_N] = { 0 };
    hs_build_pool(c, pool, pre, cnt);
    for (int i = 0; i < c.n_pool; ++i) {
        LightVtx v = load_lv(c.pool, i);
        float *o = pool20 + 20 * i;
        o[0] = v.pos.x; o[1] = v.pos.y; o[2] = v.pos.z; o[3] = v.n.x; o[4] = v.n.y; o[5] = v.n.z; o[6] = v.thr.x; o[7] = v.thr.y; o[8] = v.thr.z;
        o[9] = v.mtl.base.x; o[10] = v.mtl.base.y; o[11] = v.mtl.base.z; o[12] = v.mtl.roughness; o[13] = v.mtl.metallic; o[14] = v.mtl.eta;
        o[15] = (v.flags & B2_LV_SRC) ? 1.f : 0.f; o[16] = (v.flags & B2_LV_PARALLEL) ? 1.f : 0.f; o[17] = v.cutoff; o[18] = v.pdf_fwd; o[19] = v.pdf_rev;
    }
}
}
