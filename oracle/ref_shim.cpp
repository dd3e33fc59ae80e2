// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// A thin extern "C" face over the UNMODIFIED reference sources, which are compiled
// where they lie under /root/reference by oracle/Makefile into oracle/_ref/libref_bdpt.so:
//     src/cpu_bdpt.cpp  src/object.cpp  src/geometric.cu (as host C++)
// Nothing from those files is copied here.  The shim only (1) builds the
// std::map<int,AABB> groups exactly the way main_cli.cpp:104-122 does (new object,
// set mtl / obj_id, groups[gid].add_obj), (2) calls the reference entry points, and
// (3) recovers the hit object id (CudaHit carries none) by walking the same
// group/object order through the reference's own virtual check_intersect/intersectAABB
// (cpu_bdpt.cpp:43-61), cross-checked against cpu_find_closest_hit's returned t.
#include "cpu_bdpt.h"   // pulls bdpt_cu.cuh -> geometric.cuh -> object.h -> ray.h (none has include guards)
#include <omp.h>
#include <cstring>
#include <cstdint>

// Non-static helpers defined in the reference's cpu_bdpt.cpp (:30, :82, :112).
CudaHit cpu_find_closest_hit(float3 ray_point, float3 ray_dir, std::map<int, AABB> &groups,
                             const std::vector<CudaLight> &lights);
float3 cpu_check_visibility(float3 p1, float3 p2, std::map<int, AABB> &groups);
float cpu_calculate_mis_weight(const CudaEyeVertex *eye_path, int s_idx, const CudaLightVertex *light_path,
                               int t_idx, const float3 &dir_e_to_l, float dist2, const float3 &camera_pos);

namespace {
struct RefScene {
    std::map<int, AABB> groups;
    std::vector<CudaLight> lights;
    std::vector<Object *> owned;
};
inline float3 f3(const float *p) { return make_float3(p[0], p[1], p[2]); }
inline void st3(float *p, float3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }
inline CudaMaterial mk_mtl(const float *m) {
    CudaMaterial c;
    memset(&c, 0, sizeof(c));
    c.base_color = f3(m); c.roughness = m[3]; c.metallic = m[4]; c.eta = m[5];
    return c;
}
}  // namespace

extern "C" {

// kind[i]: 0 = sphere (geom: cx cy cz r, rest ignored), 1 = triangle (geom: 9 floats).
// mtl: 6 floats per object (r g b roughness metallic eta).  lights: 12 floats each
// (pos3 dir3 illum3 cutoff_RADIANS is_parallel ball_r), as main_cli.cpp:131-139 leaves them.
void *ref_scene_create(int n_obj, const int *kind, const int *group, const float *geom, const float *mtl,
                       int n_lights, const float *lights) {
    RefScene *s = new RefScene;
    for (int i = 0; i < n_obj; ++i) {
        const float *g = geom + 9 * i;
        const float *m = mtl + 6 * i;
        Material mm;
        mm.base_color = glm::vec3(m[0], m[1], m[2]);
        mm.roughness = m[3]; mm.metallic = m[4]; mm.eta = m[5];
        Object *o;
        if (kind[i] == 0) {
            Sphere *sp = new Sphere;
            sp->center = glm::vec3(g[0], g[1], g[2]);
            sp->r = g[3];
            sp->scale = glm::vec3(1, 1, 1);
            o = sp;
        } else {
            Triangle *t = new Triangle;
            for (int k = 0; k < 3; ++k) t->vert[k] = glm::vec3(g[3 * k], g[3 * k + 1], g[3 * k + 2]);
            o = t;
        }
        o->mtl = mm;
        o->obj_id = i;
        s->groups[group[i]].add_obj(o);
        s->owned.push_back(o);
    }
    for (int i = 0; i < n_lights; ++i) {
        const float *l = lights + 12 * i;
        CudaLight L;
        memset(&L, 0, sizeof(L));
        L.pos = f3(l); L.dir = f3(l + 3); L.illum = f3(l + 6);
        L.cutoff = l[9];
        L.is_parallel = (int) l[10];
        L.light_ball.r = l[11];
        L.light_ball.center = L.pos;
        L.light_ball.mtl_old.Kd = L.illum;
        s->lights.push_back(L);
    }
    return s;
}

void ref_scene_destroy(void *h) {
    RefScene *s = (RefScene *) h;
    for (Object *o : s->owned) delete o;
    delete s;
}

void ref_set_threads(int n) { omp_set_num_threads(n); }
int ref_max_threads(void) { return omp_get_max_threads(); }

// Group bounds as add_obj grew them (object.cpp:123-147); out: per group key order, 7 floats (key, min3, max3).
int ref_group_bounds(void *h, float *out, int cap) {
    RefScene *s = (RefScene *) h;
    int n = 0;
    for (auto &g : s->groups) {
        if (n < cap) {
            float *o = out + 7 * n;
            o[0] = (float) g.first;
            o[1] = g.second.min.x; o[2] = g.second.min.y; o[3] = g.second.min.z;
            o[4] = g.second.max.x; o[5] = g.second.max.y; o[6] = g.second.max.z;
        }
        ++n;
    }
    return n;
}

// Closest hit for n rays.  id_out: obj_id, or -2-k for light ball k, or -1 for a miss.
// Returns the number of rays for which the id-recovery walk disagreed with cpu_find_closest_hit (must be 0).
int ref_closest_hit(void *h, int n, const float *org, const float *dir, int32_t *id_out, float *t_out,
                    float *pos_out, float *nrm_out, float *mtl_out) {
    RefScene *s = (RefScene *) h;
    int bad = 0;
    for (int i = 0; i < n; ++i) {
        float3 o = f3(org + 3 * i), d = f3(dir + 3 * i);
        CudaHit hit = cpu_find_closest_hit(o, d, s->groups, s->lights);
        // id recovery: same loop order and ranges as cpu_bdpt.cpp:43-78, through the reference's own virtuals
        int id = -1;
        float best = 1e20f;
        Ray ray;
        ray.point = glm::vec3(o.x, o.y, o.z);
        ray.vec = glm::vec3(d.x, d.y, d.z);
        for (auto &g : s->groups) {
            if (!g.second.intersectAABB(ray, 1e-4f, best)) continue;
            for (Object *obj : g.second.objs) {
                float t, u, v;
                if (obj->check_intersect(ray, t, u, v, 1e-4f, best)) { best = t; id = obj->obj_id; }
            }
        }
        for (size_t k = 0; k < s->lights.size(); ++k) {
            float t;
            if (intersect_sphere(o, d, s->lights[k].light_ball, t, best)) { best = t; id = -2 - (int) k; }
        }
        if ((id != -1) != hit.hit || (hit.hit && best != hit.t) || (hit.hit && hit.is_light != (id <= -2))) ++bad;
        id_out[i] = id;
        t_out[i] = hit.hit ? hit.t : 1e20f;
        if (pos_out) st3(pos_out + 3 * i, hit.hit ? hit.pos : make_float3(0, 0, 0));
        if (nrm_out) st3(nrm_out + 3 * i, hit.hit ? hit.normal : make_float3(0, 0, 0));
        if (mtl_out) {
            float *m = mtl_out + 6 * i;
            if (hit.hit) { st3(m, hit.mtl.base_color); m[3] = hit.mtl.roughness; m[4] = hit.mtl.metallic; m[5] = hit.mtl.eta; }
            else memset(m, 0, 6 * sizeof(float));
        }
    }
    return bad;
}

void ref_visibility(void *h, int n, const float *p1, const float *p2, float *vis_out) {
    RefScene *s = (RefScene *) h;
    for (int i = 0; i < n; ++i) {
        float3 t = cpu_check_visibility(f3(p1 + 3 * i), f3(p2 + 3 * i), s->groups);
        vis_out[i] = t.x;
    }
}

// cam: eye3 look_at3 view_up3 fov_deg.  image: W*H*3 floats, row 0 = top (cpu_bdpt.cpp:345-346,479).
void ref_render(void *h, const float *cam, int eye_depth, int light_depth, int W, int H, int spp, int spl,
                float *image) {
    RefScene *s = (RefScene *) h;
    Camera c;
    c.eye = glm::vec3(cam[0], cam[1], cam[2]);
    c.look_at = glm::vec3(cam[3], cam[4], cam[5]);
    c.view_up = glm::vec3(cam[6], cam[7], cam[8]);
    c.fov = cam[9];
    std::vector<float3> img((size_t) W * H, make_float3(0, 0, 0));
    run_cpu_bdpt(c, s->groups, s->lights, img.data(), eye_depth, light_depth, W, H, spp, spl);
    memcpy(image, img.data(), sizeof(float3) * img.size());
}

void ref_bsdf_evaluate(int n, const float *mtl, const float *wo, const float *wi, const float *N, float *out) {
    for (int i = 0; i < n; ++i)
        st3(out + 3 * i, bsdf_evaluate(mk_mtl(mtl + 6 * i), f3(wo + 3 * i), f3(wi + 3 * i), f3(N + 3 * i)));
}
void ref_bsdf_pdf(int n, const float *mtl, const float *wo, const float *wi, const float *N, float *out) {
    for (int i = 0; i < n; ++i)
        out[i] = bsdf_pdf(mk_mtl(mtl + 6 * i), f3(wo + 3 * i), f3(wi + 3 * i), f3(N + 3 * i));
}
// u: 3 floats per call (u_rr,u1,u2).  Outputs are pre-zeroed so untouched reference outputs are defined.
void ref_bsdf_sample(int n, const float *mtl, const float *wo, const float *N, const float *u, const float *cur_eta,
                     float *wi_out, float *f_out, float *pdf_out, int32_t *delta_out, float *eta_out) {
    for (int i = 0; i < n; ++i) {
        float3 wi = make_float3(0, 0, 0), f = make_float3(0, 0, 0);
        float pdf = 0.0f, ne = 0.0f;
        bool isd = false;
        bsdf_sample(mk_mtl(mtl + 6 * i), f3(wo + 3 * i), f3(N + 3 * i), u[3 * i], u[3 * i + 1], u[3 * i + 2],
                    cur_eta[i], wi, f, pdf, isd, ne);
        st3(wi_out + 3 * i, wi); st3(f_out + 3 * i, f);
        pdf_out[i] = pdf; delta_out[i] = isd ? 1 : 0; eta_out[i] = ne;
    }
}

// Flat vertex layouts (floats):
//   eye  vertex (17): pos3 normal3 thr3 mtl6 pdf_fwd pdf_rev
//   light vertex(20): pos3 normal3 thr3 mtl6 is_light_source is_parallel source_cutoff pdf_fwd pdf_rev
float ref_mis_weight(const float *eye, int n_eye, int s_idx, const float *lit, int n_lit, int t_idx,
                     const float *d_vec, float dist2, const float *cam_pos) {
    std::vector<CudaEyeVertex> E(n_eye);
    std::vector<CudaLightVertex> L(n_lit);
    for (int i = 0; i < n_eye; ++i) {
        const float *p = eye + 17 * i;
        memset(&E[i], 0, sizeof(CudaEyeVertex));
        E[i].pos = f3(p); E[i].normal = f3(p + 3); E[i].throughput = f3(p + 6); E[i].mtl = mk_mtl(p + 9);
        E[i].pdf_fwd = p[15]; E[i].pdf_rev = p[16];
    }
    for (int i = 0; i < n_lit; ++i) {
        const float *p = lit + 20 * i;
        memset(&L[i], 0, sizeof(CudaLightVertex));
        L[i].pos = f3(p); L[i].normal = f3(p + 3); L[i].throughput = f3(p + 6); L[i].mtl = mk_mtl(p + 9);
        L[i].is_light_source = p[15] != 0.0f; L[i].is_parallel = p[16] != 0.0f; L[i].source_cutoff = p[17];
        L[i].pdf_fwd = p[18]; L[i].pdf_rev = p[19];
    }
    float3 d = f3(d_vec), c = f3(cam_pos);
    return cpu_calculate_mis_weight(E.data(), s_idx, L.data(), t_idx, d, dist2, c);
}

}  // extern "C"
