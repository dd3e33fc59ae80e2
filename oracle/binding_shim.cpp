// oracle/binding_shim.cpp -- TEST INFRASTRUCTURE ONLY.  Drives examples/bdpt_b200_helper.cpp (the reference-side binding over
// libb2pt.so) the way the reference's main_cli does: it builds the std::map<int,AABB> groups exactly like main_cli.cpp:104-122
// (new object, set mtl / obj_id, groups[gid].add_obj -- the reference's own object.cpp, compiled from where it lies), fills
// CudaLight like main_cli.cpp:131-139, then calls move_data_to_cuda_bdpt / run_cuda_bdpt (main_cli.cpp:201-213).
// tests/test_gpu_binding.py compares the film with the one api.Scene.render_bdpt returns for the same scene.
#include "bdpt_cu_helper.h"
#include <cstring>
#include <stdexcept>
#include <string>

void set_bdpt_camera(const Camera &c);
void release_cuda_bdpt();

static std::string g_err;

extern "C" const char *binding_last_error() { return g_err.c_str(); }

// kind 0 = sphere, 1 = triangle; 6 material floats per object; 12 floats per light (cutoff in radians); cam = eye3 look3 up3 fov
extern "C" int binding_render(int n_obj, const int *kind, const int *group, const float *geom, const float *mtl, int n_lights, const float *lights,
                              const float *cam10, int eye_depth, int light_depth, int W, int H, int spp, int spl, float *image) {
    std::map<int, AABB> groups;
    std::vector<Object *> owned;
    std::vector<CudaLight> cu_light;
    try {
        for (int i = 0; i < n_obj; ++i) {
            const float *g = geom + 9 * i, *m = mtl + 6 * i;
            Material mm; mm.base_color = glm::vec3(m[0], m[1], m[2]); mm.roughness = m[3]; mm.metallic = m[4]; mm.eta = m[5];
            Object *o;
            if (kind[i] == 0) { Sphere *sp = new Sphere; sp->center = glm::vec3(g[0], g[1], g[2]); sp->r = g[3]; sp->scale = glm::vec3(1, 1, 1); o = sp; }
            else { Triangle *t = new Triangle; for (int k = 0; k < 3; ++k) t->vert[k] = glm::vec3(g[3 * k], g[3 * k + 1], g[3 * k + 2]); o = t; }
            o->mtl = mm; o->obj_id = i;
            groups[group[i]].add_obj(o);
            owned.push_back(o);
        }
        for (int i = 0; i < n_lights; ++i) {
            const float *l = lights + 12 * i;
            CudaLight L; memset(&L, 0, sizeof(L));
            L.pos = make_float3(l[0], l[1], l[2]); L.dir = make_float3(l[3], l[4], l[5]); L.illum = make_float3(l[6], l[7], l[8]);
            L.cutoff = l[9]; L.is_parallel = (int) l[10]; L.light_ball.r = l[11]; L.light_ball.center = L.pos;
            cu_light.push_back(L);
        }
        Camera c; c.eye = glm::vec3(cam10[0], cam10[1], cam10[2]); c.look_at = glm::vec3(cam10[3], cam10[4], cam10[5]);
        c.view_up = glm::vec3(cam10[6], cam10[7], cam10[8]); c.fov = cam10[9];
        move_data_to_cuda_bdpt(groups, cu_light, spl);                                     // main_cli.cpp:201
        set_bdpt_camera(c);
        CudaCamera cu_cam; memset(&cu_cam, 0, sizeof(cu_cam));
        std::vector<float3> frame((size_t) W * H);
        run_cuda_bdpt(cu_cam, frame.data(), light_depth, eye_depth, W, H, spp, spl);        // main_cli.cpp:211 (declared order: light_depth, eye_depth)
        memcpy(image, frame.data(), sizeof(float3) * frame.size());
        release_cuda_bdpt();
    } catch (const std::exception &e) {
        g_err = e.what();
        for (Object *o : owned) delete o;
        return -1;
    }
    for (Object *o : owned) delete o;
    return 0;
}
