// examples/bdpt_b200_helper.cpp -- the reference-side binding of INTEGRATION.md section 1 as a real translation unit.
// It REPLACES the reference's src/bdpt_cu_helper.cpp: the same two entry points main_cli.cpp:200-213 calls,
//     void move_data_to_cuda_bdpt(std::map<int, AABB> groups, std::vector<CudaLight> &lights, int spl)   include/bdpt_cu_helper.h:5
//     void run_cuda_bdpt(CudaCamera cam, float3 *image, int light_depth, int eye_depth, int W, int H, int spp, int spl)      :6
// implemented over the C ABI of libb2pt.so (include/b2pt.h).  It is compiled against the reference's OWN headers
// (include/bdpt_cu_helper.h -> bdpt_cu.cuh -> geometric.cuh -> object.h, from where they lie) by oracle/Makefile into
// oracle/_ref/libbinding_b200.so and exercised by tests/test_gpu_binding.py.  object.h has no include guard: it must come
// in exactly once, through bdpt_cu_helper.h.
#include "bdpt_cu_helper.h"
#include "b2pt.h"
#include <stdexcept>

namespace {
B2Scene *g_scene = nullptr;
B2Camera g_cam;
int g_spl = 1;
void put(float *d, const glm::vec3 &v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }
void put(float *d, const float3 &v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }
}  // namespace

void move_data_to_cuda_bdpt(std::map<int, AABB> groups, std::vector<CudaLight> &lights, int spl) {
    std::vector<B2Tri> tris; std::vector<B2Sphere> sph; std::vector<B2Material> mats; std::vector<B2Light> ls;
    for (auto &g : groups)                                   // std::map key order = the CPU visiting order (cpu_bdpt.cpp:43)
        for (Object *o : g.second.objs) {
            B2Material m; put(m.base_color, o->mtl.base_color); m.roughness = o->mtl.roughness; m.metallic = o->mtl.metallic; m.eta = o->mtl.eta;
            mats.push_back(m);
            if (auto *t = dynamic_cast<Triangle *>(o)) {
                B2Tri r; put(r.v0, t->vert[0]); put(r.v1, t->vert[1]); put(r.v2, t->vert[2]);
                r.obj_id = o->obj_id; r.group = g.first; r.material = (int) mats.size() - 1; tris.push_back(r);
            } else if (auto *s = dynamic_cast<Sphere *>(o)) {
                B2Sphere r; put(r.center, s->center); r.r = s->r;
                r.obj_id = o->obj_id; r.group = g.first; r.material = (int) mats.size() - 1; sph.push_back(r);
            }
        }
    for (auto &L : lights) {
        B2Light r; put(r.pos, L.pos); put(r.dir, L.dir); put(r.illum, L.illum);
        r.cutoff = L.cutoff; r.is_parallel = L.is_parallel; r.ball_r = L.light_ball.r; ls.push_back(r);
    }
    g_spl = spl;
    if (g_scene) { b2_scene_destroy(g_scene); g_scene = nullptr; }   // the reference never clears its globals (bdpt_cu_helper.cpp:3-10); we do
    if (b2_scene_create(tris.data(), tris.size(), sph.data(), sph.size(), mats.data(), mats.size(), ls.data(), ls.size(), 0, &g_scene) ||
        b2_scene_build_bvh(g_scene, nullptr))
        throw std::runtime_error(b2_last_error());
}

// main_cli.cpp hands over CudaCamera{eye,UL,dx,dy} built with a hard-coded F = 50 (main_cli.cpp:158-166); the CPU estimator
// this library reproduces derives its basis from Camera{eye,look_at,view_up,fov} (cpu_bdpt.cpp:190-200), which is what
// B2Camera carries: one added call next to init_camera.
void set_bdpt_camera(const Camera &c) { put(g_cam.eye, c.eye); put(g_cam.look_at, c.look_at); put(g_cam.view_up, c.view_up); g_cam.fov = c.fov; }

void run_cuda_bdpt(CudaCamera, float3 *image_buffer, int light_depth, int eye_depth, int W, int H, int spp, int spl) {
    if (!g_scene) throw std::runtime_error("run_cuda_bdpt before move_data_to_cuda_bdpt");
    B2RenderParams p{}; p.width = W; p.height = H; p.spp = spp; p.spl = spl; p.eye_depth = eye_depth; p.light_depth = light_depth;
    p.seed = 1; p.sample_stride = 1; p.film_mode = B2_FILM_MEAN;
    B2Stats st;
    if (b2_render_bdpt(g_scene, &g_cam, &p, reinterpret_cast<float *>(image_buffer), &st))   // float3[W*H] == W*H*3 floats, row 0 = top
        throw std::runtime_error(b2_last_error());
}

void release_cuda_bdpt() { if (g_scene) { b2_scene_destroy(g_scene); g_scene = nullptr; } }
