// oracle/ref_cuda_shim.cpp -- TEST / BASELINE INFRASTRUCTURE ONLY (never linked into the product).
//
// extern "C" face over the reference's EXISTING CUDA integrators, compiled unmodified for sm_100a from where they lie
// under /root/reference by oracle/Makefile into oracle/_ref/libref_cuda.so:
//     src/bdpt_cu.cu  src/pt_cu.cu  src/ppm_cu.cu  (nvcc)   src/*_cu_helper.cpp  src/object.cpp  src/geometric.cu
// Nothing from those files is copied here.  Purpose ("(ii) bdpt_cu.cu compiled for sm_100a as an honesty
// number"): time the reference's own GPU kernels on the same B200 beside this repo's path, and give the PT / PPM rows of
// SURVEY.md 8(f) a reference that runs on the GPU box.  It is NOT the parity oracle of the BDPT path (the
// CUDA BDPT divides the light power by spl twice, builds lights*spl^2 light paths, seeds its RNG from the clock).
//
// The shim builds std::map<int,AABB> groups the way main_cli.cpp:104-122 does, calls move_data_to_cuda_* /
// run_cuda_* (main_cli.cpp:201-213) and empties the helpers' process-global scene vectors first (bdpt/pt never clear
// them: bdpt_cu_helper.cpp:3-10).
#include "bdpt_cu_helper.h"   // -> bdpt_cu.cuh -> geometric.cuh -> object.h (no include guards: include exactly once)
#include <chrono>
#include <cstring>
#include <cstdio>
#include <unistd.h>
#include <fcntl.h>

void move_data_to_cuda_pt(std::map<int, AABB> groups, std::vector<CudaLight> &cuda_lights, int);
void run_cuda_pt(CudaCamera cam, float3 *image_buffer, int light_depth, int eye_depth, int W, int H, int spp);
void move_data_to_cuda_ppm(std::map<int, AABB> groups, std::vector<CudaLight> &cuda_lights, int);
void run_cuda_ppm(CudaCamera cam, float3 *image_buffer, int light_depth, int eye_depth, int W, int H, int spp);

#define REF_NS(ns)                                                                                         \
    namespace ns {                                                                                         \
    extern std::vector<CudaSphere> cuda_spheres; extern std::vector<CudaTriangle> cuda_triangles;          \
    extern std::vector<CudaLight> cuda_lights; extern float3 scene_max_bound, scene_min_bound; }
REF_NS(bdpt_ns) REF_NS(pt_ns) REF_NS(ppm_ns)

namespace {
struct RefCudaScene {
    std::map<int, AABB> groups;
    std::vector<CudaLight> lights;
    std::vector<Object *> owned;
};
inline float3 f3(const float *p) { return make_float3(p[0], p[1], p[2]); }
template <class NS_S, class NS_T, class NS_L>
void reset_ns(NS_S &s, NS_T &t, NS_L &l, float3 &mx, float3 &mn, size_t light_cap) {
    s.clear(); t.clear(); l.clear();
    l.reserve(light_cap);          // bdpt_cu.cu:555-556 copies num_lights*spl records from this vector: keep the over-read inside its capacity
    mx = make_float3(-1e9f, -1e9f, -1e9f); mn = make_float3(1e9f, 1e9f, 1e9f);
}
// The reference prints progress to stdout from inside its wrappers; keep the caller's stdout clean.
struct QuietStdout {
    int saved = -1;
    explicit QuietStdout(bool on) {
        if (!on) return;
        fflush(stdout);
        saved = dup(1);
        int nul = open("/dev/null", O_WRONLY);
        if (nul >= 0) { dup2(nul, 1); close(nul); }
    }
    ~QuietStdout() { if (saved >= 0) { fflush(stdout); std::cout.flush(); dup2(saved, 1); close(saved); } }
};
// move_data_to_cuda_{bdpt,pt,ppm} build every CudaSphere / CudaTriangle in an UNINITIALISED stack local and never assign its
// mtl_old (bdpt_cu_helper.cpp:26,42, pt_cu_helper.cpp:25,41: the line is commented out), yet check_visibility decides
// opaque / transparent on mtl_old.refract and tints by mtl_old.Ks (geometric.cuh:309-320): the reference's shadow rays
// depend on whatever the caller left on the stack (seen here as a blue channel 3.5x too bright in one process and
// correct in the next).  The shim zeroes the stack region the helper's frame will occupy, so the undefined read resolves
// to "zero-initialised" = every object opaque -- the reading this repo's pt / ppm integrators and oracle document.
__attribute__((noinline)) void scrub_stack() {
    volatile unsigned char buf[1 << 16];
    for (size_t i = 0; i < sizeof(buf); ++i) buf[i] = 0;
}
}  // namespace

extern "C" {

// Same argument layout as ref_scene_create in ref_shim.cpp (kind 0 = sphere, 1 = triangle; 6 material floats per
// object; 12 floats per light with the cutoff already in radians).
void *refcuda_scene_create(int n_obj, const int *kind, const int *group, const float *geom, const float *mtl,
                           int n_lights, const float *lights) {
    RefCudaScene *s = new RefCudaScene;
    for (int i = 0; i < n_obj; ++i) {
        const float *g = geom + 9 * i, *m = mtl + 6 * i;
        Material mm;
        mm.base_color = glm::vec3(m[0], m[1], m[2]);
        mm.roughness = m[3]; mm.metallic = m[4]; mm.eta = m[5];
        Object *o;
        if (kind[i] == 0) {
            Sphere *sp = new Sphere;
            sp->center = glm::vec3(g[0], g[1], g[2]); sp->r = g[3]; sp->scale = glm::vec3(1, 1, 1);
            o = sp;
        } else {
            Triangle *t = new Triangle;
            for (int k = 0; k < 3; ++k) t->vert[k] = glm::vec3(g[3 * k], g[3 * k + 1], g[3 * k + 2]);
            o = t;
        }
        o->mtl = mm; o->obj_id = i;
        s->groups[group[i]].add_obj(o);
        s->owned.push_back(o);
    }
    for (int i = 0; i < n_lights; ++i) {
        const float *l = lights + 12 * i;
        CudaLight L;
        memset(&L, 0, sizeof(L));
        L.pos = f3(l); L.dir = f3(l + 3); L.illum = f3(l + 6);
        L.cutoff = l[9]; L.is_parallel = (int) l[10];
        L.light_ball.r = l[11]; L.light_ball.center = L.pos; L.light_ball.mtl_old.Kd = L.illum;
        s->lights.push_back(L);
    }
    return s;
}

void refcuda_scene_destroy(void *h) {
    RefCudaScene *s = (RefCudaScene *) h;
    for (Object *o : s->owned) delete o;
    delete s;
}

// mode: 0 = bdpt (run_cuda_bdpt), 1 = pt (run_cuda_pt), 2 = ppm (run_cuda_ppm).  cam: eye3 UL3 dx3 dy3 as main_cli.cpp:162-166
// fills them.  Depth arguments are passed in main_cli.cpp:211-213's order (EYE_DEPTH first).  film: W*H*3 floats, mean
// radiance as the wrapper leaves it.  Returns the wall time of run_cuda_* in milliseconds (what main_cli.cpp:207-219
// reports: device allocation, RNG initialisation, kernels and the read-back), or < 0 on a bad mode.
double refcuda_render(void *h, int mode, const float *cam12, int W, int H, int spp, int spl, int eye_depth, int light_depth,
                      float *film, int quiet) {
    RefCudaScene *s = (RefCudaScene *) h;
    CudaCamera cam = {};
    cam.eye = f3(cam12); cam.UL = f3(cam12 + 3); cam.dx = f3(cam12 + 6); cam.dy = f3(cam12 + 9);
    std::vector<float3> img((size_t) W * H, make_float3(0, 0, 0));
    const size_t cap = s->lights.size() * (size_t) (spl > 0 ? spl : 1) + 1;
    QuietStdout q(quiet != 0);
    double ms = -1.0;
    if (mode == 0) {
        reset_ns(bdpt_ns::cuda_spheres, bdpt_ns::cuda_triangles, bdpt_ns::cuda_lights, bdpt_ns::scene_max_bound, bdpt_ns::scene_min_bound, cap);
        std::map<int, AABB> groups_copy = s->groups;      // the by-value argument, built before the scrub
        scrub_stack();
        move_data_to_cuda_bdpt(std::move(groups_copy), s->lights, spl);
        const auto t0 = std::chrono::steady_clock::now();
        run_cuda_bdpt(cam, img.data(), eye_depth, light_depth, W, H, spp, spl);
        ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    } else if (mode == 1) {
        reset_ns(pt_ns::cuda_spheres, pt_ns::cuda_triangles, pt_ns::cuda_lights, pt_ns::scene_max_bound, pt_ns::scene_min_bound, cap);
        std::map<int, AABB> groups_copy = s->groups;      // the by-value argument, built before the scrub
        scrub_stack();
        move_data_to_cuda_pt(std::move(groups_copy), s->lights, spl);
        const auto t0 = std::chrono::steady_clock::now();
        run_cuda_pt(cam, img.data(), eye_depth, light_depth, W, H, spp);
        ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    } else if (mode == 2) {
        reset_ns(ppm_ns::cuda_spheres, ppm_ns::cuda_triangles, ppm_ns::cuda_lights, ppm_ns::scene_max_bound, ppm_ns::scene_min_bound, cap);
        std::map<int, AABB> groups_copy = s->groups;      // the by-value argument, built before the scrub
        scrub_stack();
        move_data_to_cuda_ppm(std::move(groups_copy), s->lights, spl);
        const auto t0 = std::chrono::steady_clock::now();
        run_cuda_ppm(cam, img.data(), eye_depth, light_depth, W, H, spp);
        ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    for (size_t i = 0; i < img.size(); ++i) { film[3 * i] = img[i].x; film[3 * i + 1] = img[i].y; film[3 * i + 2] = img[i].z; }
    return ms;
}

}  // extern "C"
