/* oracle/bdpt_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of the reference's bidirectional path-tracing sample loop
 * (/root/reference/src/cpu_bdpt.cpp, src/object.cpp, include/geometric.cuh).  It is the
 * checker for the CUDA path: only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it.  The product (libb2pt.so) never does.
 *
 * Parity status: PINNED.  The reference ships no tests or golden vectors,
 * so the restatement is pinned against the reference itself, compiled unmodified into
 * oracle/_ref/libref_bdpt.so: bit-identical hits / visibility / BSDF / MIS tables and a
 * bit-identical single-thread film (tests/test_oracle_pin.py, tests/golden/).
 * One boundary stays unpinned: glm (un-vendored, un-pinned dependency of the reference)
 * is replaced by oracle/compat/glm/glm.hpp for BOTH builds.
 */
#ifndef BDPT_ORACLE_H
#define BDPT_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct OrScene OrScene;

/* rng_mode for or_render / or_light_pool */
#define OR_RNG_MT19937 0 /* std::mt19937(1337+tid) / (9999+tid) + uniform_real_distribution<float>, cpu_bdpt.cpp:213,338 */
#define OR_RNG_PHILOX  1 /* counter-based Philox4x32-10 keyed by (seed; kind, path ids, draw block) -- the GPU's streams */

typedef struct OrRenderOpts {
    int      rng_mode;
    uint64_t seed;        /* Philox key (ignored for mt19937) */
    int      nthreads;    /* <=0: OpenMP default */
    int      max_delta;   /* cap on consecutive-free delta bounces per subpath; <=0: unbounded like the reference */
    int      sample_begin;/* Philox mode only: render samples s in [sample_begin, sample_begin+spp) of the global   */
                          /* sample index space (mt19937 mode ignores it); the film is still divided by spp.        */
} OrRenderOpts;

typedef struct OrStats {
    uint64_t eye_samples;    /* (pixel, s) iterations, cpu_bdpt.cpp:349 */
    uint64_t closest_rays;   /* cpu_find_closest_hit calls (eye + light subpaths) */
    uint64_t shadow_rays;    /* cpu_check_visibility calls, :424 */
    uint64_t path_samples;   /* connections reaching the MIS evaluation (:431) + depth-0 emitter hits (:372) */
    uint64_t delta_cap_hits; /* subpaths stopped by max_delta */
    double   seconds;        /* pool + eye phases, the interval the reference times at :210 -> :485 */
} OrStats;

/* Scene construction mirrors main_cli.cpp:104-139 (objects in file order, obj_id = index).
 * kind[i]: 0 sphere (geom: cx cy cz r), 1 triangle (geom: 9 floats); geom stride 9; mtl stride 6
 * (r g b roughness metallic eta); lights stride 12 (pos3 dir3 illum3 cutoff_RADIANS is_parallel ball_r). */
OrScene *or_scene_create(int n_obj, const int *kind, const int *group, const float *geom, const float *mtl,
                         int n_lights, const float *lights);
void or_scene_destroy(OrScene *s);
int or_group_bounds(OrScene *s, float *out7, int cap);

/* Scene grammar parser, main_cli.cpp:99-141.  Returns a malloc'ed scene description or NULL. */
typedef struct OrParsed {
    int    n_obj, n_lights;
    int   *kind, *group;
    float *geom, *mtl, *lights;
    float  cam[10]; /* eye3 look_at3 view_up3 fov */
    int    W, H;
} OrParsed;
OrParsed *or_parse_scene_file(const char *path);
void or_parsed_free(OrParsed *p);

/* id_out: obj_id, -2-k for light ball k, -1 miss; t_out = 1e20 on a miss.  pos/nrm/mtl may be NULL. */
void or_closest_hit(OrScene *s, int n, const float *org, const float *dir, int32_t *id_out, float *t_out,
                    float *pos_out, float *nrm_out, float *mtl_out);
void or_visibility(OrScene *s, int n, const float *p1, const float *p2, float *vis_out);

void or_bsdf_evaluate(int n, const float *mtl, const float *wo, const float *wi, const float *N, float *out);
void or_bsdf_pdf(int n, const float *mtl, const float *wo, const float *wi, const float *N, float *out);
void or_bsdf_sample(int n, const float *mtl, const float *wo, const float *N, const float *u, const float *cur_eta,
                    float *wi_out, float *f_out, float *pdf_out, int32_t *delta_out, float *eta_out);

/* Flat vertex layouts: eye (17) pos3 normal3 thr3 mtl6 pdf_fwd pdf_rev;
 * light (20) pos3 normal3 thr3 mtl6 is_light_source is_parallel source_cutoff pdf_fwd pdf_rev. */
float or_mis_weight(const float *eye, int n_eye, int s_idx, const float *lit, int n_lit, int t_idx,
                    const float *d_vec, float dist2, const float *cam_pos);

/* Light-vertex pool (cpu_bdpt.cpp:202-325): n_lights*spl paths x light_depth vertices, 20 floats each. */
void or_light_pool(OrScene *s, int light_depth, int spl, const OrRenderOpts *opts, float *pool20);

/* Whole render (cpu_bdpt.cpp:173-488).  cam: eye3 look_at3 view_up3 fov_deg.  image: W*H*3, row 0 = top. */
void or_render(OrScene *s, const float *cam, int eye_depth, int light_depth, int W, int H, int spp, int spl,
               const OrRenderOpts *opts, float *image, OrStats *stats);

/* The reference's other two integrators (src/pt_cu.cu:20-252, src/ppm_cu.cu:64-322), restated on the Philox streams of the
 * GPU path.  Parity UNPINNED on the CPU (the reference has no CPU implementation of them and seeds curand from the clock);
 * pinned statistically against the reference's own kernels on the GPU box (tests/test_gpu_pt_ppm.py).
 * or_render_ppm: stats->shadow_rays carries the number of hit points, path_samples the photon-to-hit-point gathers. */
void or_render_pt(OrScene *s, const float *cam, int max_depth, int W, int H, int spp, const OrRenderOpts *opts, float *image, OrStats *stats);
void or_render_ppm(OrScene *s, const float *cam, int eye_depth, int light_depth, int W, int H, int spl, float radius,
                   const OrRenderOpts *opts, float *image, OrStats *stats);

/* Philox4x32-10 block and the u32 -> [0,1) float map shared with the GPU (known-answer tested). */
void or_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
float or_u32_to_unit_float(uint32_t x);
/* mt19937 + libstdc++ uniform_real_distribution<float>(0,1) restatement, for the pin test. */
void or_mt19937_floats(uint32_t seed, int n, float *out);

#ifdef __cplusplus
}
#endif
#endif
