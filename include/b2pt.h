/* include/b2pt.h -- the drop-in C ABI of the B200-native BDPT path (libb2pt.so).
 *
 * Plain C: pointers, sizes and PODs only; no C++/torch types cross this boundary.  Every
 * entry point returns 0 on success or a negative B2_ERR_* code; b2_last_error() gives the
 * thread-local message.  There is NO CPU fallback: without a CUDA device every compute call
 * fails with B2_ERR_NO_DEVICE.
 *
 * Reference interfaces replaced (paths relative to the reference repository):
 *   b2_scene_create / b2_scene_build_bvh  <- move_data_to_cuda_bdpt(groups, lights, spl)   include/bdpt_cu_helper.h:5,
 *                                            src/bdpt_cu_helper.cpp:13-68 (flatten groups -> POD arrays)
 *   b2_render_bdpt                        <- run_cuda_bdpt(cam, image, ...)                include/bdpt_cu_helper.h:6
 *                                            and run_cpu_bdpt(camera, groups, lights, ...) include/cpu_bdpt.h:9-17
 *                                            (the estimator reproduced is the CPU one, src/cpu_bdpt.cpp:173-488)
 *   b2_render_bdpt_multi                  <- (none: the reference is single-GPU) spp sharding + NCCL film reduce
 *   b2_intersect (closest)                <- cpu_find_closest_hit                          src/cpu_bdpt.cpp:30-80
 *   b2_intersect (any)                    <- cpu_check_visibility                          src/cpu_bdpt.cpp:82-107
 *   b2_bsdf_evaluate / _pdf / _sample     <- bsdf_evaluate / bsdf_pdf / bsdf_sample        include/geometric.cuh:419-562
 *   b2_mis_weight                         <- cpu_calculate_mis_weight                      src/cpu_bdpt.cpp:112-167
 *   b2_parse_scene_file                   <- the scene grammar loop                        src/main_cli.cpp:99-141
 *   b2_film_to_rgb8 / b2_write_image      <- clamp -> gamma 2.2 -> 8-bit + cv::imwrite     src/main_cli.cpp:221-254
 */
#ifndef B2PT_H
#define B2PT_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2_OK 0
#define B2_ERR_INVALID   (-1)
#define B2_ERR_NO_DEVICE (-2)
#define B2_ERR_CUDA      (-3)
#define B2_ERR_OOM       (-4)
#define B2_ERR_IO        (-5)
#define B2_ERR_STATE     (-6)

typedef struct B2Scene B2Scene;

/* Material, include/object.h:28-33 */
typedef struct B2Material { float base_color[3]; float roughness, metallic, eta; } B2Material;
/* Triangle, include/object.h:68-80.  obj_id = the file-order counter shared with spheres (main_cli.cpp:109,118);
 * group = key of the std::map<int,AABB> group it was added to (main_cli.cpp:110,119,126). */
typedef struct B2Tri { float v0[3], v1[3], v2[3]; int32_t obj_id, group, material; } B2Tri;
/* Sphere, include/object.h:54-66 (scale is always (1,1,1) on this path, main_cli.cpp:107) */
typedef struct B2Sphere { float center[3]; float r; int32_t obj_id, group, material; } B2Sphere;
/* CudaLight, include/geometric.cuh:73-78 as main_cli.cpp:131-139 fills it; cutoff in RADIANS. */
typedef struct B2Light { float pos[3], dir[3], illum[3]; float cutoff; int32_t is_parallel; float ball_r; } B2Light;
/* Camera, include/object.h:104-110 (fov in degrees; run_cpu_bdpt uses camera.fov, cpu_bdpt.cpp:191) */
typedef struct B2Camera { float eye[3], look_at[3], view_up[3]; float fov; } B2Camera;

#define B2_BUILDER_PLOC 0  /* parallel locally-ordered clustering over the Morton order (default) */
#define B2_BUILDER_LBVH 1  /* Karras radix tree over the Morton codes (the A/B baseline) */
typedef struct B2BvhParams {
    int32_t max_leaf_tris;  /* 1..3 triangles per leaf child (0 = default 3) */
    float   pad_rel;        /* relative padding of triangle boxes (0 = default 2^-18 of the coordinate magnitude) */
    int32_t builder;        /* B2_BUILDER_* */
    int32_t ploc_radius;    /* PLOC search radius along the Morton order (0 = default 2) */
} B2BvhParams;

typedef struct B2BvhInfo {
    uint64_t n_tris, n_nodes, node_bytes, tri_bytes;
    int32_t  depth;
    float    build_ms;      /* device time of the last build (CUDA events) */
    float    sah_cost;      /* SAH cost of the wide tree over exact boxes, root area normalised: one wide-node test = 1, one triangle test = 0.3 */
    int32_t  build_passes;  /* PLOC clustering passes of the last build */
    int32_t  reserved;
} B2BvhInfo;

#define B2_FILM_MEAN 0  /* film = sum / spp  (what run_cpu_bdpt writes, cpu_bdpt.cpp:479) */
#define B2_FILM_SUM  1  /* film = raw sum of this shard's samples (for multi-GPU reduction) */
#define B2_FILM_ACCUMULATE 2 /* film += sum / spp: one progressive pass added into a running sum (main.cpp:443-446); device entry only */
#define B2_STRATEGIES_REFERENCE 0 /* exactly the strategy set of cpu_bdpt.cpp (parity mode) */
/* The three below are NOT the reference estimator (csrc/b2_strategies.cuh): physically based spot lights, s = 1 and t = 1 only. */
#define B2_STRATEGIES_FULL           1 /* 1/2 next-event estimation (s=1) + 1/2 light tracing with t=1 film splats + emitter hits */
#define B2_STRATEGIES_LIGHT_TRACING  2 /* only the t=1 splats: W*H*spp light subpaths, atomicAdd into the film */
#define B2_STRATEGIES_NEE            3 /* only the eye subpaths with next-event estimation (the estimator the splats are validated against) */
#define B2_STRATEGIES_MIS            4 /* one light subpath per eye sample, every s/t strategy (next-event, connections, t=1 splats),
                                         balance heuristic over real area densities: SURVEY 8(f) rank 2, csrc/b2_mis.cuh */

/* The other two integrators of the reference's command line (src/main_cli.cpp:201-213); every
 * render entry point below takes them through B2RenderParams.integrator:
 *   PT  <- run_cuda_pt (include/pt_cu_helper.h:6, kernel src/pt_cu.cu:20-252): eye_depth = its max_depth, spl / light_depth unused
 *   PPM <- run_cuda_ppm (include/ppm_cu_helper.h, kernels src/ppm_cu.cu:64-322): ONE pass of n_lights*spl photons of
 *          light_depth bounces over one hit point per pixel; spp is ignored as in the reference (the progressive entry point
 *          averages passes with fresh seeds); photons do not shard over GPUs (the resolve clamps the summed flux).
 * The reference has no CPU implementation of either; intersection rules follow its GPU helpers (include/geometric.cuh:240-390). */
#define B2_INTEGRATOR_BDPT 0
#define B2_INTEGRATOR_PT   1
#define B2_INTEGRATOR_PPM  2

typedef struct B2RenderParams {
    int32_t  width, height;
    int32_t  spp;            /* total samples per pixel of the whole job */
    int32_t  spl;            /* light subpaths per light per render (cpu_bdpt.cpp:202) */
    int32_t  eye_depth, light_depth;
    uint64_t seed;           /* Philox key */
    int32_t  sample_offset;  /* this call renders samples s = sample_offset + k*sample_stride < spp */
    int32_t  sample_stride;  /* (1-GPU: 0,1; G GPUs: g,G) */
    int32_t  film_mode;      /* B2_FILM_MEAN / B2_FILM_SUM */
    int32_t  strategies;     /* B2_STRATEGIES_* */
    int32_t  max_delta;      /* cap on delta bounces per subpath (0 = default 256) */
    int32_t  max_batch_paths;/* eye samples in flight per wavefront batch (0 = up to 2^26, as HBM allows) */
    int32_t  count_traversal;/* !=0: use the counting traversal variant and fill node/tri visit counters */
    int32_t  profile_stages; /* !=0: bracket every stage launch with CUDA events and fill the ms_* stage times */
    int32_t  integrator;     /* B2_INTEGRATOR_*: main_cli's --mode (0 = bdpt, the path this library is built around) */
    float    ppm_radius;     /* ppm: gather radius = grid cell size (0 = PPM_RADIUS 0.05, include/ppm_cu.cuh:5) */
    int32_t  reserved[2];
} B2RenderParams;

typedef struct B2Stats {
    uint64_t eye_samples;     /* (pixel, s) iterations, cpu_bdpt.cpp:349 */
    uint64_t closest_rays;    /* closest-hit queries (eye + light subpaths) */
    uint64_t shadow_rays;     /* visibility queries, cpu_bdpt.cpp:424 */
    uint64_t path_samples;    /* connections reaching the MIS evaluation (:431) + depth-0 emitter hits (:372) */
    uint64_t delta_cap_hits;
    uint64_t nodes_visited;   /* only with count_traversal */
    uint64_t tris_tested;     /* only with count_traversal */
    uint64_t n_batches, n_waves, n_launches;
    uint64_t pairs_tested;    /* (eye vertex, pool vertex) pairs that went through the connection pre-test */
    uint64_t eye_vertices;    /* eye vertices created (each is connected to the whole pool), cpu_bdpt.cpp:384 */
    float    ms_total;        /* device time of the call, CUDA events */
    float    ms_light_pool;   /* pool generation (always measured) */
    float    ms_primary, ms_extend, ms_connect; /* per-stage device time (first-bounce extend, later extends, connect); only with profile_stages */
    float    ms_reserved[2];
} B2Stats;

/* ---- library / device ---- */
const char *b2_last_error(void);
int b2_device_count(int *count);
const char *b2_version(void);

/* ---- scene ---- */
int b2_scene_create(const B2Tri *tris, size_t n_tris, const B2Sphere *spheres, size_t n_spheres,
                    const B2Material *materials, size_t n_materials, const B2Light *lights, size_t n_lights,
                    int device, B2Scene **out);
int b2_scene_build_bvh(B2Scene *scene, const B2BvhParams *params);
int b2_scene_bvh_info(const B2Scene *scene, B2BvhInfo *info);
void b2_scene_destroy(B2Scene *scene);

/* ---- ray queries (host buffers; copies are part of the call) ----
 * mode B2_CLOSEST: org/dir are n*3 floats; prim_id = obj_id, -2-k for light ball k, -1 miss; t = 1e20 on a miss.
 * mode B2_ANY    : org = p1, dir = p2 (segment end points, cpu_check_visibility's arguments);
 *                  prim_id = 1 if unoccluded else 0; t is not written (may be NULL).
 * mode | B2_BRUTE_FORCE uses the exhaustive kernel instead of the BVH (parity cross-check). */
#define B2_CLOSEST 0
#define B2_ANY 1
#define B2_BRUTE_FORCE 16
#define B2_COUNT_VISITS 32
#define B2_KEEP_RAY_ORDER 128 /* walk the rays in the caller's order.  Default for closest hits (batches of >= 2^18 rays against scenes of
                               * >= 100 000 triangles): the library walks them in the order of (cell of the origin, direction octant,
                               * coarse direction) - neighbouring lanes then fetch the same nodes - and writes every result to its
                               * ray's own slot; ids and t are the same either way */
#define B2_FAST_SHADOW 64   /* with B2_ANY: the arithmetic the RENDER's shadow rays use (contracted Moeller-Trumbore, approximate reciprocals,
                            * csrc/b2_isect.cuh tri_test_fast) instead of the bit-exact test -- exposed so that tests can count where the two differ */
int b2_intersect(B2Scene *scene, const float *org, const float *dir, size_t n, int mode, int32_t *prim_id, float *t);
/* Same on device pointers (float3-packed arrays already resident in HBM), asynchronous on `stream` (cudaStream_t). */
int b2_intersect_device(B2Scene *scene, const float *d_org, const float *d_dir, size_t n, int mode,
                        int32_t *d_prim_id, float *d_t, void *stream);
/* Visit counters accumulated by B2_COUNT_VISITS launches since the last reset: out[0]=rays, [1]=nodes, [2]=tris. */
int b2_intersect_counters(B2Scene *scene, uint64_t out[3], int reset);

/* ---- render ----
 * BDPT reference strategy set: when the first connection launch of a render sends >= 16 shadow rays per eye vertex, the
 * connection stage of every later wave (and of later calls with the same spl / depths / film size on this handle) runs on
 * eye-vertex records sorted by position (30-bit Morton keys, one radix sort per wave; csrc/b2_render.cu).  This changes which
 * 32 vertices share a warp, not what any of them computes: counters are equal, films differ by the order of the fp32 atomic
 * adds only.  It holds a second record buffer (80 B per eye sample in flight) on the handle.  The environment variable
 * B2PT_CONN_SORT=0 / 1 forces caller order / sorted order (tests, A/B runs). */
int b2_render_bdpt(B2Scene *scene, const B2Camera *cam, const B2RenderParams *params, float *film_rgb_host,
                   B2Stats *stats);
/* Film stays on the device (W*H*3 fp32, caller-owned, e.g. a torch tensor); work is enqueued on `stream`
 * and the call returns after the stream has drained its own work (stats need the counters). */
int b2_render_bdpt_device(B2Scene *scene, const B2Camera *cam, const B2RenderParams *params, float *d_film_rgb,
                          void *stream, B2Stats *stats);
/* One node, n_gpus devices, ONE process: scenes[g] is the same scene created on device g (b2_scene_create(..., device g)).
 * GPU g renders sample indices s = g (mod n_gpus); the raw-sum films are summed onto scenes[0]'s device with one NCCL
 * reduce (NCCL is dlopen'ed on first use) and resolved per params->film_mode.  stats: counters summed, ms_total = max. */
int b2_render_bdpt_multi(B2Scene **scenes, int n_gpus, const B2Camera *cam, const B2RenderParams *params,
                         float *film_rgb_host, B2Stats *stats);
/* Optional: create the NCCL communicators for this device list ahead of the first b2_render_bdpt_multi (ncclCommInitAll
 * takes seconds on an 8-GPU node; without this call the first multi-GPU render pays for it).  No reference counterpart. */
int b2_multi_init(const int *devices, int n_gpus);
/* Progressive rendering, the one non-graphics behaviour of the reference viewer (src/main.cpp:399-531), headless:
 * `passes` renders of params->spp samples per pixel, pass k with Philox key params->seed + k (so every pass draws a new
 * light pool), added into a running sum; film_rgb_host receives sum / passes.  rms (may be NULL, else `passes` floats):
 * rms[k] = sqrt(sum over pixels and channels of (cur8 - last8)^2) / 255 between the tone-mapped 8-bit frames after pass
 * k and k-1 (the viewer's convergence curve, main.cpp:502-531; rms[0] = 0).  on_pass (may be NULL) is called after
 * every pass with the running average (host memory, valid during the call).  stats: counters and device time summed. */
typedef void (*B2PassCallback)(int32_t pass, const float *film_avg_rgb, int32_t width, int32_t height, void *user);
int b2_render_bdpt_progressive(B2Scene *scene, const B2Camera *cam, const B2RenderParams *params, int32_t passes,
                               float *film_rgb_host, float *rms, B2PassCallback on_pass, void *user, B2Stats *stats);
/* The viewer's frame-to-frame RMS on two 8-bit frames of n bytes each (host only). */
int b2_frame_rms(const uint8_t *cur_rgb8, const uint8_t *last_rgb8, size_t n, float *out);
/* The light-vertex pool the render would use: n_lights*spl*light_depth vertices x 20 floats
 * (pos3 normal3 thr3 mtl6 is_light_source is_parallel source_cutoff pdf_fwd pdf_rev), cpu_bdpt.cpp:202-325. */
int b2_light_pool(B2Scene *scene, const B2RenderParams *params, float *pool20_host);
/* The same pool for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Calls no cudaMalloc /
 * cudaFree / allocating library routine: the scratch b2_light_pool keeps in the scene's workspace (80-byte pool records +
 * counters) is one file-scope device array of B2_NOALLOC_MAX_POOL_VERTICES records (2.6 MB) that every GPU that loads the
 * library holds until exit.  d_pool20: DEVICE buffer of n_lights*spl*light_depth x 20 floats (layout as above), caller-owned.
 * cuda_stream: the caller's stream (the call is asynchronous on it), or NULL = the scene's own stream, blocking.
 * A pool larger than the maximum is rejected with B2_ERR_INVALID (use b2_light_pool).  Every call shares the array and
 * clears the part it uses first; calls that overlap on the GPU (other streams, other handles on the device) race on it. */
#define B2_NOALLOC_MAX_POOL_VERTICES 32768
int b2_light_pool_device_noalloc(B2Scene *scene, const B2RenderParams *params, float *d_pool20, void *cuda_stream);

/* ---- BSDF / MIS / RNG tables (device kernels over host arrays; parity tests) ---- */
int b2_bsdf_evaluate(size_t n, const float *mtl6, const float *wo, const float *wi, const float *N, float *out3);
int b2_bsdf_pdf(size_t n, const float *mtl6, const float *wo, const float *wi, const float *N, float *out);
int b2_bsdf_sample(size_t n, const float *mtl6, const float *wo, const float *N, const float *u3, const float *cur_eta,
                   float *wi_out, float *f_out, float *pdf_out, int32_t *delta_out, float *eta_out);
int b2_mis_weight(size_t n, const float *eye17, const int32_t *n_eye, const int32_t *s_idx, const float *lit20,
                  const int32_t *n_lit, const int32_t *t_idx, const float *d_vec, const float *dist2,
                  const float *cam_pos, int32_t max_verts, float *out);
int b2_philox(size_t n, const uint32_t *ctr4, const uint32_t *key2, uint32_t *out4, float *unit_out4);

/* ---- roofline denominators measured in the run: non-tensor FP32 FMA peak of `device` in TFLOP/s, and the
 * read bandwidth of a buffer of `bytes` (0 = 32 MiB) that stays resident in L2, in GB/s.  A few milliseconds each. ---- */
int b2_measure_fp32_peak(int device, float *tflops);
int b2_measure_l2_bandwidth(int device, size_t bytes, float *gbs);

/* ---- host-side surface kept from the reference (no GPU needed) ---- */
typedef struct B2ParsedScene {
    size_t n_tris, n_spheres, n_materials, n_lights;
    B2Tri *tris; B2Sphere *spheres; B2Material *materials; B2Light *lights;
    B2Camera camera;
    int32_t width, height;  /* the scene's R line */
} B2ParsedScene;
int b2_parse_scene_file(const char *path, B2ParsedScene **out);
void b2_parsed_scene_free(B2ParsedScene *p);

/* ---- OBJ / MTL meshes (north_star: tiny_obj_loader surface).  The reference vendors include/tiny_obj_loader.h (v2.0.0)
 * but has no call site and no OBJ -> Triangle adapter; this is that adapter (semantics in csrc/b2_obj.cpp, SURVEY.md
 * appendix D): the file is read by tinyobj::ObjReader::ParseFromFile (tiny_obj_loader.h:546, compiled from where the
 * header lies into csrc/b2_tinyobj.cpp), faces -> Triangle objects in file order, Kd * map_Kd texel baked per triangle
 * into Material.base_color, Pr -> roughness, Pm -> metallic, Ni -> eta for transparent illum models, triangles binned
 * into spatial `G` groups. */
#define B2_OBJ_READER_AUTO    0  /* tinyobjloader when the library was built with it (b2_obj_has_tinyobj()), else the own reader */
#define B2_OBJ_READER_OWN     1  /* the repo's own small OBJ / MTL reader (pinned against tinyobjloader in tests/test_obj.py) */
#define B2_OBJ_READER_TINYOBJ 2  /* tinyobjloader, error if it was not compiled in */
typedef struct B2ObjOptions {
    int32_t first_obj_id;     /* obj_id of the first face */
    int32_t first_group;      /* `G` key of the first spatial group; the mesh uses first_group .. first_group + grid*grid - 1 */
    int32_t group_grid;       /* g: triangles are binned into g x g groups over the x-z extent; 0 = round(cbrt(n / 2)), <= 256 */
    float   default_color[3]; /* faces without a usemtl material */
    float   default_roughness;
    int32_t reader;           /* B2_OBJ_READER_* */
    int32_t reserved[3];
} B2ObjOptions;
void b2_obj_default_options(B2ObjOptions *opt);
int b2_obj_has_tinyobj(void); /* 1 if libb2pt.so was built with the reference's vendored tiny_obj_loader.h */
/* Triangles + materials of one OBJ file (camera, lights, spheres of the result are empty). */
int b2_load_obj(const char *obj_path, const B2ObjOptions *opt, B2ParsedScene **out);
/* The reference grammar plus one directive, `O <path.obj>` (path relative to the scene file): the mesh is spliced in at
 * that point (obj ids continue, groups start at the current `G` value).  Only this entry point knows the directive:
 * the reference's own loop (main_cli.cpp:99-141, mirrored by b2_parse_scene_file) would read the path's letters as
 * directives. */
#define B2_PARSE_OBJ_DIRECTIVE 1
int b2_parse_scene_file_ex(const char *path, int flags, B2ParsedScene **out);
/* clamp [0,1] -> pow(1/2.2) -> *255 truncated, RGB order, rows top to bottom (main_cli.cpp:225-243) */
int b2_film_to_rgb8(const float *film_rgb, int32_t width, int32_t height, uint8_t *rgb8);
/* format from the extension: .ppm (P6), .png (zlib deflate), .pfm (raw float) */
int b2_write_image(const char *path, const float *film_rgb, int32_t width, int32_t height);

#ifdef __cplusplus
}
#endif
#endif /* B2PT_H */
