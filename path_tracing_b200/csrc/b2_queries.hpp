// b2_queries.hpp -- host-side declarations of the batch query / table kernels (b2_queries.cu).
#pragma once
#include <algorithm>
#include "b2_cuda_util.hpp"
#include "b2_scene.cuh"

// mode: B2_CLOSEST / B2_ANY, optionally | B2_BRUTE_FORCE, | B2_COUNT_VISITS (d_counters: rays, nodes, tris).
// All pointers are device pointers; asynchronous on `st`.
// Scratch of the ray reordering (keys, ray indices, sorted copies, CUB's temporary storage); owned by the scene, grown on demand.
struct RaySortScratch { DevBuf<uint32_t> keys, keys_sorted, idx, idx_sorted; DevBuf<unsigned char> tmp; };
// (key, index) pairs of `ss.keys` / `ss.idx` sorted by the low `end_bit` key bits into `ss.keys_sorted` / `ss.idx_sorted`;
// b2_sort_reserve sizes the five buffers for n pairs first (the connection stage's eye-vertex reordering uses the same scratch type)
int b2_sort_reserve(RaySortScratch &ss, size_t n);
int b2_sort_pairs(RaySortScratch &ss, size_t n, int end_bit, cudaStream_t st);
int b2_intersect_launch(const DevScene &sc, int bvh_depth, const float *d_org, const float *d_dir, size_t n, int mode,
                        int32_t *d_prim_id, float *d_t, unsigned long long *d_counters, cudaStream_t st, RaySortScratch *sort_scratch = nullptr);

int b2_tables_bsdf_evaluate(size_t n, const float *mtl6, const float *wo, const float *wi, const float *N, float *out3);
int b2_tables_bsdf_pdf(size_t n, const float *mtl6, const float *wo, const float *wi, const float *N, float *out);
int b2_tables_bsdf_sample(size_t n, const float *mtl6, const float *wo, const float *N, const float *u3, const float *cur_eta,
                          float *wi_out, float *f_out, float *pdf_out, int32_t *delta_out, float *eta_out);
int b2_tables_mis_weight(size_t n, const float *eye17, const int32_t *n_eye, const int32_t *s_idx, const float *lit20, const int32_t *n_lit,
                         const int32_t *t_idx, const float *d_vec, const float *dist2, const float *cam_pos, int32_t max_verts, float *out);
int b2_tables_philox(size_t n, const uint32_t *ctr4, const uint32_t *key2, uint32_t *out4, float *unit_out4);
