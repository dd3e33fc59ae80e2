// b2_rng.cuh -- counter-based per-path RNG (Philox4x32-10, Salmon et al. SC'11, Random123 constants).
//
// The reference draws from one std::mt19937 per OpenMP thread (cpu_bdpt.cpp:213,338), which makes its
// sample streams depend on thread scheduling.  Here every draw is a pure function of
//     key = (seed_lo, seed_hi),  ctr = (id0, id1, block, kind)
// so 1-GPU and 8-GPU runs (and any batch split) see identical streams:
//   kind 1 = light subpath: id0 = light-path index (idx of cpu_bdpt.cpp:218), id1 = 0
//   kind 2 = eye subpath  : id0 = pixel index py*W+px,  id1 = global sample index s (cpu_bdpt.cpp:349)
//   block 0,   lanes 0,1   : emission draws (cpu_bdpt.cpp:236 / :248) or pixel jitter (:350-351)
//   block 1+k, lanes 0,1,2 : u_rr,u1,u2 of the k-th bsdf_sample call of that subpath, delta bounces
//                            included -- the reference also burns 3 uniforms per call (:297, :448)
// u32 -> float: top 24 bits * 2^-24, in [0,1).  The CPU oracle (oracle/bdpt_oracle.c) uses the same schedule.
#pragma once
#include "b2_common.cuh"

#define B2_RNG_LIGHT 1u
#define B2_RNG_EYE 2u

B2_HD uint32_t mulhi32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t) (((uint64_t) a * b) >> 32);
#endif
}

B2_HD uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = mulhi32(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        uint32_t hi1 = mulhi32(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
    }
    return c;
}
B2_HD float u32_to_unit(uint32_t x) { return (float) (x >> 8) * 5.9604644775390625e-8f; }

B2_HD void rng_draw2(uint2 key, uint32_t kind, uint32_t id0, uint32_t id1, float &a, float &b) {
    uint4 o = philox4x32_10(make_uint4(id0, id1, 0u, kind), key);
    a = u32_to_unit(o.x); b = u32_to_unit(o.y);
}
B2_HD void rng_draw3(uint2 key, uint32_t kind, uint32_t id0, uint32_t id1, uint32_t bounce, float &a, float &b, float &c) {
    uint4 o = philox4x32_10(make_uint4(id0, id1, 1u + bounce, kind), key);
    a = u32_to_unit(o.x); b = u32_to_unit(o.y); c = u32_to_unit(o.z);
}
