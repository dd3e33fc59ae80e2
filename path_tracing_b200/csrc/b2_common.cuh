// b2_common.cuh -- small vector algebra and exact-arithmetic helpers shared by all kernels.
//
// Two families of float3 operations exist on purpose, because the reference mixes two maths
// libraries whose rounding differs and the intersector has to be bit-exact against it:
//   g3_*  : glm semantics (object.cpp / cpu_bdpt.cpp camera + object tests): normalize = v * (1/sqrt(dot))
//   f3_*  : geometric.cuh:90-107 float3 operators: normalize = v / sqrt(dot)
// "x_*" scalar helpers never contract into FMA (the CPU reference is un-contracted x86-64 SSE).
// The file also compiles as plain host C++ (tests/hostsim: a test-only build of the device
// headers used to debug logic without a GPU; it is never part of libb2pt.so).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define B2_HD __host__ __device__ __forceinline__
#define B2_D __device__ __forceinline__
#define B2_NOINLINE static __host__ __device__ __noinline__
#else
#define B2_HD inline
#define B2_D inline
#define B2_NOINLINE static inline
#endif

#define B2_EPSILON 1e-4f                 // geometric.cuh:6
#define B2_PI 3.14159265358979323846f    // geometric.cuh:7
#define B2_T_INF 1e20f                   // cpu_bdpt.cpp:36

// ---- un-contracted IEEE fp32 primitives ----
#if defined(__CUDA_ARCH__)
B2_HD float x_mul(float a, float b) { return __fmul_rn(a, b); }
B2_HD float x_add(float a, float b) { return __fadd_rn(a, b); }
B2_HD float x_sub(float a, float b) { return __fsub_rn(a, b); }
B2_HD float x_div(float a, float b) { return __fdiv_rn(a, b); }
B2_HD float x_sqrt(float a) { return __fsqrt_rn(a); }
#else
B2_HD float x_mul(float a, float b) { return a * b; }
B2_HD float x_add(float a, float b) { return a + b; }
B2_HD float x_sub(float a, float b) { return a - b; }
B2_HD float x_div(float a, float b) { return a / b; }
B2_HD float x_sqrt(float a) { return sqrtf(a); }
#endif

B2_HD float3 mk3(float x, float y, float z) { return make_float3(x, y, z); }
B2_HD float3 mk3(float4 v) { return make_float3(v.x, v.y, v.z); }
B2_HD float4 mk4(float3 v, float w) { return make_float4(v.x, v.y, v.z, w); }

// ---- exact (un-contracted) vector ops: used wherever bits must match the CPU reference ----
B2_HD float3 x3_add(float3 a, float3 b) { return mk3(x_add(a.x, b.x), x_add(a.y, b.y), x_add(a.z, b.z)); }
B2_HD float3 x3_sub(float3 a, float3 b) { return mk3(x_sub(a.x, b.x), x_sub(a.y, b.y), x_sub(a.z, b.z)); }
B2_HD float3 x3_muls(float3 a, float s) { return mk3(x_mul(a.x, s), x_mul(a.y, s), x_mul(a.z, s)); }
B2_HD float3 x3_divs(float3 a, float s) { return mk3(x_div(a.x, s), x_div(a.y, s), x_div(a.z, s)); }
// dot = (x*x' + y*y') + z*z'   (glm compute_dot<vec3> and geometric.cuh:95 agree on this order)
B2_HD float x3_dot(float3 a, float3 b) { return x_add(x_add(x_mul(a.x, b.x), x_mul(a.y, b.y)), x_mul(a.z, b.z)); }
// cross: glm's (x.y*y.z - y.y*x.z, ...) and geometric.cuh:96 produce the same bits (commutative products)
B2_HD float3 x3_cross(float3 a, float3 b) {
    return mk3(x_sub(x_mul(a.y, b.z), x_mul(b.y, a.z)), x_sub(x_mul(a.z, b.x), x_mul(b.z, a.x)), x_sub(x_mul(a.x, b.y), x_mul(b.x, a.y)));
}
B2_HD float x3_length(float3 a) { return x_sqrt(x3_dot(a, a)); }
B2_HD float3 g3_normalize(float3 v) { return x3_muls(v, x_div(1.0f, x_sqrt(x3_dot(v, v)))); }  // glm::normalize
B2_HD float3 f3_normalize(float3 v) { return x3_divs(v, x3_length(v)); }                       // geometric.cuh:98
B2_HD float3 x3_neg(float3 a) { return mk3(-a.x, -a.y, -a.z); }

// ---- fast scalar ops for the shading maths (parity there is to ~1e-6 relative, not bitwise) ----
// One MUFU instruction each on the device (~1-2 ulp, denormals flushed: every caller clamps far above 1e-38); the
// compiler's own x / y and sqrtf(x) carry range fix-ups that cost 4-7 instructions apiece, and the connection stage
// evaluates ~50 of them per pair.
#if defined(__CUDA_ARCH__)
B2_HD float fast_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
B2_HD float fast_rsqrt(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
B2_HD float fast_sqrt(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
#else
B2_HD float fast_rcp(float x) { return 1.0f / x; }
B2_HD float fast_rsqrt(float x) { return 1.0f / sqrtf(x); }
B2_HD float fast_sqrt(float x) { return sqrtf(x); }
#endif
#if defined(__CUDA_ARCH__)
B2_HD float fast_div(float a, float b) { return a * fast_rcp(b); }
#else
B2_HD float fast_div(float a, float b) { return a / b; }
#endif
// sin / cos of 2 pi u, u in [0, 1): sincospif has no slow range-reduction path (sinf / cosf of an angle carry one each)
B2_HD void sincos_2pi(float u, float &s, float &c) {
#if defined(__CUDA_ARCH__)
    sincospif(2.0f * u, &s, &c);
#else
    const float phi = 2.0f * B2_PI * u; s = sinf(phi); c = cosf(phi);
#endif
}

// ---- fast vector ops (FMA contraction allowed): shading maths, where parity is statistical ----
B2_HD float3 operator+(float3 a, float3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
B2_HD float3 operator-(float3 a, float3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
B2_HD float3 operator*(float3 a, float3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
B2_HD float3 operator*(float3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
#if defined(__CUDA_ARCH__)
B2_HD float3 operator/(float3 a, float s) { const float r = fast_rcp(s); return mk3(a.x * r, a.y * r, a.z * r); }
#else      // the host build (tests/hostsim) keeps the reference's own operation order: it is compared to the oracle at 1e-5
B2_HD float3 operator/(float3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
#endif
B2_HD float dot3(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
B2_HD float3 cross3(float3 a, float3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
B2_HD float length3(float3 a) { return fast_sqrt(dot3(a, a)); }
#if defined(__CUDA_ARCH__)
B2_HD float3 normalize3(float3 a) { const float r = fast_rsqrt(dot3(a, a)); return mk3(a.x * r, a.y * r, a.z * r); }
#else
B2_HD float3 normalize3(float3 a) { return a / length3(a); }
#endif

// a / sqrt(x): one reciprocal square root on the device
B2_HD float3 div_sqrt3(float3 a, float x) {
#if defined(__CUDA_ARCH__)
    const float r = fast_rsqrt(x); return mk3(a.x * r, a.y * r, a.z * r);
#else
    return a / sqrtf(x);
#endif
}

// cpu_bdpt.cpp:12-16
B2_HD bool valid_color(float3 c) {
    return !(isnan(c.x) || isnan(c.y) || isnan(c.z) || isinf(c.x) || isinf(c.y) || isinf(c.z) || c.x < 0.0f || c.y < 0.0f || c.z < 0.0f);
}
// cpu_bdpt.cpp:18-25
B2_HD float3 clamp_radiance(float3 c, float max_val) {
    float m = c.x; if (m < c.y) m = c.y; if (m < c.z) m = c.z;
    if (m > max_val) { float s = max_val / m; return mk3(c.x * s, c.y * s, c.z * s); }
    return c;
}

B2_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    union { float f; uint32_t u; } c; c.f = f; return c.u;
#endif
}
B2_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}
B2_HD int popc32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}
B2_HD int ffs32(uint32_t x) {   // 1-based index of the lowest set bit, 0 if none
#if defined(__CUDA_ARCH__)
    return __ffs((int) x);
#else
    return __builtin_ffs((int) x);
#endif
}
B2_HD int clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __clzll((long long) x);
#else
    return x ? __builtin_clzll(x) : 64;
#endif
}
B2_HD int clz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __clz((int) x);
#else
    return x ? __builtin_clz(x) : 32;
#endif
}

// ---- atomics (the host flavour is sequential: hostsim runs "threads" one after another) ----
B2_HD void atomic_add_f32(float *p, float v) {
#if defined(__CUDA_ARCH__)
    atomicAdd(p, v);
#else
    *p += v;
#endif
}
B2_HD unsigned long long atomic_add_u64(unsigned long long *p, unsigned long long v) {
#if defined(__CUDA_ARCH__)
    return atomicAdd(p, v);
#else
    unsigned long long o = *p; *p += v; return o;
#endif
}
B2_HD uint32_t atomic_add_u32(uint32_t *p, uint32_t v) {
#if defined(__CUDA_ARCH__)
    return atomicAdd(p, v);
#else
    uint32_t o = *p; *p += v; return o;
#endif
}
