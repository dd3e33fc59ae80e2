// b2_warp.cuh -- warp-level helpers shared by the wavefront kernels (b2_render.cu, b2_integrators.cu): ballot + prefix
// popcount compaction with one atomic per warp.
#pragma once
#include <cstdint>

__device__ __forceinline__ uint32_t lanemask_lt() { uint32_t m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m; }

// warp-reduce a per-thread count and add it to a global 64-bit counter with one atomic per warp
__device__ __forceinline__ void warp_count_add(unsigned long long *counter, uint32_t v) {
    v = __reduce_add_sync(0xffffffffu, v);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(counter, (unsigned long long) v);
}
// warp-aggregated slot allocation: lanes with `want` get consecutive indices from *counter (one atomic per warp)
__device__ __forceinline__ int warp_alloc(int *counter, bool want) {
    const uint32_t m = __ballot_sync(0xffffffffu, want);
    if (m == 0u) return -1;
    int base = 0;
    if ((threadIdx.x & 31) == (__ffs((int) m) - 1)) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(0xffffffffu, base, __ffs((int) m) - 1);
    return want ? base + __popc(m & lanemask_lt()) : -1;
}
