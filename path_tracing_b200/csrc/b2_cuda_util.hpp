// b2_cuda_util.hpp -- error plumbing and a tiny RAII device buffer for the host side of libb2pt.so.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <string>
#include "../../include/b2pt.h"

void b2_set_error(const char *fmt, ...);            // thread-local message behind b2_last_error()

#define B2_CUDA(call)                                                                                       \
    do {                                                                                                    \
        cudaError_t e__ = (call);                                                                           \
        if (e__ != cudaSuccess) {                                                                           \
            b2_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__);      \
            return e__ == cudaErrorMemoryAllocation ? B2_ERR_OOM : B2_ERR_CUDA;                             \
        }                                                                                                   \
    } while (0)
#define B2_TRY(expr) do { int rc__ = (expr); if (rc__ != B2_OK) return rc__; } while (0)

template <class T>
struct DevBuf {
    T *p = nullptr;
    size_t n = 0;
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    ~DevBuf() { release(); }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
    int alloc(size_t count) {
        release();
        if (count == 0) count = 1;
        cudaError_t e = cudaMalloc((void **) &p, count * sizeof(T));
        if (e != cudaSuccess) { p = nullptr; b2_set_error("cudaMalloc(%zu bytes) failed: %s", count * sizeof(T), cudaGetErrorString(e)); return B2_ERR_OOM; }
        n = count;
        return B2_OK;
    }
    int ensure(size_t count) { return (count <= n && p) ? B2_OK : alloc(count); }
    T *release_ptr() { T *q = p; p = nullptr; n = 0; return q; }
};

// Traversal stack of one CTA.  The stack holds at most one group of unvisited siblings per level 2 .. L - 1 while a node of
// level L is current (a descent from level L pushes the rest of level L's group; the root has no siblings), so bvh_depth - 2
// entries are the most a walk can stack; one spare.
#ifndef B2_STACK_SPARE
#define B2_STACK_SPARE 1
#endif
inline size_t stack_entries(int bvh_depth) { return (size_t) (bvh_depth - 2 + B2_STACK_SPARE > 2 ? bvh_depth - 2 + B2_STACK_SPARE : 2); }
inline size_t stack_bytes(int threads, int bvh_depth) { return (size_t) threads * stack_entries(bvh_depth) * sizeof(uint2); }

// RAII CUDA event: destroyed on every return path (error returns used to leak the timing events of a render)
struct CudaEvent {
    cudaEvent_t e = nullptr;
    CudaEvent() = default;
    CudaEvent(const CudaEvent &) = delete;
    CudaEvent &operator=(const CudaEvent &) = delete;
    ~CudaEvent() { if (e) cudaEventDestroy(e); }
    cudaError_t create(unsigned flags = cudaEventDefault) { return e ? cudaSuccess : cudaEventCreateWithFlags(&e, flags); }
    operator cudaEvent_t() const { return e; }
};

static inline unsigned blocks_for(size_t n, unsigned threads) { return (unsigned) ((n + threads - 1) / threads); }
