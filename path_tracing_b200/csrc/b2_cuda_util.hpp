// b2_cuda_util.hpp -- error plumbing and a tiny RAII device buffer for the host side of libb2pt.so.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <mutex>
#include <string>
#include <vector>
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

// Device blocks that a QUIESCENT owner hands back (b2_scene_destroy, after its stream has drained) instead of freeing them,
// for the next allocation on the same device that fits.  cudaMalloc / cudaFree take driver-wide locks: with several ranks of one
// node creating and destroying scenes in step, nine cudaFree calls of a scene handle were measured at 0.46 - 0.56 s on some
// ranks (bench.py e2e at N = 2 and 4, profiles/r02_bench_n{2,4}_v5.json), against 2 - 6 ms otherwise.  Bounded: at most 64
// blocks and 1 GiB per process; a block serves requests of at least half its size; an allocation failure flushes the cache.
struct DevBlockCache {
    struct Blk { void *p; size_t bytes; int dev; };
    static std::mutex &mtx() { static std::mutex m; return m; }
    static std::vector<Blk> &blocks() { static std::vector<Blk> v; return v; }
    static size_t &held() { static size_t h = 0; return h; }
    static void *take(int dev, size_t bytes, size_t &got) {
        std::lock_guard<std::mutex> lk(mtx());
        std::vector<Blk> &v = blocks();
        int best = -1;
        for (int i = 0; i < (int) v.size(); ++i)
            if (v[i].dev == dev && v[i].bytes >= bytes && v[i].bytes <= 2 * bytes + 4096 && (best < 0 || v[i].bytes < v[best].bytes)) best = i;
        if (best < 0) return nullptr;
        void *p = v[best].p; got = v[best].bytes; held() -= got;
        v.erase(v.begin() + best);
        return p;
    }
    static bool give(int dev, void *p, size_t bytes) {
        std::lock_guard<std::mutex> lk(mtx());
        if (blocks().size() >= 64 || held() + bytes > ((size_t) 1 << 30)) return false;
        blocks().push_back(Blk{ p, bytes, dev }); held() += bytes;
        return true;
    }
    static void flush() {
        std::lock_guard<std::mutex> lk(mtx());
        int cur = 0; cudaGetDevice(&cur);
        for (const Blk &b : blocks()) { cudaSetDevice(b.dev); cudaFree(b.p); }
        cudaSetDevice(cur);
        blocks().clear(); held() = 0;
    }
};

template <class T>
struct DevBuf {
    T *p = nullptr;
    size_t n = 0;
    size_t bytes = 0;          // size of the block behind p (>= n * sizeof(T) when it came from the cache)
    int dev = -1;
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    ~DevBuf() { release(); }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; bytes = 0; }
    // for owners whose work on the block has completed (stream drained): keep the block for the next allocation that fits
    void release_cached() { if (p && !DevBlockCache::give(dev, p, bytes)) cudaFree(p); p = nullptr; n = 0; bytes = 0; }
    int alloc(size_t count) {
        release();
        if (count == 0) count = 1;
        const size_t want = count * sizeof(T);
        if (cudaGetDevice(&dev) != cudaSuccess) dev = -1;
        size_t got = 0;
        if (void *q = DevBlockCache::take(dev, want, got)) { p = (T *) q; n = count; bytes = got; return B2_OK; }
        cudaError_t e = cudaMalloc((void **) &p, want);
        if (e != cudaSuccess) { cudaGetLastError(); DevBlockCache::flush(); e = cudaMalloc((void **) &p, want); }
        if (e != cudaSuccess) { p = nullptr; b2_set_error("cudaMalloc(%zu bytes) failed: %s", want, cudaGetErrorString(e)); return B2_ERR_OOM; }
        n = count; bytes = want;
        return B2_OK;
    }
    int ensure(size_t count) { return (count <= n && p) ? B2_OK : alloc(count); }
    T *release_ptr() { T *q = p; p = nullptr; n = 0; bytes = 0; return q; }
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
