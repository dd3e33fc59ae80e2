// b2_peaks.cu -- in-run microbenchmarks for the roofline denominators that MEASURED_PEAKS.json does not carry
// ("FP32_peak measured by an in-run FMA microbench"; traversal of a scene whose nodes + triangles fit in the
// 126 MB L2 is bounded by L2, not HBM, bandwidth).  Called by bench.py / bench_traversal.py through the C ABI; each takes
// a few milliseconds.
#include "b2_cuda_util.hpp"

namespace {

// 8 independent FMA chains per thread: no memory traffic, enough ILP to issue one FFMA per cycle per lane
__global__ void __launch_bounds__(256) k_fma_peak(float *out, float a, float b, int iters) {
    float x0 = threadIdx.x * 1e-3f, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f, x7 = x0 + 7.f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
        }
    }
    const float s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 123.456f) out[0] = s;              // never true: keeps the chains alive
}

// every thread streams 128-bit loads over a buffer that stays resident in L2 (re-read `reps` times)
__global__ void __launch_bounds__(256) k_l2_read(const uint4 *__restrict__ buf, size_t n16, int reps, unsigned *sink) {
    unsigned acc = 0u;
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    for (int r = 0; r < reps; ++r)
        for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
            uint4 v;
            asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(buf + i));   // .cg: L2 only, no L1
            acc ^= v.x ^ v.y ^ v.z ^ v.w;
        }
    if (acc == 0x12345u) sink[0] = acc;
}

int pick_device(int device, int &sms) {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || device < 0 || device >= n_dev) { b2_set_error("no CUDA device %d", device); return B2_ERR_NO_DEVICE; }
    B2_CUDA(cudaSetDevice(device));
    sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    return B2_OK;
}

}  // namespace

extern "C" int b2_measure_fp32_peak(int device, float *tflops) {
    if (!tflops) { b2_set_error("NULL argument"); return B2_ERR_INVALID; }
    int sms = 148;
    B2_TRY(pick_device(device, sms));
    DevBuf<float> out; B2_TRY(out.alloc(4));
    CudaEvent e0, e1; B2_CUDA(e0.create()); B2_CUDA(e1.create());
    const int iters = 2048, grid = sms * 16;            // 8 chains x 16 x iters FMAs per thread
    float best = 0.0f;
    for (int rep = 0; rep < 4; ++rep) {                 // first pass warms up (clock ramp); best of the rest
        B2_CUDA(cudaEventRecord(e0));
        k_fma_peak<<<grid, 256>>>(out.p, 1.0000001f, 1e-7f, iters);
        B2_CUDA(cudaEventRecord(e1));
        B2_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f; B2_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double flops = 2.0 * 8.0 * 16.0 * iters * 256.0 * grid;
        if (rep > 0 && ms > 0.f) best = fmaxf(best, (float) (flops / (ms * 1e-3) / 1e12));
    }
    B2_CUDA(cudaGetLastError());
    *tflops = best;
    return B2_OK;
}

extern "C" int b2_measure_l2_bandwidth(int device, size_t bytes, float *gbs) {
    if (!gbs) { b2_set_error("NULL argument"); return B2_ERR_INVALID; }
    int sms = 148;
    B2_TRY(pick_device(device, sms));
    if (bytes == 0) bytes = (size_t) 32 << 20;          // well inside the L2 of either die
    const size_t n16 = bytes / sizeof(uint4);
    DevBuf<uint4> buf; B2_TRY(buf.alloc(n16));
    DevBuf<unsigned> sink; B2_TRY(sink.alloc(4));
    B2_CUDA(cudaMemset(buf.p, 1, n16 * sizeof(uint4)));
    CudaEvent e0, e1; B2_CUDA(e0.create()); B2_CUDA(e1.create());
    const int reps = 32, grid = sms * 8;
    float best = 0.0f;
    for (int rep = 0; rep < 4; ++rep) {
        B2_CUDA(cudaEventRecord(e0));
        k_l2_read<<<grid, 256>>>(buf.p, n16, reps, sink.p);
        B2_CUDA(cudaEventRecord(e1));
        B2_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f; B2_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms > 0.f) best = fmaxf(best, (float) ((double) n16 * sizeof(uint4) * reps / (ms * 1e-3) / 1e9));
    }
    B2_CUDA(cudaGetLastError());
    *gbs = best;
    return B2_OK;
}
