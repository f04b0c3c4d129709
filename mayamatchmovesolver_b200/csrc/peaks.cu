// Measured FP64 ceilings of the device: the roofline denominators of the compute-bound
// kernels (SURVEY.md section 8d asks for them; MEASURED_PEAKS.json only has HBM and bf16).
//   DFMA: 8 independent fma chains per thread, 1024 threads x 2 CTAs per SM.
//   DMMA: 8 independent mma.sync.m8n8k4.f64 accumulator pairs per warp.
#include <cuda_runtime.h>

#include <algorithm>
#include <string>
#include <vector>

#include "device_problem.h"

namespace mmb {

__global__ void __launch_bounds__(1024) k_peak_dfma(double *out, int iters, double a, double b) {
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
        x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
        x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
    const double s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 123.456) out[0] = s;  // keeps the chains alive
}

__global__ void __launch_bounds__(256) k_peak_dmma(double *out, int iters, double a, double b) {
    double c[8][2];
#pragma unroll
    for (int t = 0; t < 8; ++t) { c[t][0] = threadIdx.x + t; c[t][1] = t; }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int t = 0; t < 8; ++t)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                         : "+d"(c[t][0]), "+d"(c[t][1])
                         : "d"(a), "d"(b));
    }
    double s = 0.0;
#pragma unroll
    for (int t = 0; t < 8; ++t) s += c[t][0] + c[t][1];
    if (s == 123.456) out[0] = s;
}

mmb_status measure_fp64_peaks(int device, double *dfma_tflops, double *dmma_tflops) {
    MMB_CUDA_CHECK(cudaSetDevice(device));
    int sms = 0;
    MMB_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    double *out = nullptr;
    MMB_CUDA_CHECK(cudaMalloc((void **)&out, sizeof(double)));
    cudaEvent_t e0, e1;
    MMB_CUDA_CHECK(cudaEventCreate(&e0));
    MMB_CUDA_CHECK(cudaEventCreate(&e1));
    const int iters = 20000;
    double best_fma = 0.0, best_mma = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        float ms = 0.f;
        cudaEventRecord(e0);
        k_peak_dfma<<<sms * 2, 1024>>>(out, iters, 0.999999, 1e-9);
        cudaEventRecord(e1);
        MMB_CUDA_CHECK(cudaEventSynchronize(e1));
        cudaEventElapsedTime(&ms, e0, e1);
        const double fl = (double)sms * 2 * 1024 * 8.0 * 2.0 * iters;  // 8 FMAs = 16 flops per thread per iteration
        if (rep) best_fma = std::max(best_fma, fl / (ms * 1e-3) / 1e12);
        cudaEventRecord(e0);
        k_peak_dmma<<<sms * 4, 256>>>(out, iters, 0.5, 1e-9);
        cudaEventRecord(e1);
        MMB_CUDA_CHECK(cudaEventSynchronize(e1));
        cudaEventElapsedTime(&ms, e0, e1);
        const double flm = (double)sms * 4 * 8 * 8.0 * 512.0 * iters;  // 8 warps x 8 DMMAs x (8*8*4*2) flops
        if (rep) best_mma = std::max(best_mma, flm / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(out);
    if (dfma_tflops) *dfma_tflops = best_fma;
    if (dmma_tflops) *dmma_tflops = best_mma;
    return MMB_OK;
}

// What a plain streaming kernel reaches at the SIZE of one Jacobian launch: reads `read_bytes`,
// writes `write_bytes` (16-byte accesses, one wave of CTAs), L2 flushed before every repetition
// exactly as bench.py does.  MEASURED_PEAKS.json's copy figure is taken over 2 GiB; a 96 MB
// transfer spends a visible part of its time ramping up and draining.
__global__ void __launch_bounds__(256) k_stream_rw(const double2 *__restrict__ in, long long nr, double2 *__restrict__ out,
                                                   long long nw) {
    const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x, step = (long long)gridDim.x * blockDim.x;
    double s = 0.0;
    for (long long i = t; i < nr; i += step) { const double2 v = in[i]; s += v.x + v.y; }
    const double2 val = make_double2(s, 1.5);
    for (long long i = t; i < nw; i += step) out[i] = val;
}

mmb_status measure_stream_floor(int device, long long read_bytes, long long write_bytes, int reps, double *ms_median) {
    MMB_CUDA_CHECK(cudaSetDevice(device));
    if (read_bytes < 0 || write_bytes <= 0 || reps < 1) return MMB_ERR_INVALID_ARGUMENT;
    int sms = 0;
    MMB_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    const size_t flush_bytes = (size_t)256 << 20;
    double2 *in = nullptr, *out = nullptr;
    void *flush = nullptr;
    MMB_CUDA_CHECK(cudaMalloc((void **)&in, (size_t)std::max<long long>(read_bytes, 16)));
    MMB_CUDA_CHECK(cudaMalloc((void **)&out, (size_t)write_bytes));
    MMB_CUDA_CHECK(cudaMalloc(&flush, flush_bytes));
    MMB_CUDA_CHECK(cudaMemset(in, 0, (size_t)std::max<long long>(read_bytes, 16)));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    std::vector<float> ms(reps);
    cudaError_t e = cudaSuccess;
    for (int r = 0; r < reps + 2 && e == cudaSuccess; ++r) {
        cudaMemsetAsync(flush, r, flush_bytes);
        cudaEventRecord(e0);
        k_stream_rw<<<sms * 8, 256>>>(in, read_bytes / 16, out, write_bytes / 16);
        cudaEventRecord(e1);
        e = cudaEventSynchronize(e1);
        if (r >= 2) cudaEventElapsedTime(&ms[r - 2], e0, e1);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(in); cudaFree(out); cudaFree(flush);
    if (e != cudaSuccess) { set_last_error(std::string("stream floor: ") + cudaGetErrorString(e)); return MMB_ERR_CUDA; }
    std::sort(ms.begin(), ms.end());
    if (ms_median) *ms_median = ms[reps / 2];
    return MMB_OK;
}

}  // namespace mmb
