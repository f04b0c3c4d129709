// dev microbenchmark: what an event-bracketed kernel launch costs on this box
#include <cstdio>
#include <cuda_runtime.h>
__global__ void empty_k(int *p) { if (p && threadIdx.x == 9999) *p = 1; }
__global__ void busy_k(float *p, int n) { int i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) p[i] = p[i] * 1.0001f + 1.f; }
int main() {
    cudaStream_t st; cudaStreamCreate(&st);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float *buf; cudaMalloc(&buf, 64 << 20);
    cudaFuncSetAttribute(empty_k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 << 10);
    auto run = [&](const char *name, int smem, int k, bool prior, int grid) {
        float tot = 0; const int reps = 50;
        for (int r = 0; r < reps + 5; ++r) {
            if (prior) busy_k<<<4096, 256, 0, st>>>(buf, 1 << 20);
            cudaEventRecord(a, st);
            for (int i = 0; i < k; ++i) empty_k<<<grid, 128, smem, st>>>(nullptr);
            cudaEventRecord(b, st);
            cudaStreamSynchronize(st);
            float ms; cudaEventElapsedTime(&ms, a, b);
            if (r >= 5) tot += ms;
        }
        printf("%-40s %.2f us per bracket, %.2f us per kernel\n", name, tot / reps * 1e3, k ? tot / reps * 1e3 / k : 0.0);
    };
    run("no kernel", 0, 0, false, 1);
    run("1 empty grid 1", 0, 1, false, 1);
    run("1 empty grid 576 smem 0", 0, 1, false, 576);
    run("1 empty grid 576 smem 50K", 50 << 10, 1, false, 576);
    run("prior + 1 empty grid 576 smem 0", 0, 1, true, 576);
    run("prior + 1 empty grid 576 smem 50K", 50 << 10, 1, true, 576);
    run("10 empty grid 576 smem 50K", 50 << 10, 10, false, 576);
    run("prior + 10 empty grid 576 smem 50K", 50 << 10, 10, true, 576);
    return 0;
}
