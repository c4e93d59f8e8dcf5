// Development microbenchmark (GPU box): FP64 dependent-op latencies and graph kernel-node cost.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_lat(double *out, long long *clk, double a, double b) {
    double x = a; long long t0, t1;
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 256; i++) x = __fma_rn(x, b, a);
    t1 = clock64(); clk[0] = t1 - t0;
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 256; i++) x = a / (x + b);
    t1 = clock64(); clk[1] = t1 - t0;
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 256; i++) x = sqrt(x + b);
    t1 = clock64(); clk[2] = t1 - t0;
    float f = (float)x;
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 256; i++) f = (float)((double)f + b);
    t1 = clock64(); clk[3] = t1 - t0;
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 64; i++) x = sin(x) + b;
    t1 = clock64(); clk[4] = t1 - t0;
    const double *p = out + 1;
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 64; i++) p = out + 1 + (long long)(*(volatile double *)p);
    t1 = clock64(); clk[5] = t1 - t0;
    long long j = (long long)(x * 1e6) + 123456789;
    int d = (int)b + 287;
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 64; i++) j = j / d + 1234567;
    t1 = clock64(); clk[6] = t1 - t0;
    out[0] = x + f + (double)j + (double)(p - out);
}
__global__ void k_empty(int *p) { if (p && threadIdx.x == 9999) *p = 1; }
int main() {
    double *out; long long *clk;
    cudaMalloc(&out, 1024); cudaMalloc(&clk, 64); cudaMemset(out, 0, 1024);
    k_lat<<<1, 1>>>(out, clk, 1.0000001, 0.9999999);
    long long h[8]; cudaMemcpy(h, clk, 64, cudaMemcpyDeviceToHost);
    printf("cycles per dependent op: dfma %.1f  ddiv(+add) %.1f  dsqrt(+add) %.1f  cvt-add-cvt %.1f  sin(+add) %.1f  L2/L1 ld chain %.1f  i64div %.1f\n",
           h[0] / 256., h[1] / 256., h[2] / 256., h[3] / 256., h[4] / 64., h[5] / 64., h[6] / 64.);
    cudaStream_t st; cudaStreamCreate(&st);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int blocks : {1, 148, 1024}) {
        cudaGraph_t g; cudaGraphExec_t ge;
        cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
        for (int i = 0; i < 64; i++) k_empty<<<blocks, 256, 0, st>>>(nullptr);
        cudaStreamEndCapture(st, &g); cudaGraphInstantiate(&ge, g, 0);
        cudaGraphLaunch(ge, st); cudaStreamSynchronize(st);
        cudaEventRecord(e0, st);
        for (int r = 0; r < 20; r++) cudaGraphLaunch(ge, st);
        cudaEventRecord(e1, st); cudaStreamSynchronize(st);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("graph of 64 empty kernels x %d blocks: %.2f us per kernel node\n", blocks, ms * 1e3 / (20 * 64));
        cudaEventRecord(e0, st);
        for (int r = 0; r < 20 * 64; r++) k_empty<<<blocks, 256, 0, st>>>(nullptr);
        cudaEventRecord(e1, st); cudaStreamSynchronize(st);
        cudaEventElapsedTime(&ms, e0, e1);
        printf("stream launches of empty kernels x %d blocks: %.2f us per kernel\n", blocks, ms * 1e3 / (20 * 64));
    }
    return 0;
}
