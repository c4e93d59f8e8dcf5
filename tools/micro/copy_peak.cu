// Development microbenchmark (GPU box): what a pure copy achieves on this part with the access
// pattern of the LB kernels — 9 read streams and 9 write streams of 8192-row columns, 128- or
// 256-bit accesses, no arithmetic.  The ceiling lb_fused_v2 / lb_fused_v4 can be held against
// (MEASURED_PEAKS.json is a torch copy_, which both exceed).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/copy_peak tools/micro/copy_peak.cu
#include <cstdio>
#include <cuda_runtime.h>

struct __align__(32) d4 { double x, y, z, w; };
__device__ __forceinline__ d4 ld4(const double *p) {
    d4 v;
    asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void st4(double *p, d4 v) {
    asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v.x), "d"(v.y), "d"(v.z), "d"(v.w) : "memory");
}

// one thread: 9 planes x 2 (or 4) doubles, like the LB kernels (loads first, then stores)
__global__ void __launch_bounds__(256) copy9_v2(const double *src, double *dst, long long plane, long long n2) {
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n2) return;
    double2 v[9];
#pragma unroll
    for (int i = 0; i < 9; i++) v[i] = __ldg(reinterpret_cast<const double2 *>(src + i * plane) + g);
#pragma unroll
    for (int i = 0; i < 9; i++) reinterpret_cast<double2 *>(dst + i * plane)[g] = v[i];
}
__global__ void __launch_bounds__(256) copy9_v4(const double *src, double *dst, long long plane, long long n4) {
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n4) return;
    d4 v[9];
#pragma unroll
    for (int i = 0; i < 9; i++) v[i] = ld4(src + i * plane + 4 * g);
#pragma unroll
    for (int i = 0; i < 9; i++) st4(dst + i * plane + 4 * g, v[i]);
}
// plain one-stream copy, 256-bit, grid-stride
__global__ void __launch_bounds__(256) copy1_v4(const double *src, double *dst, long long n4) {
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n4; g += (long long)gridDim.x * blockDim.x)
        st4(dst + 4 * g, ld4(src + 4 * g));
}
// in place: read nine planes, write the same addresses back (what an AA-pattern / esoteric-pull
// LB step would do to HBM: one buffer, reads and writes in the same DRAM pages)
__global__ void __launch_bounds__(256) inplace9_v4(double *buf, long long plane, long long n4) {
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n4) return;
    d4 v[9];
#pragma unroll
    for (int i = 0; i < 9; i++) {
        asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[i].x), "=d"(v[i].y), "=d"(v[i].z), "=d"(v[i].w) : "l"(buf + i * plane + 4 * g));
    }
#pragma unroll
    for (int i = 0; i < 9; i++) {
        v[i].x += 1.;
        st4(buf + i * plane + 4 * g, v[i]);
    }
}
__global__ void read9_v4(const double *src, double *dst, long long plane, long long n4) {
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n4) return;
    double s = 0;
#pragma unroll
    for (int i = 0; i < 9; i++) { d4 v = ld4(src + i * plane + 4 * g); s += v.x + v.y + v.z + v.w; }
    if (s == 1.2345e300) dst[g] = s;
}

int main() {
    const long long nodes = 8192ll * 8192, plane = nodes;
    double *a, *b;
    cudaMalloc(&a, 9 * plane * 8);
    cudaMalloc(&b, 9 * plane * 8);
    cudaMemset(a, 0, 9 * plane * 8);
    cudaMemset(b, 0, 9 * plane * 8);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const double bytes = 144.0 * nodes;
    for (int variant = 0; variant < 5; variant++) {
        float best = 1e9;
        for (int rep = 0; rep < 12; rep++) {
            cudaEventRecord(e0);
            if (variant == 0) copy9_v2<<<(unsigned)((nodes / 2 + 255) / 256), 256>>>(a, b, plane, nodes / 2);
            if (variant == 1) copy9_v4<<<(unsigned)((nodes / 4 + 255) / 256), 256>>>(a, b, plane, nodes / 4);
            if (variant == 2) copy1_v4<<<148 * 16, 256>>>(a, b, 9 * nodes / 4);
            if (variant == 3) read9_v4<<<(unsigned)((nodes / 4 + 255) / 256), 256>>>(a, b, plane, nodes / 4);
            if (variant == 4) inplace9_v4<<<(unsigned)((nodes / 4 + 255) / 256), 256>>>(a, plane, nodes / 4);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (rep >= 2 && ms < best) best = ms;
        }
        const char *names[5] = {"copy 9 planes, 128-bit, loads then stores", "copy 9 planes, 256-bit, loads then stores",
                                "copy one stream, 256-bit, grid-stride", "read only 9 planes, 256-bit",
                                "in place 9 planes, 256-bit (read + write back)"};
        const double by = variant == 3 ? bytes / 2 : bytes;
        printf("%-45s %.4f ms  %.1f GB/s\n", names[variant], best, by / best / 1e6);
    }
    printf("cuda error: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
