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

// Bulk-copy engine variant (the TMA unit's 1-D form, SASS UBLKCP): no thread touches the data.  One
// elected thread per CTA moves chunks of CH doubles of each of the nine planes global -> shared
// (cp.async.bulk + mbarrier complete_tx) and shared -> global (cp.async.bulk bulk_group) through a ring
// of S stages, LA = S - 2 loads ahead.  Answers whether a TMA-staged LB kernel could see more HBM
// bandwidth than the LDG.256 / STG.256 form.
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
template <int CH, int S>
__global__ void __launch_bounds__(32) bulk9(const double *src, double *dst, long long plane, long long nchunks) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *buf = reinterpret_cast<double *>(smem_raw);  // [S][9][CH]
    __shared__ __align__(8) unsigned long long mbar[S];
    if (threadIdx.x != 0) return;
    for (int s = 0; s < S; s++)
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar[s])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    const long long first = blockIdx.x, stride = gridDim.x;
    const long long mine = first < nchunks ? (nchunks - first + stride - 1) / stride : 0;
    constexpr unsigned kBytes = CH * 8;
    constexpr int LA = S - 2;
    auto issue_load = [&](long long j) {
        const int s = (int)(j % S);
        const long long c = first + j * stride;
        const unsigned mb = smem_u32(&mbar[s]);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(9u * kBytes) : "memory");
#pragma unroll
        for (int i = 0; i < 9; i++)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(smem_u32(buf + ((long long)s * 9 + i) * CH)), "l"(src + i * plane + c * CH),
                           "r"(kBytes), "r"(mb) : "memory");
    };
    for (long long j = 0; j < LA && j < mine; j++) issue_load(j);
    for (long long k = 0; k < mine; k++) {
        if (k + LA < mine) {
            // stage (k-2) % S is loaded next: its stores (group k-2) must have read it out
            asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            issue_load(k + LA);
        }
        const int s = (int)(k % S);
        const unsigned mb = smem_u32(&mbar[s]), parity = (unsigned)((k / S) & 1);
        asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@!p bra W;\n}"
                     ::"r"(mb), "r"(parity) : "memory");
        const long long c = first + k * stride;
#pragma unroll
        for (int i = 0; i < 9; i++)
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                         ::"l"(dst + i * plane + c * CH), "r"(smem_u32(buf + ((long long)s * 9 + i) * CH)),
                           "r"(kBytes) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
template <int CH, int S>
static float run_bulk(const double *a, double *b, long long plane, long long nodes, int ctas_per_sm,
                      cudaEvent_t e0, cudaEvent_t e1) {
    const int smem = S * 9 * CH * 8;
    cudaFuncSetAttribute(bulk9<CH, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    float best = 1e9;
    for (int rep = 0; rep < 12; rep++) {
        cudaEventRecord(e0);
        bulk9<CH, S><<<148 * ctas_per_sm, 32, smem>>>(a, b, plane, nodes / CH);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep >= 2 && ms < best) best = ms;
    }
    return best;
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
    {
        // bulk-copy engine: (chunk per plane, stages, CTAs per SM)
        float t;
        t = run_bulk<512, 6>(a, b, plane, nodes, 1, e0, e1);   // 4 KB x 9 x 6 = 216 KB, 1 CTA/SM
        printf("%-45s %.4f ms  %.1f GB/s\n", "cp.async.bulk 9 planes, 4 KB chunks, 6 stages, 1 CTA/SM", t, bytes / t / 1e6);
        t = run_bulk<256, 6>(a, b, plane, nodes, 2, e0, e1);   // 2 KB x 9 x 6 = 108 KB, 2 CTAs/SM
        printf("%-45s %.4f ms  %.1f GB/s\n", "cp.async.bulk 9 planes, 2 KB chunks, 6 stages, 2 CTA/SM", t, bytes / t / 1e6);
        t = run_bulk<1024, 3>(a, b, plane, nodes, 1, e0, e1);  // 8 KB x 9 x 3 = 216 KB, 1 CTA/SM
        printf("%-45s %.4f ms  %.1f GB/s\n", "cp.async.bulk 9 planes, 8 KB chunks, 3 stages, 1 CTA/SM", t, bytes / t / 1e6);
        t = run_bulk<128, 8>(a, b, plane, nodes, 3, e0, e1);   // 1 KB x 9 x 8 = 72 KB, 3 CTAs/SM
        printf("%-45s %.4f ms  %.1f GB/s\n", "cp.async.bulk 9 planes, 1 KB chunks, 8 stages, 3 CTA/SM", t, bytes / t / 1e6);
        // correctness of the last variant: b must equal a on a sampled plane
        cudaMemset(b, 0xff, 9 * plane * 8);
        bulk9<512, 6><<<148, 32, 6 * 9 * 512 * 8>>>(a, b, plane, nodes / 512);
        double ha[4], hb[4];
        bool same = true;
        for (int i = 0; i < 9; i += 4) {
            cudaMemcpy(ha, a + i * plane + nodes - 4, 32, cudaMemcpyDeviceToHost);
            cudaMemcpy(hb, b + i * plane + nodes - 4, 32, cudaMemcpyDeviceToHost);
            for (int k = 0; k < 4; k++) same = same && ha[k] == hb[k];
        }
        printf("bulk copy check (tail of planes 0, 4, 8): %s\n", same ? "ok" : "MISMATCH");
    }
    printf("cuda error: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
