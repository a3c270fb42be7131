// Scratch micro-benchmark: do the fp64 tensor-core instructions (mma.sync ... f64) and plain DFMA run on separate
// pipes on sm_100a, and at what rates?   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ubench_dmma ubench_dmma.cu
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
#ifdef BIG_SHAPES
__device__ __forceinline__ void dmma1688(double* c, const double* a, const double* b) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
#endif

// MODE 0: DFMA only (NF chains / thread); 1: DMMA only (NM accumulator pairs / warp); 2: both interleaved
template <int MODE, int NF, int NM>
__global__ void __launch_bounds__(256) k(double* out, int iters, double seed) {
    double f[NF > 0 ? NF : 1];
    double c[NM > 0 ? 2 * NM : 2];
    const double a = seed + threadIdx.x * 1e-9, b = 1.0 - 1e-9 * threadIdx.x;
#pragma unroll
    for (int i = 0; i < NF; ++i) f[i] = seed * i;
#pragma unroll
    for (int i = 0; i < 2 * NM; ++i) c[i] = seed * i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            if (MODE != 1) {
#pragma unroll
                for (int i = 0; i < NF; ++i) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(f[i]) : "d"(a), "d"(b));
            }
            if (MODE != 0) {
#pragma unroll
                for (int i = 0; i < NM; ++i) dmma884(c[2 * i], c[2 * i + 1], a, b);
            }
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NF; ++i) s += f[i];
#pragma unroll
    for (int i = 0; i < 2 * NM; ++i) s += c[i];
    if (s == 12345.678) out[0] = s;
}

#ifdef BIG_SHAPES
__global__ void __launch_bounds__(256) kbig(double* out, int iters, double seed) {
    double c[4][4];
    double a[4], b[2];
    for (int i = 0; i < 4; ++i) { a[i] = seed + i + threadIdx.x * 1e-9; for (int j = 0; j < 4; ++j) c[i][j] = seed * (i + j); }
    b[0] = 1.0 - 1e-9 * threadIdx.x; b[1] = 0.5;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int i = 0; i < 4; ++i) dmma1688(c[i], a, b);
    }
    double s = 0;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
    if (s == 12345.678) out[0] = s;
}
#endif

template <typename F> static float timeit(F f) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}

int main() {
    double* out; CK(cudaMalloc(&out, 8));
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    const int nsm = p.multiProcessorCount;
    const int iters = 20000;
    for (int bps = 1; bps <= 4; bps *= 2) {
        const int grid = nsm * bps, warps = grid * 8;
        const double nthreads = (double)grid * 256;
        float t;
        t = timeit([&] { k<0, 8, 0><<<grid, 256>>>(out, iters, 1.0); });
        printf("blocks/SM %d (warps/SMSP %d): DFMA only x8 chains        %8.3f ms  %7.2f TFLOP/s\n", bps, 2 * bps, t, 2.0 * nthreads * 8 * 4 * iters / t / 1e9);
        t = timeit([&] { k<1, 0, 4><<<grid, 256>>>(out, iters, 1.0); });
        printf("blocks/SM %d                 : DMMA m8n8k4 only x4 accs     %8.3f ms  %7.2f TFLOP/s\n", bps, t, 2.0 * warps * 256.0 * 4 * 4 * iters / t / 1e9);
        t = timeit([&] { k<1, 0, 8><<<grid, 256>>>(out, iters, 1.0); });
        printf("blocks/SM %d                 : DMMA m8n8k4 only x8 accs     %8.3f ms  %7.2f TFLOP/s\n", bps, t, 2.0 * warps * 256.0 * 8 * 4 * iters / t / 1e9);
        t = timeit([&] { k<2, 8, 4><<<grid, 256>>>(out, iters, 1.0); });
        printf("blocks/SM %d                 : both (8 DFMA + 4 DMMA)       %8.3f ms  DFMA %7.2f + DMMA %7.2f TFLOP/s\n", bps, t,
               2.0 * nthreads * 8 * 4 * iters / t / 1e9, 2.0 * warps * 256.0 * 4 * 4 * iters / t / 1e9);
        t = timeit([&] { k<2, 4, 4><<<grid, 256>>>(out, iters, 1.0); });
        printf("blocks/SM %d                 : both (4 DFMA + 4 DMMA)       %8.3f ms  DFMA %7.2f + DMMA %7.2f TFLOP/s\n", bps, t,
               2.0 * nthreads * 4 * 4 * iters / t / 1e9, 2.0 * warps * 256.0 * 4 * 4 * iters / t / 1e9);
#ifdef BIG_SHAPES
        t = timeit([&] { kbig<<<grid, 256>>>(out, iters, 1.0); });
        printf("blocks/SM %d                 : DMMA m16n8k8 x4 accs         %8.3f ms  %7.2f TFLOP/s\n", bps, t, 2.0 * warps * 1024.0 * 4 * 4 * iters / t / 1e9);
#endif
    }
    CK(cudaDeviceSynchronize());
    return 0;
}
