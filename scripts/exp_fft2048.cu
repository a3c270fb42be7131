// EXPERIMENT, not part of libpcsb200 (kept for the record; see DESIGN.md 6.1, C4, and profiles/r02_fft_experiment.txt).
// Built into the library for one measurement (`python -m pycsou_b200._build` with this file in csrc/, driven by
// scripts/probe_fft.py through two extra C entry points): results are exact (1.7e-7 relative error against a complex128
// transform, cuFFT: 3.1e-7) but the row pass takes 29.7 us and the column pass 68.6 us where cuFFT takes 23.5 / 29.7 us,
// so the C4 iteration keeps calling cuFFT.  What the column pass would need: tiles of 4-8 columns per block (32-64 bytes
// of every row instead of 16) with the lanes exchanged in rounds.
// 2-D complex64 DFT of 2048 x 2048 arrays for the compressed-sensing MRI iteration (C4; the reference's DFT2 is numpy's
// pocketfft through linop/transform -- see pycsou_b200/linop/transform.py): two passes of 2048-point transforms, rows then
// columns, each reading and writing the array once.  Unnormalised in both directions (the 1/sqrt(N) factors live in the
// kernels of pcs_cps.cu).  cuFFT's two passes take 22 us each at this size (3 TB/s); these are written for the HBM rate.
//
// One 2048-point transform = 8 x 8 x 8 x 4, decimation in frequency, 256 threads with 8 points each in registers,
// three exchanges through shared memory.  With N = R M, n = M n2 + n1, k = R k1 + k2:
//     X[R k1 + k2] = sum_n1 W_M^(n1 k1) [ W_N^(n1 k2) sum_n2 x[M n2 + n1] W_R^(n2 k2) ]
// applied three times (R = 8; M = 256, 32, 4) and a last 4-point transform:
//   stage 1  thread n1 = t                 x[256 n2 + t]         -> y[k2] W_2048^(t k2)      -> S1[k2][t]
//   stage 2  thread (k2, m1), t = m1+32 k2 S1[k2][32 m2 + m1]    -> z[j2] W_256^(m1 j2)      -> S2[k2, j2][m1]
//   stage 3  thread (k2, j2, p1)           S2[k2, j2][4 p2 + p1] -> .[q2] W_32^(p1 q2)       -> S3[k2, j2, q2][p1]
//   stage 4  thread g = k2 + 8 j2 + 64 q2 (two groups a thread)  4-point transform over p1   -> X[512 q1 + g]
// so both the loads of stage 1 and the stores of stage 4 are contiguous over the threads.  Shared-memory layouts are
// padded / skewed so that every exchange but the last read is free of bank conflicts (see idx2, idx3).
#include "pcs_common.cuh"

#include <math.h>
#include <mutex>

using namespace pcs;

namespace {

constexpr int FN = 2048, FT = 256;
constexpr int F_A = 2176, F_B = 2112;                 // words (float2) of the two exchange buffers of one transform

__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul(float2 a, float2 w) { return make_float2(a.x * w.x - a.y * w.y, a.x * w.y + a.y * w.x); }
// multiplication by -i (forward) or +i (inverse)
template <bool INV> __device__ __forceinline__ float2 rot90(float2 a) { return INV ? make_float2(-a.y, a.x) : make_float2(a.y, -a.x); }
template <bool INV> __device__ __forceinline__ float2 tw(const float2* __restrict__ table, int k) {
    const float2 w = __ldg(table + k);                 // table holds exp(-2 pi i k / 2048)
    return INV ? make_float2(w.x, -w.y) : w;
}

// 4-point transform in place: c[m] <- sum_n c[n] W_4^(n m)
template <bool INV> __device__ __forceinline__ void dft4(float2& c0, float2& c1, float2& c2, float2& c3) {
    const float2 d0 = cadd(c0, c2), d2 = csub(c0, c2), d1 = cadd(c1, c3), d3 = rot90<INV>(csub(c1, c3));
    c0 = cadd(d0, d1); c2 = csub(d0, d1); c1 = cadd(d2, d3); c3 = csub(d2, d3);
}
// 8-point transform in place, natural order: a[k] <- sum_n a[n] W_8^(n k)
template <bool INV> __device__ __forceinline__ void dft8(float2 (&a)[8]) {
    constexpr float H = 0.70710678118654752440f;
    float2 b0 = cadd(a[0], a[4]), b4 = csub(a[0], a[4]);
    float2 b1 = cadd(a[1], a[5]), b5 = csub(a[1], a[5]);
    float2 b2 = cadd(a[2], a[6]), b6 = csub(a[2], a[6]);
    float2 b3 = cadd(a[3], a[7]), b7 = csub(a[3], a[7]);
    // W_8^1 = (1 - i)/sqrt 2, W_8^2 = -i, W_8^3 = (-1 - i)/sqrt 2 (conjugates for the inverse)
    b5 = INV ? make_float2(H * (b5.x - b5.y), H * (b5.x + b5.y)) : make_float2(H * (b5.x + b5.y), H * (b5.y - b5.x));
    b6 = rot90<INV>(b6);
    b7 = INV ? make_float2(-H * (b7.x + b7.y), H * (b7.x - b7.y)) : make_float2(H * (b7.y - b7.x), -H * (b7.x + b7.y));
    dft4<INV>(b0, b1, b2, b3);
    dft4<INV>(b4, b5, b6, b7);
    a[0] = b0; a[2] = b1; a[4] = b2; a[6] = b3;
    a[1] = b4; a[3] = b5; a[5] = b6; a[7] = b7;
}

__device__ __forceinline__ int idx2(int sa, int m1) { return sa * 33 + m1; }
__device__ __forceinline__ int idx3(int k2, int j2, int q2, int p1) {
    return 4 * k2 + p1 + 32 * j2 + (j2 & 3) + 8 * (j2 >> 2) + 272 * q2;
}

// Stages 1 (after its loads) to 4 (before its stores) of NL independent transforms carried by the same 256 threads:
// v[l] = the thread's 8 points x[256 n2 + t] of transform l on entry, on exit v[l][q1 + 4 h] = X[512 q1 + t + 256 h]
// (h = 0, 1).  bufA / bufB: NL x F_A / NL x F_B words of shared memory.
template <bool INV, int NL>
__device__ __forceinline__ void fft2048_core(float2 (&v)[NL][8], float2* bufA, float2* bufB, const float2* __restrict__ table, int t) {
    // stage 1
#pragma unroll
    for (int l = 0; l < NL; ++l) dft8<INV>(v[l]);
#pragma unroll
    for (int k2 = 0; k2 < 8; ++k2) {
        const float2 w = tw<INV>(table, t * k2);
#pragma unroll
        for (int l = 0; l < NL; ++l) bufA[l * F_A + k2 * 256 + t] = k2 ? cmul(v[l][k2], w) : v[l][0];
    }
    __syncthreads();
    // stage 2
    {
        const int m1 = t & 31, k2 = t >> 5;
#pragma unroll
        for (int l = 0; l < NL; ++l) {
#pragma unroll
            for (int m2 = 0; m2 < 8; ++m2) v[l][m2] = bufA[l * F_A + k2 * 256 + 32 * m2 + m1];
            dft8<INV>(v[l]);
        }
#pragma unroll
        for (int j2 = 0; j2 < 8; ++j2) {
            const float2 w = tw<INV>(table, 8 * m1 * j2);
#pragma unroll
            for (int l = 0; l < NL; ++l) bufB[l * F_B + idx2(k2 * 8 + j2, m1)] = j2 ? cmul(v[l][j2], w) : v[l][0];
        }
    }
    __syncthreads();
    // stage 3
    {
        const int sa = t & 63, p1 = t >> 6, k2 = sa >> 3, j2 = sa & 7;
#pragma unroll
        for (int l = 0; l < NL; ++l) {
#pragma unroll
            for (int p2 = 0; p2 < 8; ++p2) v[l][p2] = bufB[l * F_B + idx2(sa, 4 * p2 + p1)];
            dft8<INV>(v[l]);
        }
#pragma unroll
        for (int q2 = 0; q2 < 8; ++q2) {
            const float2 w = tw<INV>(table, 64 * p1 * q2);
#pragma unroll
            for (int l = 0; l < NL; ++l) bufA[l * F_A + idx3(k2, j2, q2, p1)] = q2 ? cmul(v[l][q2], w) : v[l][0];
        }
    }
    __syncthreads();
    // stage 4: groups g = t and t + 256
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int g = t + 256 * h, k2 = g & 7, j2 = (g >> 3) & 7, q2 = g >> 6;
        const int base = idx3(k2, j2, q2, 0);
#pragma unroll
        for (int l = 0; l < NL; ++l) {
            float2 c0 = bufA[l * F_A + base], c1 = bufA[l * F_A + base + 1], c2 = bufA[l * F_A + base + 2], c3 = bufA[l * F_A + base + 3];
            dft4<INV>(c0, c1, c2, c3);
            v[l][4 * h] = c0; v[l][4 * h + 1] = c1; v[l][4 * h + 2] = c2; v[l][4 * h + 3] = c3;
        }
    }
}

// Row pass: one block per row of 2048 complex values.
template <bool INV>
__global__ void __launch_bounds__(FT) fft_rows_kernel(float2* __restrict__ out, const float2* __restrict__ in, const float2* __restrict__ table) {
    __shared__ float2 bufA[F_A], bufB[F_B];
    const int t = threadIdx.x;
    const size_t row = (size_t)blockIdx.x * FN;
    float2 v[1][8];
#pragma unroll
    for (int n2 = 0; n2 < 8; ++n2) v[0][n2] = in[row + 256 * n2 + t];
    fft2048_core<INV, 1>(v, bufA, bufB, table, t);
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int q1 = 0; q1 < 4; ++q1) out[row + 512 * q1 + t + 256 * h] = v[0][q1 + 4 * h];
}

// Column pass: one block per PAIR of adjacent columns of one 2048 x 2048 array (16 bytes of every row: one float4).
template <bool INV>
__global__ void __launch_bounds__(FT) fft_cols_kernel(float2* __restrict__ out, const float2* __restrict__ in, const float2* __restrict__ table) {
    extern __shared__ __align__(16) unsigned char fsm[];
    float2* bufA = reinterpret_cast<float2*>(fsm);
    float2* bufB = bufA + 2 * F_A;
    const int t = threadIdx.x;
    const size_t img = (size_t)(blockIdx.x / (FN / 2)) * FN * FN;
    const int c = (blockIdx.x % (FN / 2)) * 2;
    float2 v[2][8];
#pragma unroll
    for (int n2 = 0; n2 < 8; ++n2) {
        const float4 a = *reinterpret_cast<const float4*>(in + img + (size_t)(256 * n2 + t) * FN + c);
        v[0][n2] = make_float2(a.x, a.y);
        v[1][n2] = make_float2(a.z, a.w);
    }
    fft2048_core<INV, 2>(v, bufA, bufB, table, t);
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int q1 = 0; q1 < 4; ++q1)
            *reinterpret_cast<float4*>(out + img + (size_t)(512 * q1 + t + 256 * h) * FN + c) =
                make_float4(v[0][q1 + 4 * h].x, v[0][q1 + 4 * h].y, v[1][q1 + 4 * h].x, v[1][q1 + 4 * h].y);
}

// The table exp(-2 pi i k / 2048), k < 2048, is a static __device__ array (the library never allocates), filled in double
// precision by a kernel the first time a device is used.
__global__ void fft_twiddle_init_kernel(float2* table) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < FN) {
        double s, c;
        sincospi(-2.0 * (double)k / (double)FN, &s, &c);
        table[k] = make_float2((float)c, (float)s);
    }
}
__device__ float2 g_twiddle[FN];
struct FftDeviceState { bool table_ready, attr_fwd, attr_inv; };
FftDeviceState g_fft_dev[64];
std::mutex g_fft_mu;

const float2* twiddle_table(cudaStream_t s, bool inverse, size_t smem_cols, const void* cols_kernel) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    float2* table = nullptr;
    if (cudaGetSymbolAddress((void**)&table, g_twiddle) != cudaSuccess) return nullptr;
    std::lock_guard<std::mutex> lock(g_fft_mu);
    FftDeviceState& st = g_fft_dev[dev];
    if (!st.table_ready) {
        fft_twiddle_init_kernel<<<FN / 256, 256, 0, s>>>(table);
        if (cudaStreamSynchronize(s) != cudaSuccess) return nullptr;      // once per device: other streams may use it from now on
        st.table_ready = true;
    }
    bool& attr = inverse ? st.attr_inv : st.attr_fwd;
    if (!attr) {
        if (cudaFuncSetAttribute(cols_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_cols) != cudaSuccess) return nullptr;
        attr = true;
    }
    return table;
}

template <bool INV> int fft2_launch(float2* out, float2* tmp, const float2* in, int64_t batch, cudaStream_t s) {
    const size_t smem_cols = (size_t)2 * (F_A + F_B) * sizeof(float2);
    const float2* table = twiddle_table(s, INV, smem_cols, (const void*)fft_cols_kernel<INV>);
    if (!table) return check_launch("pcs_fft2_2048_c64");
    fft_rows_kernel<INV><<<(unsigned)(batch * FN), FT, 0, s>>>(tmp, in, table);
    fft_cols_kernel<INV><<<(unsigned)(batch * FN / 2), FT, smem_cols, s>>>(out, tmp, table);
    return check_launch("pcs_fft2_2048_c64");
}

}  // namespace

// out = DFT2(in) (inverse != 0: the inverse transform), unnormalised, for `batch` arrays of 2048 x 2048 complex64 stored one
// after the other; tmp: scratch of the same size (the array after the row pass); out may be the same buffer as in, tmp not.
extern "C" int pcs_fft2_2048_c64(void* out, void* tmp, const void* in, int64_t batch, int inverse, void* stream) {
    const char* what = "pcs_fft2_2048_c64";
    PCS_REQUIRE(out && tmp && in, "%s: null pointer", what);
    PCS_REQUIRE(batch >= 1 && batch <= 1024, "%s: batch must be in [1, 1024]", what);
    PCS_REQUIRE(tmp != in && tmp != out, "%s: tmp must not alias in or out", what);
    PCS_REQUIRE(((uintptr_t)out | (uintptr_t)tmp | (uintptr_t)in) % 16 == 0, "%s: arrays must be 16-byte aligned", what);
    cudaStream_t s = (cudaStream_t)stream;
    return inverse ? fft2_launch<true>((float2*)out, (float2*)tmp, (const float2*)in, batch, s)
                   : fft2_launch<false>((float2*)out, (float2*)tmp, (const float2*)in, batch, s);
}

// One pass alone: 2048-point transforms along the rows (axis = 1) or the columns (axis = 0) of `batch` 2048 x 2048 arrays;
// out may be in.
extern "C" int pcs_fft_2048_pass_c64(void* out, const void* in, int64_t batch, int axis, int inverse, void* stream) {
    const char* what = "pcs_fft_2048_pass_c64";
    PCS_REQUIRE(out && in, "%s: null pointer", what);
    PCS_REQUIRE(batch >= 1 && batch <= 1024, "%s: batch must be in [1, 1024]", what);
    PCS_REQUIRE(axis == 0 || axis == 1, "%s: axis must be 0 or 1", what);
    PCS_REQUIRE(((uintptr_t)out | (uintptr_t)in) % 16 == 0, "%s: arrays must be 16-byte aligned", what);
    cudaStream_t s = (cudaStream_t)stream;
    const size_t smem_cols = (size_t)2 * (F_A + F_B) * sizeof(float2);
    const float2* table = inverse ? twiddle_table(s, true, smem_cols, (const void*)fft_cols_kernel<true>)
                                  : twiddle_table(s, false, smem_cols, (const void*)fft_cols_kernel<false>);
    if (!table) return check_launch(what);
    float2* o = (float2*)out;
    const float2* i = (const float2*)in;
    if (axis == 1) {
        if (inverse) fft_rows_kernel<true><<<(unsigned)(batch * FN), FT, 0, s>>>(o, i, table);
        else fft_rows_kernel<false><<<(unsigned)(batch * FN), FT, 0, s>>>(o, i, table);
    } else {
        if (inverse) fft_cols_kernel<true><<<(unsigned)(batch * FN / 2), FT, smem_cols, s>>>(o, i, table);
        else fft_cols_kernel<false><<<(unsigned)(batch * FN / 2), FT, smem_cols, s>>>(o, i, table);
    }
    return check_launch(what);
}
