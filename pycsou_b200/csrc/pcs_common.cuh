// Shared device/host helpers for libpcsb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define PCS_OK 0
#define PCS_ERR_ARG 1
#define PCS_ERR_CUDA 2

namespace pcs {

// thread-local error string, read through pcs_last_error()
void set_error(const char* fmt, ...);
int check_launch(const char* what);
int num_sms();

#define PCS_REQUIRE(cond, ...)            \
    do {                                  \
        if (!(cond)) {                    \
            pcs::set_error(__VA_ARGS__);  \
            return PCS_ERR_ARG;           \
        }                                 \
    } while (0)

// Grid sizing: memory-bound elementwise kernels run a grid-stride loop over a grid that is a multiple of the SM
// count (148 on B200) so that every SM carries the same number of resident CTAs.
inline int grid_for(int64_t work_items, int block, int ctas_per_sm) {
    int64_t need = (work_items + block - 1) / block;
    int64_t cap = (int64_t)num_sms() * ctas_per_sm;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Sum K doubles per thread over the whole block; result valid in thread 0.  Uses warp shuffles + one smem hop.
template <int K, int BLOCK>
__device__ __forceinline__ void block_sum(double (&v)[K], double* smem /* K * (BLOCK/32) doubles */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        v[k] = warp_sum(v[k]);
        if (lane == 0) smem[k * (BLOCK / 32) + warp] = v[k];
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
        for (int k = 0; k < K; ++k) {
            double t = (lane < BLOCK / 32) ? smem[k * (BLOCK / 32) + lane] : 0.0;
            v[k] = warp_sum(t);
        }
    }
}

// Deterministic grid reduction of K partial sums: every block writes its partials to `partials[k*gridDim + block]`,
// the last block to arrive (atomic ticket) sums them in fixed order into out[0..K) and resets the ticket.
// `ticket` must be zero before the first launch (workspace is allocated zeroed by the host).
// Returns (block-uniform) whether this block was the last one.  With `sys_scope` the arrival fence is system wide, so
// that stores this grid made to PEER memory are ordered before anything the last block publishes afterwards.
template <int K, int BLOCK>
__device__ __forceinline__ bool grid_sum_finish(double (&v)[K], double* partials, unsigned int* ticket, double* out,
                                                double* smem, bool sys_scope = false) {
    __shared__ bool is_last;
    block_sum<K, BLOCK>(v, smem);
    const int nb = gridDim.x * gridDim.y * gridDim.z;
    const int bid = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < K; ++k) partials[(size_t)k * nb + bid] = v[k];
        if (sys_scope) __threadfence_system();
        else __threadfence();
        unsigned int t = atomicAdd(ticket, 1u);
        is_last = (t == (unsigned int)nb - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        double acc[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            acc[k] = 0.0;
            for (int i = threadIdx.x; i < nb; i += BLOCK) acc[k] += __ldcg(&partials[(size_t)k * nb + i]);
        }
        __syncthreads();
        block_sum<K, BLOCK>(acc, smem);
        if (threadIdx.x == 0) {
#pragma unroll
            for (int k = 0; k < K; ++k) out[k] = acc[k];
            *ticket = 0u;
        }
    }
    return is_last;
}

template <typename T> struct Vec;   // widest aligned vector (16 bytes)
template <> struct Vec<float> { typedef float4 type; static constexpr int N = 4; };
template <> struct Vec<double> { typedef double2 type; static constexpr int N = 2; };

template <typename T> __device__ __forceinline__ T tabs(T v);
template <> __device__ __forceinline__ float tabs<float>(float v) { return fabsf(v); }
template <> __device__ __forceinline__ double tabs<double>(double v) { return fabs(v); }
template <typename T> __device__ __forceinline__ T tsqrt(T v);
template <> __device__ __forceinline__ float tsqrt<float>(float v) { return sqrtf(v); }
template <> __device__ __forceinline__ double tsqrt<double>(double v) { return sqrt(v); }
template <typename T> __device__ __forceinline__ T tmax(T a, T b) { return a > b ? a : b; }
template <typename T> __device__ __forceinline__ T tmin(T a, T b) { return a < b ? a : b; }

// product rounded on its own (never contracted into an FMA)
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }

// ---- proximal maps exactly as the reference evaluates them -------------------------------------------------------
// L1: LpNorm.prox with the l-infinity ball (func/base.py:239-240, math/prox.py:253-256): x - tau*clip(x/tau,-1,1)
template <typename T> __device__ __forceinline__ T prox_l1(T x, T tau) {
    T q = x / tau;
    q = q > T(1) ? T(1) : q;
    q = q < T(-1) ? T(-1) : q;
    return x - mul_rn(tau, q);   // separately rounded product (no FMA): |x| < tau must cancel as it does in numpy
}
// group shrink factor of L21Norm.prox (func/penalty.py:554): clip(1 - tau/||g||, 0); ||g||==0 -> 1-inf -> 0
template <typename T> __device__ __forceinline__ T l21_factor(T nrm, T tau) {
    T f = T(1) - tau / nrm;  // nrm == 0 -> -inf
    return f > T(0) ? f : T(0);
}

}  // namespace pcs

// Generates the two extern "C" wrappers <name>_f32 / <name>_f64 around a templated host launcher `impl<T>`.
#define PCS_EXPORT_REAL(name, params, args)                                    \
    extern "C" int name##_f32 params { typedef float T; return name##_impl<T> args; }  \
    extern "C" int name##_f64 params { typedef double T; return name##_impl<T> args; }
