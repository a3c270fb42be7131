// Device stencils shared by the stand-alone difference kernels and the fused solver kernels.
// `p` points at the element at axis position `i` (0 <= i < n); neighbours along the axis are `st` elements apart.
// kind: 0 forward, 1 centered, 2 backward (pylops v1.x FirstDerivative; restated in oracle/pylops_standin.py).
#pragma once
#include "pcs_common.cuh"

#define PCS_KIND_FORWARD 0
#define PCS_KIND_CENTERED 1
#define PCS_KIND_BACKWARD 2

namespace pcs {

template <typename T>
__device__ __forceinline__ T d1_fwd(const T* p, int64_t i, int64_t n, int64_t st, int kind, int edge, T inv_h) {
    if (kind == PCS_KIND_FORWARD) return (i < n - 1) ? (p[st] - p[0]) * inv_h : T(0);
    if (kind == PCS_KIND_BACKWARD) return (i >= 1) ? (p[0] - p[-st]) * inv_h : T(0);
    if (i >= 1 && i <= n - 2) return (T(0.5) * p[st] - T(0.5) * p[-st]) * inv_h;
    if (!edge) return T(0);
    return (i == 0) ? (p[st] - p[0]) * inv_h : (p[0] - p[-st]) * inv_h;
}

// gather form of the transpose: out[i] = sum_j D[j,i] z[j]
template <typename T>
__device__ __forceinline__ T d1_adj(const T* p, int64_t i, int64_t n, int64_t st, int kind, int edge, T inv_h) {
    T v = T(0);
    if (kind == PCS_KIND_FORWARD) {
        if (i <= n - 2) v -= p[0] * inv_h;
        if (i >= 1) v += p[-st] * inv_h;
        return v;
    }
    if (kind == PCS_KIND_BACKWARD) {
        if (i <= n - 2) v -= p[st] * inv_h;
        if (i >= 1) v += p[0] * inv_h;
        return v;
    }
    // centered: interior rows j in [1, n-2] scatter -0.5 z[j] to j-1 and +0.5 z[j] to j+1
    if (i + 1 >= 1 && i + 1 <= n - 2) v -= (T(0.5) * p[st]) * inv_h;
    if (i - 1 >= 1 && i - 1 <= n - 2) v += (T(0.5) * p[-st]) * inv_h;
    if (edge) {
        if (i == 0) v -= p[0] * inv_h;           // row 0:  (x1 - x0)/h
        if (i == 1) v += p[-st] * inv_h;
        if (i == n - 2) v -= p[st] * inv_h;      // row n-1: (x[n-1] - x[n-2])/h
        if (i == n - 1) v += p[0] * inv_h;
    }
    return v;
}

template <typename T>
__device__ __forceinline__ T d2_fwd(const T* p, int64_t i, int64_t n, int64_t st, int edge, T inv_h2) {
    if (i >= 1 && i <= n - 2) return (p[st] - T(2) * p[0] + p[-st]) * inv_h2;
    if (!edge) return T(0);
    if (i == 0) return (p[0] - T(2) * p[st] + p[2 * st]) * inv_h2;
    return (p[-2 * st] - T(2) * p[-st] + p[0]) * inv_h2;
}

template <typename T>
__device__ __forceinline__ T d2_adj(const T* p, int64_t i, int64_t n, int64_t st, int edge, T inv_h2) {
    T v = T(0);
    if (i + 1 >= 1 && i + 1 <= n - 2) v += p[st];
    if (i >= 1 && i <= n - 2) v -= T(2) * p[0];
    if (i - 1 >= 1 && i - 1 <= n - 2) v += p[-st];
    if (edge) {
        // row 0: x0 - 2 x1 + x2 ; row n-1: x[n-3] - 2 x[n-2] + x[n-1]
        if (i == 0) v += p[0];
        if (i == 1) v -= T(2) * p[-st];
        if (i == 2) v += p[-2 * st];
        if (i == n - 3) v += p[2 * st];
        if (i == n - 2) v -= T(2) * p[st];
        if (i == n - 1) v += p[0];
    }
    return v * inv_h2;
}

}  // namespace pcs
