// Row-tiled 2-D stencil machinery shared by the stand-alone Gradient kernels (pcs_diff.cu) and the two-kernel Condat-Vu
// update for any Gradient kind (pcs_pds_grad.cu).
//
// Layout of the work: a thread owns V consecutive columns (V = 16 bytes / sizeof(T) when rows and bases are 16-byte
// aligned, else 1) and marches down T2_ROWS rows of a tile; a warp therefore reads 32 x 16 contiguous bytes per row
// (coalesced 128-bit accesses) and there is no per-element index arithmetic (no 64-bit division).  The row-direction
// stencil runs on a three-row register window (every row is loaded once per tile, + 2 halo rows per T2_ROWS); the
// column-direction neighbours come from the adjacent lanes by warp shuffles (lane 0 / 31: one L1-served scalar load).
//
// Finite differences follow pylops v1.x FirstDerivative (pycsou/linop/diff.py:107-121; oracle/pylops_standin.py): each
// row of D has at most three entries (fm, fc, fp) on (i-1, i, i+1); the adjoint gathers with the transposed entries.
#pragma once
#include <initializer_list>
#include <stdint.h>
#include "pcs_diff.cuh"

namespace pcs {

#ifndef PCS_T2_BLOCK
#define PCS_T2_BLOCK 128
#endif
#ifndef PCS_T2_ROWS
#define PCS_T2_ROWS 16
#endif
#ifndef PCS_T2_GROUP
#define PCS_T2_GROUP 1
#endif
#ifndef PCS_T2_GRID_PER_SM
#define PCS_T2_GRID_PER_SM 16
#endif
constexpr int T2_BLOCK = PCS_T2_BLOCK;     // threads per block, all along the row
constexpr int T2_ROWS = PCS_T2_ROWS;       // rows marched by a thread per tile
constexpr int T2_GROUP = PCS_T2_GROUP;     // rows whose loads are issued together (independent 128-bit loads in flight)
constexpr int T2_GRID_PER_SM = PCS_T2_GRID_PER_SM;   // persistent blocks per SM walking the tiles in row-major order
static_assert(T2_ROWS % T2_GROUP == 0, "tile height must be a whole number of load groups");

template <typename T> struct Coef3 { T m, c, p; };

// entries of row i of the first-derivative matrix on columns (i-1, i, i+1); inv_h = 1 / sampling step.  KIND is a
// compile-time constant: evaluated per element with run-time kind / edge the coefficient logic alone made the kernels
// issue bound (50-60 % of the HBM rate).
template <typename T, int KIND>
__device__ __forceinline__ Coef3<T> d1_row(int edge, int i, int n, T inv_h) {
    Coef3<T> r = {T(0), T(0), T(0)};
    if (i < 0 || i >= n) return r;
    if (KIND == PCS_KIND_FORWARD) {
        if (i < n - 1) { r.c = -inv_h; r.p = inv_h; }
    } else if (KIND == PCS_KIND_BACKWARD) {
        if (i >= 1) { r.m = -inv_h; r.c = inv_h; }
    } else {
        if (i >= 1 && i <= n - 2) { r.m = T(-0.5) * inv_h; r.p = T(0.5) * inv_h; }
        else if (edge && n >= 2) {
            if (i == 0) { r.c = -inv_h; r.p = inv_h; }
            else { r.m = -inv_h; r.c = inv_h; }
        }
    }
    return r;
}
// entries of row i of the TRANSPOSE on columns (i-1, i, i+1): D^T[i, i-1] = D[i-1, i] etc.
template <typename T, int KIND>
__device__ __forceinline__ Coef3<T> d1_row_adj(int edge, int i, int n, T inv_h) {
    Coef3<T> r;
    r.m = d1_row<T, KIND>(edge, i - 1, n, inv_h).p;
    r.c = d1_row<T, KIND>(edge, i, n, inv_h).c;
    r.p = d1_row<T, KIND>(edge, i + 1, n, inv_h).m;
    return r;
}
// coefficients of the V columns a thread owns (fixed for the whole tile: computed once, outside the row loop)
template <typename T, int V> struct ColCoef { Coef3<T> k[V]; };
template <typename T, int V, int KIND, bool ADJ>
__device__ __forceinline__ ColCoef<T, V> col_coef(int edge, int j0, int W, T inv_h) {
    ColCoef<T, V> c;
#pragma unroll
    for (int e = 0; e < V; ++e)
        c.k[e] = ADJ ? d1_row_adj<T, KIND>(edge, j0 + e, W, inv_h) : d1_row<T, KIND>(edge, j0 + e, W, inv_h);
    return c;
}
// c * v accumulated only where the matrix has an entry (a zero entry never touches v: inf / nan outside the stencil stay out)
template <typename T> __device__ __forceinline__ T acc3(T c, T v, T acc) { return c != T(0) ? fma(c, v, acc) : acc; }
template <> __device__ __forceinline__ float acc3<float>(float c, float v, float acc) { return c != 0.f ? fmaf(c, v, acc) : acc; }

template <typename T, int V> struct Pack { T v[V]; };

template <typename T, int V> __device__ __forceinline__ Pack<T, V> pack_zero() {
    Pack<T, V> p;
#pragma unroll
    for (int e = 0; e < V; ++e) p.v[e] = T(0);
    return p;
}
// row pointer + first column of the thread; columns >= W read as zero (V > 1 implies W % V == 0: all-or-nothing)
template <typename T, int V>
__device__ __forceinline__ Pack<T, V> pack_load(const T* __restrict__ row, int j0, int W) {
    Pack<T, V> p;
    if (j0 >= W) return pack_zero<T, V>();
    if constexpr (V == 1) {
        p.v[0] = row[j0];
    } else if constexpr (sizeof(T) * V == 16) {
        const uint4 t = *reinterpret_cast<const uint4*>(row + j0);
        *reinterpret_cast<uint4*>(p.v) = t;
    } else {
#pragma unroll
        for (int e = 0; e < V; ++e) p.v[e] = row[j0 + e];
    }
    return p;
}
template <typename T, int V>
__device__ __forceinline__ void pack_store(T* __restrict__ row, int j0, int W, const Pack<T, V>& p) {
    if (j0 >= W) return;
    if constexpr (V == 1) {
        row[j0] = p.v[0];
    } else if constexpr (sizeof(T) * V == 16) {
        *reinterpret_cast<uint4*>(row + j0) = *reinterpret_cast<const uint4*>(p.v);
    } else {
#pragma unroll
        for (int e = 0; e < V; ++e) row[j0 + e] = p.v[e];
    }
}

template <typename T> __device__ __forceinline__ T shfl_up1(T v) { return __shfl_up_sync(0xffffffffu, v, 1); }
template <typename T> __device__ __forceinline__ T shfl_dn1(T v) { return __shfl_down_sync(0xffffffffu, v, 1); }

// left / right neighbours (columns j0 - 1 and j0 + V) of this thread's pack of one row.  ALL 32 lanes must call it.
template <typename T, int V>
__device__ __forceinline__ void row_neighbours(const T* __restrict__ row, const Pack<T, V>& c, int j0, int W, int lane,
                                               T& left, T& right) {
    left = shfl_up1<T>(c.v[V - 1]);
    right = shfl_dn1<T>(c.v[0]);
    if (lane == 0) left = (j0 >= 1 && j0 - 1 < W) ? row[j0 - 1] : T(0);
    if (lane == 31) right = (j0 + V < W) ? row[j0 + V] : T(0);
}

// column-direction stencil on a pack: out[e] = sum over (j-1, j, j+1) of coef(j) * value
template <typename T, int V>
__device__ __forceinline__ Pack<T, V> col_stencil(const Pack<T, V>& c, T left, T right, const ColCoef<T, V>& kc) {
    Pack<T, V> o;
#pragma unroll
    for (int e = 0; e < V; ++e) {
        const T vm = e == 0 ? left : c.v[e == 0 ? 0 : e - 1];
        const T vp = e == V - 1 ? right : c.v[e == V - 1 ? V - 1 : e + 1];
        o.v[e] = acc3<T>(kc.k[e].p, vp, acc3<T>(kc.k[e].c, c.v[e], acc3<T>(kc.k[e].m, vm, T(0))));
    }
    return o;
}
// row-direction stencil on the three-row window
template <typename T, int V>
__device__ __forceinline__ Pack<T, V> row_stencil(const Pack<T, V>& m, const Pack<T, V>& c, const Pack<T, V>& p,
                                                  const Coef3<T>& k) {
    Pack<T, V> o;
#pragma unroll
    for (int e = 0; e < V; ++e) o.v[e] = acc3<T>(k.p, p.v[e], acc3<T>(k.c, c.v[e], acc3<T>(k.m, m.v[e], T(0))));
    return o;
}

// guarded row load: rows outside [0, H) read as zero
template <typename T, int V>
__device__ __forceinline__ Pack<T, V> row_load(const T* __restrict__ base, int i, int H, int W, int j0) {
    return (i >= 0 && i < H) ? pack_load<T, V>(base + (int64_t)i * W, j0, W) : pack_zero<T, V>();
}

// tiles of T2_ROWS rows x (T2_BLOCK * V) columns, numbered row-major; a grid-stride loop walks them
struct Tiles {
    int ntx, nty;
    long long n;
};
template <int V> __host__ __device__ inline Tiles tiles_of(int H, int W) {
    Tiles t;
    t.ntx = (W + T2_BLOCK * V - 1) / (T2_BLOCK * V);
    t.nty = (H + T2_ROWS - 1) / T2_ROWS;
    t.n = (long long)t.ntx * t.nty;
    return t;
}

// run-time kind -> compile-time KIND
#define PCS_KIND_SWITCH(kind, ...)                                                        \
    switch (kind) {                                                                       \
        case PCS_KIND_FORWARD: { constexpr int KIND = PCS_KIND_FORWARD; __VA_ARGS__; } break;   \
        case PCS_KIND_BACKWARD: { constexpr int KIND = PCS_KIND_BACKWARD; __VA_ARGS__; } break; \
        default: { constexpr int KIND = PCS_KIND_CENTERED; __VA_ARGS__; } break;          \
    }

// 16-byte vector path possible?  (rows, every plane base and the plane stride must be 16-byte multiples)
template <typename T> inline bool t2_vec_ok(int64_t W, std::initializer_list<const void*> ptrs) {
    if ((W * (int64_t)sizeof(T)) % 16 != 0) return false;
    for (const void* p : ptrs)
        if (p != nullptr && ((uintptr_t)p % 16) != 0) return false;
    return true;
}

}  // namespace pcs
