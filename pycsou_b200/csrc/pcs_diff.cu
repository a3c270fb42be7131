// libpcsb200: finite-difference stencils (FirstDerivative / SecondDerivative / Gradient / Laplacian), sm_100a.
//
// Replaces pylops v1.x FirstDerivative / SecondDerivative / Gradient / Laplacian reached from
// pycsou/linop/diff.py:128-130, 218-219, 882, 957.  Boundary rules restated in
// oracle/pylops_standin.py.  Any n-D C-order array differentiated along one axis is viewed as
// (outer, n, inner); consecutive threads walk consecutive flat indices, so every access is coalesced and the
// +-inner neighbours are served by L1/L2 (HBM sees each element once).
#include "pcs_tile2d.cuh"
#include <algorithm>

using namespace pcs;

constexpr int DF_BLOCK = 256;
constexpr int DF_CTAS_PER_SM = 8;

template <typename T, bool ADJ, bool SECOND>
__global__ void __launch_bounds__(DF_BLOCK) deriv_kernel(T* __restrict__ out, const T* __restrict__ x, int64_t outer,
                                                         int64_t n, int64_t inner, int kind, int edge, T inv_h,
                                                         int accumulate) {
    const int64_t total = outer * n * inner;
    const int64_t stride = (int64_t)gridDim.x * DF_BLOCK;
    for (int64_t f = (int64_t)blockIdx.x * DF_BLOCK + threadIdx.x; f < total; f += stride) {
        const int64_t i = (f / inner) % n;
        T v;
        if (SECOND) v = ADJ ? d2_adj<T>(x + f, i, n, inner, edge, inv_h) : d2_fwd<T>(x + f, i, n, inner, edge, inv_h);
        else v = ADJ ? d1_adj<T>(x + f, i, n, inner, kind, edge, inv_h) : d1_fwd<T>(x + f, i, n, inner, kind, edge, inv_h);
        out[f] = accumulate ? out[f] + v : v;
    }
}

// ---- first derivative along one axis, row-tiled (pcs_tile2d.cuh) ------------------------------------------------------
// (outer, n, inner) with inner > 1: `outer` images of n rows x inner columns, stencil down the rows (three-row register
// window); inner == 1: outer rows of n samples, stencil along the row (lane shuffles).  128-bit accesses where the rows are
// 16-byte multiples.  ADJ: gather form of the transpose.  accumulate: out += (n-D Gradient adjoint sums its axes).
constexpr int DERIV_GROUP = 2;
template <typename T, int V, int KIND, bool ADJ>
__global__ void __launch_bounds__(T2_BLOCK) deriv_rows_kernel(T* __restrict__ out, const T* __restrict__ x, int64_t outer,
                                                              int n, int inner, int edge, T ih, int accumulate) {
    const Tiles tl = tiles_of<V>(n, inner);
    const long long total = (long long)outer * tl.n;
    for (long long t = blockIdx.x; t < total; t += gridDim.x) {
        const long long img = t / tl.n, tt = t - img * tl.n;
        const int ty = (int)(tt / tl.ntx), tx = (int)(tt - (long long)ty * tl.ntx);
        const int j0 = (tx * T2_BLOCK + (int)threadIdx.x) * V;
        const int r0 = ty * T2_ROWS, r1 = min(n, r0 + T2_ROWS);
        const T* __restrict__ xi = x + img * (int64_t)n * inner;
        T* __restrict__ oi = out + img * (int64_t)n * inner;
        Pack<T, V> m = row_load<T, V>(xi, r0 - 1, n, inner, j0);
        Pack<T, V> c = row_load<T, V>(xi, r0, n, inner, j0);
        for (int i0 = r0; i0 < r1; i0 += DERIV_GROUP) {
            Pack<T, V> nx[DERIV_GROUP], acc[DERIV_GROUP];
#pragma unroll
            for (int g = 0; g < DERIV_GROUP; ++g) {
                nx[g] = row_load<T, V>(xi, i0 + g + 1, n, inner, j0);
                acc[g] = (accumulate && i0 + g < r1) ? row_load<T, V>(oi, i0 + g, n, inner, j0) : pack_zero<T, V>();
            }
#pragma unroll
            for (int g = 0; g < DERIV_GROUP; ++g) {
                const int i = i0 + g;
                if (i < r1) {
                    const Coef3<T> k = ADJ ? d1_row_adj<T, KIND>(edge, i, n, ih) : d1_row<T, KIND>(edge, i, n, ih);
                    Pack<T, V> o = row_stencil<T, V>(m, c, nx[g], k);
#pragma unroll
                    for (int e = 0; e < V; ++e) o.v[e] += acc[g].v[e];
                    pack_store<T, V>(oi + (int64_t)i * inner, j0, inner, o);
                }
                m = c;
                c = nx[g];
            }
        }
    }
}

template <typename T, int V, int KIND, bool ADJ>
__global__ void __launch_bounds__(T2_BLOCK) deriv_cols_kernel(T* __restrict__ out, const T* __restrict__ x, int rows, int n,
                                                              int edge, T ih, int accumulate) {
    const Tiles tl = tiles_of<V>(rows, n);
    const int lane = threadIdx.x & 31;
    for (long long t = blockIdx.x; t < tl.n; t += gridDim.x) {
        const int ty = (int)(t / tl.ntx), tx = (int)(t - (long long)ty * tl.ntx);
        const int j0 = (tx * T2_BLOCK + (int)threadIdx.x) * V;
        const int r0 = ty * T2_ROWS, r1 = min(rows, r0 + T2_ROWS);
        const ColCoef<T, V> kc = col_coef<T, V, KIND, ADJ>(edge, j0, n, ih);
        for (int i0 = r0; i0 < r1; i0 += DERIV_GROUP) {
            Pack<T, V> c[DERIV_GROUP], acc[DERIV_GROUP];
#pragma unroll
            for (int g = 0; g < DERIV_GROUP; ++g) {
                c[g] = row_load<T, V>(x, i0 + g, rows, n, j0);
                acc[g] = (accumulate && i0 + g < r1) ? row_load<T, V>(out, i0 + g, rows, n, j0) : pack_zero<T, V>();
            }
#pragma unroll
            for (int g = 0; g < DERIV_GROUP; ++g) {
                const int i = i0 + g;
                if (i < r1) {                                   // (uniform over the block: the shuffles are safe)
                    T l, r;
                    row_neighbours<T, V>(x + (int64_t)i * n, c[g], j0, n, lane, l, r);
                    Pack<T, V> o = col_stencil<T, V>(c[g], l, r, kc);
#pragma unroll
                    for (int e = 0; e < V; ++e) o.v[e] += acc[g].v[e];
                    pack_store<T, V>(out + (int64_t)i * n, j0, n, o);
                }
            }
        }
    }
}

template <typename T, bool ADJ>
static bool deriv_tiled(T* out, const T* x, int64_t outer, int64_t n, int64_t inner, int kind, int edge, T ih,
                        int accumulate, cudaStream_t s) {
    constexpr int VEC = 16 / (int)sizeof(T);
    const int64_t lim = (1ll << 31) - 4096;
    if (n >= lim || inner >= lim) return false;
    if (inner > 1) {
        const bool vec = t2_vec_ok<T>(inner, {out, x});
        const Tiles tl = vec ? tiles_of<VEC>((int)n, (int)inner) : tiles_of<1>((int)n, (int)inner);
        const int grid = (int)std::min<long long>(std::max<long long>(outer * tl.n, 1), (long long)num_sms() * T2_GRID_PER_SM);
        if (vec) {
            PCS_KIND_SWITCH(kind, (deriv_rows_kernel<T, VEC, KIND, ADJ><<<grid, T2_BLOCK, 0, s>>>(out, x, outer, (int)n, (int)inner, edge, ih, accumulate)));
        } else {
            PCS_KIND_SWITCH(kind, (deriv_rows_kernel<T, 1, KIND, ADJ><<<grid, T2_BLOCK, 0, s>>>(out, x, outer, (int)n, (int)inner, edge, ih, accumulate)));
        }
        return true;
    }
    if (outer >= lim) return false;
    const bool vec = t2_vec_ok<T>(n, {out, x});
    const Tiles tl = vec ? tiles_of<VEC>((int)outer, (int)n) : tiles_of<1>((int)outer, (int)n);
    const int grid = (int)std::min<long long>(std::max<long long>(tl.n, 1), (long long)num_sms() * T2_GRID_PER_SM);
    if (vec) {
        PCS_KIND_SWITCH(kind, (deriv_cols_kernel<T, VEC, KIND, ADJ><<<grid, T2_BLOCK, 0, s>>>(out, x, (int)outer, (int)n, edge, ih, accumulate)));
    } else {
        PCS_KIND_SWITCH(kind, (deriv_cols_kernel<T, 1, KIND, ADJ><<<grid, T2_BLOCK, 0, s>>>(out, x, (int)outer, (int)n, edge, ih, accumulate)));
    }
    return true;
}

template <typename T, bool ADJ, bool SECOND>
static int deriv_launch(T* out, const T* x, int64_t outer, int64_t n, int64_t inner, int kind, int edge, double step,
                        int accumulate, cudaStream_t s, const char* what) {
    PCS_REQUIRE(outer >= 0 && n >= 0 && inner >= 0, "%s: negative dimension", what);
    PCS_REQUIRE(kind >= 0 && kind <= 2, "%s: kind must be 0 (forward), 1 (centered) or 2 (backward)", what);
    PCS_REQUIRE(step != 0.0, "%s: zero step", what);
    const int64_t total = outer * n * inner;
    if (total == 0) return PCS_OK;
    PCS_REQUIRE(out && x, "%s: null pointer", what);
    PCS_REQUIRE(out != x, "%s: in-place stencils are not allowed (caller ping-pongs)", what);
    PCS_REQUIRE(!SECOND || !edge || n >= 3, "%s: edge=True needs at least 3 samples along the axis", what);
    PCS_REQUIRE(SECOND || kind != 1 || !edge || n >= 2, "%s: edge=True needs at least 2 samples along the axis", what);
    const T inv_h = SECOND ? (T)(1.0 / (step * step)) : (T)(1.0 / step);
    if constexpr (!SECOND) {
        if (deriv_tiled<T, ADJ>(out, x, outer, n, inner, kind, edge, inv_h, accumulate, s)) return check_launch(what);
    }
    deriv_kernel<T, ADJ, SECOND><<<grid_for(total, DF_BLOCK, DF_CTAS_PER_SM), DF_BLOCK, 0, s>>>(
        out, x, outer, n, inner, kind, edge, inv_h, accumulate);
    return check_launch(what);
}

#define DERIV_PARAMS                                                                                               \
    (void* out, const void* x, int64_t outer, int64_t n, int64_t inner, int kind, int edge, double step,           \
     int accumulate, void* stream)
#define DERIV_ARGS ((T*)out, (const T*)x, outer, n, inner, kind, edge, step, accumulate, (cudaStream_t)stream)

template <typename T>
static int pcs_firstderiv_fwd_impl(T* o, const T* x, int64_t a, int64_t n, int64_t b, int k, int e, double h, int acc,
                                   cudaStream_t s) {
    return deriv_launch<T, false, false>(o, x, a, n, b, k, e, h, acc, s, "pcs_firstderiv_fwd");
}
PCS_EXPORT_REAL(pcs_firstderiv_fwd, DERIV_PARAMS, DERIV_ARGS)
template <typename T>
static int pcs_firstderiv_adj_impl(T* o, const T* x, int64_t a, int64_t n, int64_t b, int k, int e, double h, int acc,
                                   cudaStream_t s) {
    return deriv_launch<T, true, false>(o, x, a, n, b, k, e, h, acc, s, "pcs_firstderiv_adj");
}
PCS_EXPORT_REAL(pcs_firstderiv_adj, DERIV_PARAMS, DERIV_ARGS)
template <typename T>
static int pcs_secondderiv_fwd_impl(T* o, const T* x, int64_t a, int64_t n, int64_t b, int k, int e, double h, int acc,
                                    cudaStream_t s) {
    return deriv_launch<T, false, true>(o, x, a, n, b, 0, e, h, acc, s, "pcs_secondderiv_fwd");
}
PCS_EXPORT_REAL(pcs_secondderiv_fwd, DERIV_PARAMS, DERIV_ARGS)
template <typename T>
static int pcs_secondderiv_adj_impl(T* o, const T* x, int64_t a, int64_t n, int64_t b, int k, int e, double h, int acc,
                                    cudaStream_t s) {
    return deriv_launch<T, true, true>(o, x, a, n, b, 0, e, h, acc, s, "pcs_secondderiv_adj");
}
PCS_EXPORT_REAL(pcs_secondderiv_adj, DERIV_PARAMS, DERIV_ARGS)

// ------------------------------------------------------------------------------------------------ 2-D gradient
// out = [d/daxis0 ; d/daxis1] (two planes of H*W), pylops Gradient = VStack of per-axis FirstDerivative.
// (two rows of loads in flight per thread: measured 4.7-5.0 TB/s at 4096^2 against 4.2-4.4 with one, profiles/r02_tile_sweep.txt;
// the Condat-Vu pair of pcs_pds_grad.cu, which carries more registers per row, is fastest with one)
constexpr int GRAD_GROUP = 2;
// Row-tiled kernels (pcs_tile2d.cuh): 128-bit coalesced accesses, three-row register window, lane shuffles for the
// column neighbours, no per-element index arithmetic.  Algorithmic traffic: forward 3 N s, adjoint 3 N s.
template <typename T, int V, int KIND>
__global__ void __launch_bounds__(T2_BLOCK) grad2d_fwd_kernel(T* __restrict__ out, const T* __restrict__ x, int H, int W,
                                                              int edge, T ih0, T ih1) {
    const Tiles tl = tiles_of<V>(H, W);
    const int lane = threadIdx.x & 31;
    T* __restrict__ out1 = out + (int64_t)H * W;
    for (long long t = blockIdx.x; t < tl.n; t += gridDim.x) {
        const int ty = (int)(t / tl.ntx), tx = (int)(t - (long long)ty * tl.ntx);
        const int j0 = (tx * T2_BLOCK + (int)threadIdx.x) * V;
        const int r0 = ty * T2_ROWS, r1 = min(H, r0 + T2_ROWS);
        const ColCoef<T, V> kc = col_coef<T, V, KIND, false>(edge, j0, W, ih1);
        Pack<T, V> m = row_load<T, V>(x, r0 - 1, H, W, j0);
        Pack<T, V> c = row_load<T, V>(x, r0, H, W, j0);
        for (int i0 = r0; i0 < r1; i0 += GRAD_GROUP) {
            Pack<T, V> nx[GRAD_GROUP];
#pragma unroll
            for (int g = 0; g < GRAD_GROUP; ++g) nx[g] = row_load<T, V>(x, i0 + g + 1, H, W, j0);
#pragma unroll
            for (int g = 0; g < GRAD_GROUP; ++g) {
                const int i = i0 + g;
                if (i < r1) {                                   // (uniform over the block: the shuffles below are safe)
                    const T* __restrict__ row = x + (int64_t)i * W;
                    pack_store<T, V>(out + (int64_t)i * W, j0, W,
                                     row_stencil<T, V>(m, c, nx[g], d1_row<T, KIND>(edge, i, H, ih0)));
                    T l, r;
                    row_neighbours<T, V>(row, c, j0, W, lane, l, r);
                    pack_store<T, V>(out1 + (int64_t)i * W, j0, W, col_stencil<T, V>(c, l, r, kc));
                }
                m = c;
                c = nx[g];
            }
        }
    }
}

template <typename T, int V, int KIND>
__global__ void __launch_bounds__(T2_BLOCK) grad2d_adj_kernel(T* __restrict__ out, const T* __restrict__ z, int H, int W,
                                                              int edge, T ih0, T ih1) {
    const Tiles tl = tiles_of<V>(H, W);
    const int lane = threadIdx.x & 31;
    const T* __restrict__ z1 = z + (int64_t)H * W;
    for (long long t = blockIdx.x; t < tl.n; t += gridDim.x) {
        const int ty = (int)(t / tl.ntx), tx = (int)(t - (long long)ty * tl.ntx);
        const int j0 = (tx * T2_BLOCK + (int)threadIdx.x) * V;
        const int r0 = ty * T2_ROWS, r1 = min(H, r0 + T2_ROWS);
        const ColCoef<T, V> kc = col_coef<T, V, KIND, true>(edge, j0, W, ih1);
        Pack<T, V> m = row_load<T, V>(z, r0 - 1, H, W, j0);
        Pack<T, V> c = row_load<T, V>(z, r0, H, W, j0);
        for (int i0 = r0; i0 < r1; i0 += GRAD_GROUP) {
            Pack<T, V> nx[GRAD_GROUP], c1[GRAD_GROUP];
#pragma unroll
            for (int g = 0; g < GRAD_GROUP; ++g) {
                nx[g] = row_load<T, V>(z, i0 + g + 1, H, W, j0);
                c1[g] = row_load<T, V>(z1, i0 + g, H, W, j0);
            }
#pragma unroll
            for (int g = 0; g < GRAD_GROUP; ++g) {
                const int i = i0 + g;
                if (i < r1) {
                    Pack<T, V> o = row_stencil<T, V>(m, c, nx[g], d1_row_adj<T, KIND>(edge, i, H, ih0));
                    T l, r;
                    row_neighbours<T, V>(z1 + (int64_t)i * W, c1[g], j0, W, lane, l, r);
                    const Pack<T, V> a1 = col_stencil<T, V>(c1[g], l, r, kc);
#pragma unroll
                    for (int e = 0; e < V; ++e) o.v[e] += a1.v[e];
                    pack_store<T, V>(out + (int64_t)i * W, j0, W, o);
                }
                m = c;
                c = nx[g];
            }
        }
    }
}

template <typename T>
static int grad2d_check(const void* out, const void* x, int64_t H, int64_t W, int kind, int edge, double s0, double s1,
                        const char* what) {
    PCS_REQUIRE(H >= 0 && W >= 0, "%s: negative dimension", what);
    PCS_REQUIRE(kind >= 0 && kind <= 2, "%s: kind must be 0 (forward), 1 (centered) or 2 (backward)", what);
    PCS_REQUIRE(s0 != 0.0 && s1 != 0.0, "%s: zero step", what);
    if (H * W == 0) return PCS_OK;
    PCS_REQUIRE(out && x && out != x, "%s: null or aliased pointer", what);
    PCS_REQUIRE(kind != 1 || !edge || (H >= 2 && W >= 2), "%s: centered+edge needs >= 2 samples per axis", what);
    return PCS_OK;
}

template <typename T>
static int pcs_grad2d_fwd_impl(T* out, const T* x, int64_t H, int64_t W, int kind, int edge, double s0, double s1,
                               cudaStream_t s) {
    int rc = grad2d_check<T>(out, x, H, W, kind, edge, s0, s1, "pcs_grad2d_fwd");
    if (rc || H * W == 0) return rc;
    PCS_REQUIRE(H < (1ll << 31) - 64 && W < (1ll << 31) - 4096, "pcs_grad2d_fwd: image side too large");
    constexpr int VEC = 16 / (int)sizeof(T);
    const T ih0 = (T)(1.0 / s0), ih1 = (T)(1.0 / s1);
    if (t2_vec_ok<T>(W, {out, x})) {
        const Tiles tl = tiles_of<VEC>((int)H, (int)W);
        const int grid = (int)std::min<long long>(tl.n, (long long)num_sms() * T2_GRID_PER_SM);
        PCS_KIND_SWITCH(kind, (grad2d_fwd_kernel<T, VEC, KIND><<<grid, T2_BLOCK, 0, s>>>(out, x, (int)H, (int)W, edge, ih0, ih1)));
    } else {
        const Tiles tl = tiles_of<1>((int)H, (int)W);
        const int grid = (int)std::min<long long>(tl.n, (long long)num_sms() * T2_GRID_PER_SM);
        PCS_KIND_SWITCH(kind, (grad2d_fwd_kernel<T, 1, KIND><<<grid, T2_BLOCK, 0, s>>>(out, x, (int)H, (int)W, edge, ih0, ih1)));
    }
    return check_launch("pcs_grad2d_fwd");
}
PCS_EXPORT_REAL(pcs_grad2d_fwd,
                (void* out, const void* x, int64_t H, int64_t W, int kind, int edge, double step0, double step1,
                 void* stream),
                ((T*)out, (const T*)x, H, W, kind, edge, step0, step1, (cudaStream_t)stream))

template <typename T>
static int pcs_grad2d_adj_impl(T* out, const T* z, int64_t H, int64_t W, int kind, int edge, double s0, double s1,
                               cudaStream_t s) {
    int rc = grad2d_check<T>(out, z, H, W, kind, edge, s0, s1, "pcs_grad2d_adj");
    if (rc || H * W == 0) return rc;
    PCS_REQUIRE(H < (1ll << 31) - 64 && W < (1ll << 31) - 4096, "pcs_grad2d_adj: image side too large");
    constexpr int VEC = 16 / (int)sizeof(T);
    const T ih0 = (T)(1.0 / s0), ih1 = (T)(1.0 / s1);
    if (t2_vec_ok<T>(W, {out, z})) {
        const Tiles tl = tiles_of<VEC>((int)H, (int)W);
        const int grid = (int)std::min<long long>(tl.n, (long long)num_sms() * T2_GRID_PER_SM);
        PCS_KIND_SWITCH(kind, (grad2d_adj_kernel<T, VEC, KIND><<<grid, T2_BLOCK, 0, s>>>(out, z, (int)H, (int)W, edge, ih0, ih1)));
    } else {
        const Tiles tl = tiles_of<1>((int)H, (int)W);
        const int grid = (int)std::min<long long>(tl.n, (long long)num_sms() * T2_GRID_PER_SM);
        PCS_KIND_SWITCH(kind, (grad2d_adj_kernel<T, 1, KIND><<<grid, T2_BLOCK, 0, s>>>(out, z, (int)H, (int)W, edge, ih0, ih1)));
    }
    return check_launch("pcs_grad2d_adj");
}
PCS_EXPORT_REAL(pcs_grad2d_adj,
                (void* out, const void* z, int64_t H, int64_t W, int kind, int edge, double step0, double step1,
                 void* stream),
                ((T*)out, (const T*)z, H, W, kind, edge, step0, step1, (cudaStream_t)stream))

// ------------------------------------------------------------------------------------------------ 2-D Laplacian
// w0 * D2(axis0) + w1 * D2(axis1)  (pylops Laplacian; pycsou/linop/diff.py:946-957).  Not symmetric at the edges
// when edge=True, hence a distinct adjoint stencil.
template <typename T, bool ADJ>
__global__ void __launch_bounds__(DF_BLOCK) lap2d_kernel(T* __restrict__ out, const T* __restrict__ x, int64_t H,
                                                         int64_t W, int edge, T w0ih, T w1ih) {
    const int64_t N = H * W, stride = (int64_t)gridDim.x * DF_BLOCK;
    for (int64_t f = (int64_t)blockIdx.x * DF_BLOCK + threadIdx.x; f < N; f += stride) {
        const int64_t r = f / W, c = f - r * W;
        T a = ADJ ? d2_adj<T>(x + f, r, H, W, edge, T(1)) : d2_fwd<T>(x + f, r, H, W, edge, T(1));
        T b = ADJ ? d2_adj<T>(x + f, c, W, 1, edge, T(1)) : d2_fwd<T>(x + f, c, W, 1, edge, T(1));
        out[f] = w0ih * a + w1ih * b;
    }
}

// Row-tiled Laplacian: the interior stencil (1, -2, 1) per axis runs on the tiles (rows / columns 0 and n-1 produce
// zero, i.e. pylops' edge=False rule; the transpose gathers with the transposed entries); with edge=True the three
// outermost rows / columns differ (one-sided second differences reach two samples inwards) and a frame kernel
// recomputes those few pixels with the scalar formulas of pcs_diff.cuh.
template <typename T> __device__ __forceinline__ Coef3<T> d2_row(int i, int n) {
    Coef3<T> r = {T(0), T(0), T(0)};
    if (i >= 1 && i <= n - 2) { r.m = T(1); r.c = T(-2); r.p = T(1); }
    return r;
}
template <typename T, bool ADJ> __device__ __forceinline__ Coef3<T> d2_coef(int i, int n) {
    if (!ADJ) return d2_row<T>(i, n);
    Coef3<T> r;
    r.m = d2_row<T>(i - 1, n).p;
    r.c = d2_row<T>(i, n).c;
    r.p = d2_row<T>(i + 1, n).m;
    return r;
}
template <typename T, int V, bool ADJ>
__global__ void __launch_bounds__(T2_BLOCK) lap2d_tiled_kernel(T* __restrict__ out, const T* __restrict__ x, int H, int W,
                                                               T w0ih, T w1ih) {
    const Tiles tl = tiles_of<V>(H, W);
    const int lane = threadIdx.x & 31;
    for (long long t = blockIdx.x; t < tl.n; t += gridDim.x) {
        const int ty = (int)(t / tl.ntx), tx = (int)(t - (long long)ty * tl.ntx);
        const int j0 = (tx * T2_BLOCK + (int)threadIdx.x) * V;
        const int r0 = ty * T2_ROWS, r1 = min(H, r0 + T2_ROWS);
        ColCoef<T, V> kc;
#pragma unroll
        for (int e = 0; e < V; ++e) kc.k[e] = d2_coef<T, ADJ>(j0 + e, W);
        Pack<T, V> m = row_load<T, V>(x, r0 - 1, H, W, j0);
        Pack<T, V> c = row_load<T, V>(x, r0, H, W, j0);
        for (int i0 = r0; i0 < r1; i0 += GRAD_GROUP) {
            Pack<T, V> nx[GRAD_GROUP];
#pragma unroll
            for (int g = 0; g < GRAD_GROUP; ++g) nx[g] = row_load<T, V>(x, i0 + g + 1, H, W, j0);
#pragma unroll
            for (int g = 0; g < GRAD_GROUP; ++g) {
                const int i = i0 + g;
                if (i < r1) {
                    const Pack<T, V> a = row_stencil<T, V>(m, c, nx[g], d2_coef<T, ADJ>(i, H));
                    T l, r;
                    row_neighbours<T, V>(x + (int64_t)i * W, c, j0, W, lane, l, r);
                    const Pack<T, V> b = col_stencil<T, V>(c, l, r, kc);
                    Pack<T, V> o;
#pragma unroll
                    for (int e = 0; e < V; ++e) o.v[e] = w0ih * a.v[e] + w1ih * b.v[e];
                    pack_store<T, V>(out + (int64_t)i * W, j0, W, o);
                }
                m = c;
                c = nx[g];
            }
        }
    }
}
// edge=True: the pixels within 3 samples of an image edge, exactly as the scalar kernel computes them
template <typename T, bool ADJ>
__global__ void __launch_bounds__(DF_BLOCK) lap2d_frame_kernel(T* __restrict__ out, const T* __restrict__ x, int64_t H,
                                                               int64_t W, T w0ih, T w1ih) {
    const int64_t nrow = 6 * W, total = nrow + 6 * H, stride = (int64_t)gridDim.x * DF_BLOCK;
    for (int64_t q = (int64_t)blockIdx.x * DF_BLOCK + threadIdx.x; q < total; q += stride) {
        int64_t r, c;
        if (q < nrow) {
            const int64_t k = q / W;
            c = q - k * W;
            r = k < 3 ? k : H - 6 + k;
        } else {
            const int64_t k = (q - nrow) / H;
            r = (q - nrow) - k * H;
            c = k < 3 ? k : W - 6 + k;
        }
        if (r < 0 || r >= H || c < 0 || c >= W) continue;
        const int64_t f = r * W + c;
        const T a = ADJ ? d2_adj<T>(x + f, r, H, W, 1, T(1)) : d2_fwd<T>(x + f, r, H, W, 1, T(1));
        const T b = ADJ ? d2_adj<T>(x + f, c, W, 1, 1, T(1)) : d2_fwd<T>(x + f, c, W, 1, 1, T(1));
        out[f] = w0ih * a + w1ih * b;
    }
}

template <typename T, bool ADJ>
static int lap2d_launch(T* out, const T* x, int64_t H, int64_t W, double w0, double w1, double s0, double s1, int edge,
                        cudaStream_t s, const char* what) {
    PCS_REQUIRE(H >= 0 && W >= 0 && s0 != 0.0 && s1 != 0.0, "%s: bad dimensions or zero step", what);
    if (H * W == 0) return PCS_OK;
    PCS_REQUIRE(out && x && out != x, "%s: null or aliased pointer", what);
    PCS_REQUIRE(!edge || (H >= 3 && W >= 3), "%s: edge=True needs >= 3 samples per axis", what);
    const T a0 = (T)(w0 / (s0 * s0)), a1 = (T)(w1 / (s1 * s1));
    if (H < (1ll << 31) - 64 && W < (1ll << 31) - 4096 && (!edge || (H >= 6 && W >= 6))) {
        constexpr int VEC = 16 / (int)sizeof(T);
        if (t2_vec_ok<T>(W, {out, x})) {
            const Tiles tl = tiles_of<VEC>((int)H, (int)W);
            const int grid = (int)std::min<long long>(tl.n, (long long)num_sms() * T2_GRID_PER_SM);
            lap2d_tiled_kernel<T, VEC, ADJ><<<grid, T2_BLOCK, 0, s>>>(out, x, (int)H, (int)W, a0, a1);
        } else {
            const Tiles tl = tiles_of<1>((int)H, (int)W);
            const int grid = (int)std::min<long long>(tl.n, (long long)num_sms() * T2_GRID_PER_SM);
            lap2d_tiled_kernel<T, 1, ADJ><<<grid, T2_BLOCK, 0, s>>>(out, x, (int)H, (int)W, a0, a1);
        }
        if (edge) lap2d_frame_kernel<T, ADJ><<<grid_for(6 * (H + W), DF_BLOCK, DF_CTAS_PER_SM), DF_BLOCK, 0, s>>>(out, x, H, W, a0, a1);
        return check_launch(what);
    }
    lap2d_kernel<T, ADJ><<<grid_for(H * W, DF_BLOCK, DF_CTAS_PER_SM), DF_BLOCK, 0, s>>>(out, x, H, W, edge, a0, a1);
    return check_launch(what);
}
#define LAP_PARAMS                                                                                              \
    (void* out, const void* x, int64_t H, int64_t W, double w0, double w1, double step0, double step1, int edge, \
     void* stream)
#define LAP_ARGS ((T*)out, (const T*)x, H, W, w0, w1, step0, step1, edge, (cudaStream_t)stream)
template <typename T>
static int pcs_laplacian2d_fwd_impl(T* o, const T* x, int64_t H, int64_t W, double w0, double w1, double s0, double s1,
                                    int e, cudaStream_t s) {
    return lap2d_launch<T, false>(o, x, H, W, w0, w1, s0, s1, e, s, "pcs_laplacian2d_fwd");
}
PCS_EXPORT_REAL(pcs_laplacian2d_fwd, LAP_PARAMS, LAP_ARGS)
template <typename T>
static int pcs_laplacian2d_adj_impl(T* o, const T* x, int64_t H, int64_t W, double w0, double w1, double s0, double s1,
                                    int e, cudaStream_t s) {
    return lap2d_launch<T, true>(o, x, H, W, w0, w1, s0, s1, e, s, "pcs_laplacian2d_adj");
}
PCS_EXPORT_REAL(pcs_laplacian2d_adj, LAP_PARAMS, LAP_ARGS)

// ---- Integration1D: pylops CausalIntegration(halfcurrent=False) behind pycsou/linop/diff.py:1071-1137 ------------------
// forward y[i] = step * sum_{j <= i} x[j] along the axis; adjoint y[i] = step * sum_{j >= i} x[j] (the scan run backwards).
// Lines with inner > 1 (the axis is not the fastest one): one thread per line, coalesced across `inner`.  Contiguous
// lines (inner == 1): one block per line, chunks of 1024 samples scanned with warp shuffles and a running carry.
namespace {
template <typename T, bool REV>
__global__ void __launch_bounds__(256) cumsum_strided_kernel(T* __restrict__ out, const T* __restrict__ x, int64_t outer,
                                                             int64_t n, int64_t inner, T step) {
    const int64_t lines = outer * inner;
    for (int64_t l = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; l < lines; l += (int64_t)gridDim.x * blockDim.x) {
        const int64_t o = l / inner, i = l - o * inner;
        const int64_t base = o * n * inner + i;
        T acc = T(0);
        for (int64_t k = 0; k < n; ++k) {
            const int64_t kk = REV ? n - 1 - k : k;
            acc += x[base + kk * inner];
            out[base + kk * inner] = step * acc;
        }
    }
}
template <typename T, bool REV>
__global__ void __launch_bounds__(256) cumsum_line_kernel(T* __restrict__ out, const T* __restrict__ x, int64_t n, T step) {
    constexpr int ITEMS = 4, NT = 256, CHUNK = ITEMS * NT;
    __shared__ T warp_tot[NT / 32];
    __shared__ T s_carry;
    const T* xl = x + (int64_t)blockIdx.x * n;
    T* ol = out + (int64_t)blockIdx.x * n;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = T(0);
    __syncthreads();
    for (int64_t c0 = 0; c0 < n; c0 += CHUNK) {
        T v[ITEMS];
        T run = T(0);
#pragma unroll
        for (int e = 0; e < ITEMS; ++e) {
            const int64_t k = c0 + (int64_t)tid * ITEMS + e;
            const T xv = k < n ? xl[REV ? n - 1 - k : k] : T(0);
            run += xv;
            v[e] = run;
        }
        T incl = run;                                             // inclusive scan of the thread totals inside the warp
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const T t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        T woff = T(0);
        for (int w = 0; w < warp; ++w) woff += warp_tot[w];
        const T carry = s_carry;
        const T off = carry + woff + (incl - run);
#pragma unroll
        for (int e = 0; e < ITEMS; ++e) {
            const int64_t k = c0 + (int64_t)tid * ITEMS + e;
            if (k < n) ol[REV ? n - 1 - k : k] = step * (off + v[e]);
        }
        __syncthreads();
        if (tid == NT - 1) s_carry = off + run;
        __syncthreads();
    }
}
template <typename T, bool REV>
int cumsum_launch(T* out, const T* x, int64_t outer, int64_t n, int64_t inner, double step, cudaStream_t s, const char* what) {
    PCS_REQUIRE(outer >= 0 && n >= 0 && inner >= 0 && outer < (1ll << 31), "%s: bad dimensions", what);
    if (outer * n * inner == 0) return PCS_OK;
    PCS_REQUIRE(out && x && out != x, "%s: null or aliased pointer", what);
    if (inner == 1 && n >= 2048) {
        cumsum_line_kernel<T, REV><<<(unsigned)outer, 256, 0, s>>>(out, x, n, (T)step);
    } else {
        const int64_t lines = outer * inner;
        cumsum_strided_kernel<T, REV><<<grid_for(lines, 256, 8), 256, 0, s>>>(out, x, outer, n, inner, (T)step);
    }
    return check_launch(what);
}
}  // namespace
#define CUMSUM_PARAMS (void* out, const void* x, int64_t outer, int64_t n, int64_t inner, double step, void* stream)
template <typename T>
static int pcs_cumsum_fwd_impl(T* o, const T* x, int64_t outer, int64_t n, int64_t inner, double step, cudaStream_t s) {
    return cumsum_launch<T, false>(o, x, outer, n, inner, step, s, "pcs_cumsum_fwd");
}
PCS_EXPORT_REAL(pcs_cumsum_fwd, CUMSUM_PARAMS, ((T*)out, (const T*)x, outer, n, inner, step, (cudaStream_t)stream))
template <typename T>
static int pcs_cumsum_adj_impl(T* o, const T* x, int64_t outer, int64_t n, int64_t inner, double step, cudaStream_t s) {
    return cumsum_launch<T, true>(o, x, outer, n, inner, step, s, "pcs_cumsum_adj");
}
PCS_EXPORT_REAL(pcs_cumsum_adj, CUMSUM_PARAMS, ((T*)out, (const T*)x, outer, n, inner, step, (cudaStream_t)stream))
