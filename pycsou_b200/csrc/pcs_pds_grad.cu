// libpcsb200: Condat-Vu (PDS) iteration for total-variation problems with ANY 2-D Gradient -- forward, backward or pycsou's
// default `centered` differences, with or without pylops' edge rows -- in TWO kernels per update (sm_100a).
//
// The one-kernel update of pcs_pds.cu is built around the forward difference (its dual update reads u one row / column
// ahead only).  Centred and backward differences reach one row / column BEHIND as well, so they take this path instead
// of falling back to ~15 stand-alone launches (PrimalDualSplitting.update_iterand + update_diagnostics,
// pycsou/opt/proxalgs.py:343-394, with K = Gradient(shape, kind=...), pycsou/linop/diff.py:777-882):
//
//   primal kernel:  a = K^T z,   xt = prox_{tau G}(x - tau g - tau a),   u = 2 xt - x,   x' = rho xt + (1 - rho) x,
//                   sums ||x' - x||^2, ||x||^2                      g = 0 | x - y | caller-supplied gradient of F
//   dual kernel:    zt = fenchel_prox_{sigma H}(z + sigma K u),   z' = rho zt + (1 - rho) z,   sums ||z' - z||^2, ||z||^2
//
// Both are row-tiled stencil kernels (pcs_tile2d.cuh: 128-bit coalesced accesses, three-row register window, lane
// shuffles for the column neighbours).  HBM traffic: primal reads x, g-term, z0, z1 and writes x', u (6 N s; 5 N s
// without a data term); dual reads u, z0, z1 and writes z0', z1' (5 N s): 11 N s against the 7 N s of the one-kernel
// update -- u makes one round trip through HBM / L2.
#include "pcs_tile2d.cuh"
#include <algorithm>

using namespace pcs;

namespace {

enum { DM_NONE = 0, DM_IDENTITY = 1, DM_GIVEN = 3 };
enum { GM_NONE = 0, GM_NONNEG = 1, GM_SEGMENT = 2, GM_L1 = 3 };
enum { HM_L21 = 0, HM_L1 = 1 };
constexpr int PG_MAX_BLOCKS = 1 << 16;       // partial-sum slots of the Condat-Vu workspace (pcs_pds_tv_workspace_bytes)

// prox of G evaluated as the reference does (math/prox.py:295-297, 340-343; func/base.py:239-240)
template <typename T> __device__ __forceinline__ T prox_g(T v, int mode, T ga, T gb) {
    if (mode == GM_NONNEG) return v < T(0) ? T(0) : v;
    if (mode == GM_SEGMENT) {
        const T r = v < ga ? ga : v;
        return r > gb ? gb : r;
    }
    if (mode == GM_L1) return prox_l1<T>(v, ga);
    return v;
}
template <typename T> __device__ __forceinline__ T trsq(T v);
template <> __device__ __forceinline__ float trsq<float>(float v) { return rsqrtf(v); }
template <> __device__ __forceinline__ double trsq<double>(double v) { return rsqrt(v); }
template <typename T> __device__ __forceinline__ T tiny_pos();
template <> __device__ __forceinline__ float tiny_pos<float>() { return 1e-30f; }
template <> __device__ __forceinline__ double tiny_pos<double>() { return 1e-300; }

template <typename T> struct PrimalArgs {
    T* x_out; T* u_out; const T* x; const T* b; const T* z;
    int H, W, kind, edge, data_mode, g_mode;
    T ih0, ih1, tau, rho, omr, ga, gb;
    double* partials; unsigned int* ticket; double* diag_out;
};

template <typename T, int V, int KIND>
__global__ void __launch_bounds__(T2_BLOCK) pds_grad_primal_kernel(const PrimalArgs<T> A) {
    __shared__ double red[2 * (T2_BLOCK / 32)];
    const int H = A.H, W = A.W;
    const Tiles tl = tiles_of<V>(H, W);
    const int lane = threadIdx.x & 31;
    const T* __restrict__ z0 = A.z;
    const T* __restrict__ z1 = A.z + (int64_t)H * W;
    double s_dx = 0.0, s_x = 0.0;
    for (long long t = blockIdx.x; t < tl.n; t += gridDim.x) {
        const int ty = (int)(t / tl.ntx), tx = (int)(t - (long long)ty * tl.ntx);
        const int j0 = (tx * T2_BLOCK + (int)threadIdx.x) * V;
        const int r0 = ty * T2_ROWS, r1 = min(H, r0 + T2_ROWS);
        const ColCoef<T, V> kc = col_coef<T, V, KIND, true>(A.edge, j0, W, A.ih1);
        Pack<T, V> m = row_load<T, V>(z0, r0 - 1, H, W, j0);
        Pack<T, V> c = row_load<T, V>(z0, r0, H, W, j0);
        T p_dx = T(0), p_x = T(0);
        for (int i0 = r0; i0 < r1; i0 += T2_GROUP) {
            // all loads of the group first: 4 x (z0, z1, x, data term) independent 128-bit loads in flight per thread
            Pack<T, V> nx[T2_GROUP], c1[T2_GROUP], xv[T2_GROUP], bv[T2_GROUP];
#pragma unroll
            for (int g = 0; g < T2_GROUP; ++g) {
                nx[g] = row_load<T, V>(z0, i0 + g + 1, H, W, j0);
                c1[g] = row_load<T, V>(z1, i0 + g, H, W, j0);
                xv[g] = row_load<T, V>(A.x, i0 + g, H, W, j0);
                bv[g] = A.data_mode != DM_NONE ? row_load<T, V>(A.b, i0 + g, H, W, j0) : pack_zero<T, V>();
            }
#pragma unroll
            for (int g = 0; g < T2_GROUP; ++g) {
                const int i = i0 + g;
                if (i < r1) {                                   // (uniform over the block: the shuffles below are safe)
                    const int64_t ro = (int64_t)i * W;
                    const Pack<T, V> a0 = row_stencil<T, V>(m, c, nx[g], d1_row_adj<T, KIND>(A.edge, i, H, A.ih0));
                    T l, r;
                    row_neighbours<T, V>(z1 + ro, c1[g], j0, W, lane, l, r);
                    const Pack<T, V> a1 = col_stencil<T, V>(c1[g], l, r, kc);
                    Pack<T, V> xn, un;
#pragma unroll
                    for (int e = 0; e < V; ++e) {
                        const T x = xv[g].v[e];
                        const T gg = A.data_mode == DM_NONE ? T(0) : (A.data_mode == DM_IDENTITY ? x - bv[g].v[e] : bv[g].v[e]);
                        const T v = fma(-A.tau, gg + (a0.v[e] + a1.v[e]), x);
                        const T xt = prox_g<T>(v, A.g_mode, A.ga, A.gb);
                        un.v[e] = fma(T(2), xt, -x);
                        xn.v[e] = fma(A.rho, xt, A.omr * x);
                        if (j0 + e < W) {
                            const T d = xn.v[e] - x;
                            p_dx = fma(d, d, p_dx);
                            p_x = fma(x, x, p_x);
                        }
                    }
                    pack_store<T, V>(A.x_out + ro, j0, W, xn);
                    pack_store<T, V>(A.u_out + ro, j0, W, un);
                }
                m = c;
                c = nx[g];
            }
        }
        s_dx += (double)p_dx;
        s_x += (double)p_x;
    }
    double v[2] = {s_dx, s_x};
    grid_sum_finish<2, T2_BLOCK>(v, A.partials, A.ticket, A.diag_out, red);
}

template <typename T> struct DualArgs {
    T* z_out; const T* u; const T* z;
    int H, W, kind, edge, h_mode;
    T ih0, ih1, sigma, rho, omr, lam;
    double* partials; unsigned int* ticket; double* diag_out;
};

template <typename T, int V, int KIND>
__global__ void __launch_bounds__(T2_BLOCK) pds_grad_dual_kernel(const DualArgs<T> A) {
    __shared__ double red[2 * (T2_BLOCK / 32)];
    const int H = A.H, W = A.W;
    const Tiles tl = tiles_of<V>(H, W);
    const int lane = threadIdx.x & 31;
    const int64_t plane = (int64_t)H * W;
    double s_dz = 0.0, s_z = 0.0;
    for (long long t = blockIdx.x; t < tl.n; t += gridDim.x) {
        const int ty = (int)(t / tl.ntx), tx = (int)(t - (long long)ty * tl.ntx);
        const int j0 = (tx * T2_BLOCK + (int)threadIdx.x) * V;
        const int r0 = ty * T2_ROWS, r1 = min(H, r0 + T2_ROWS);
        const ColCoef<T, V> kc = col_coef<T, V, KIND, false>(A.edge, j0, W, A.ih1);
        Pack<T, V> m = row_load<T, V>(A.u, r0 - 1, H, W, j0);
        Pack<T, V> c = row_load<T, V>(A.u, r0, H, W, j0);
        T p_dz = T(0), p_z = T(0);
        for (int i0 = r0; i0 < r1; i0 += T2_GROUP) {
            Pack<T, V> nx[T2_GROUP], z0v[T2_GROUP], z1v[T2_GROUP];
#pragma unroll
            for (int g = 0; g < T2_GROUP; ++g) {
                nx[g] = row_load<T, V>(A.u, i0 + g + 1, H, W, j0);
                z0v[g] = row_load<T, V>(A.z, i0 + g, H, W, j0);
                z1v[g] = row_load<T, V>(A.z + plane, i0 + g, H, W, j0);
            }
#pragma unroll
            for (int g = 0; g < T2_GROUP; ++g) {
                const int i = i0 + g;
                if (i < r1) {
                    const int64_t ro = (int64_t)i * W;
                    const Pack<T, V> k0 = row_stencil<T, V>(m, c, nx[g], d1_row<T, KIND>(A.edge, i, H, A.ih0));
                    T l, r;
                    row_neighbours<T, V>(A.u + ro, c, j0, W, lane, l, r);
                    const Pack<T, V> k1 = col_stencil<T, V>(c, l, r, kc);
                    Pack<T, V> n0, n1;
#pragma unroll
                    for (int e = 0; e < V; ++e) {
                        const T zz0 = z0v[g].v[e], zz1 = z1v[g].v[e];
                        const T w0 = fma(A.sigma, k0.v[e], zz0), w1 = fma(A.sigma, k1.v[e], zz1);
                        // fenchel_prox of lam*L21 (core/functional.py:207, func/penalty.py:551-557) = projection of w onto
                        // the lam-ball per pixel; of lam*L1 = clamp to [-lam, lam] per component (Moreau identity)
                        T t0, t1;
                        if (A.h_mode == HM_L21) {
                            const T ss = fma(w1, w1, fma(w0, w0, tiny_pos<T>()));
                            const T gsc = tmin<T>(A.lam * trsq<T>(ss), T(1));
                            t0 = w0 * gsc;
                            t1 = w1 * gsc;
                        } else {
                            t0 = tmin<T>(tmax<T>(w0, -A.lam), A.lam);
                            t1 = tmin<T>(tmax<T>(w1, -A.lam), A.lam);
                        }
                        n0.v[e] = fma(A.rho, t0, A.omr * zz0);
                        n1.v[e] = fma(A.rho, t1, A.omr * zz1);
                        if (j0 + e < W) {
                            const T e0 = n0.v[e] - zz0, e1 = n1.v[e] - zz1;
                            p_dz = fma(e0, e0, fma(e1, e1, p_dz));
                            p_z = fma(zz0, zz0, fma(zz1, zz1, p_z));
                        }
                    }
                    pack_store<T, V>(A.z_out + ro, j0, W, n0);
                    pack_store<T, V>(A.z_out + plane + ro, j0, W, n1);
                }
                m = c;
                c = nx[g];
            }
        }
        s_dz += (double)p_dz;
        s_z += (double)p_z;
    }
    double v[2] = {s_dz, s_z};
    grid_sum_finish<2, T2_BLOCK>(v, A.partials, A.ticket, A.diag_out, red);
}

template <int V> int pg_grid(int H, int W) {
    const Tiles tl = tiles_of<V>(H, W);
    long long g = std::min<long long>(tl.n, (long long)num_sms() * T2_GRID_PER_SM);
    g = std::min<long long>(g, PG_MAX_BLOCKS / 2);
    return (int)(g < 1 ? 1 : g);
}

int pg_check(const char* what, int64_t H, int64_t W, int kind, int edge, double s0, double s1) {
    PCS_REQUIRE(H >= 1 && W >= 1 && H < (1ll << 31) - 64 && W < (1ll << 31) - 4096, "%s: bad image size", what);
    PCS_REQUIRE(kind >= 0 && kind <= 2, "%s: kind must be 0 (forward), 1 (centered) or 2 (backward)", what);
    PCS_REQUIRE(s0 != 0.0 && s1 != 0.0, "%s: zero step", what);
    PCS_REQUIRE(kind != 1 || !edge || (H >= 2 && W >= 2), "%s: centered+edge needs >= 2 samples per axis", what);
    return PCS_OK;
}

}  // namespace

template <typename T>
static int pcs_pds_grad_primal_step_impl(T* x_out, T* u_out, const T* x, const T* b, const T* z, int64_t H, int64_t W,
                                         int kind, int edge, int data_mode, int g_mode, double g_a, double g_b, double tau,
                                         double rho, double step0, double step1, double* diag_out, void* workspace,
                                         cudaStream_t s) {
    const char* what = "pcs_pds_grad_primal_step";
    const int rc = pg_check(what, H, W, kind, edge, step0, step1);
    if (rc) return rc;
    PCS_REQUIRE(x_out && u_out && x && z && diag_out && workspace, "%s: null pointer", what);
    PCS_REQUIRE(x_out != x && u_out != x && u_out != x_out, "%s: outputs must not alias the inputs or each other", what);
    PCS_REQUIRE(data_mode == DM_NONE || data_mode == DM_IDENTITY || data_mode == DM_GIVEN, "%s: bad data_mode %d", what, data_mode);
    PCS_REQUIRE(data_mode == DM_NONE || b, "%s: data term needs its array", what);
    PCS_REQUIRE(g_mode >= GM_NONE && g_mode <= GM_L1 && tau > 0, "%s: bad prox mode / step", what);
    PrimalArgs<T> A;
    A.x_out = x_out; A.u_out = u_out; A.x = x; A.b = b; A.z = z;
    A.H = (int)H; A.W = (int)W; A.kind = kind; A.edge = edge; A.data_mode = data_mode; A.g_mode = g_mode;
    A.ih0 = (T)(1.0 / step0); A.ih1 = (T)(1.0 / step1); A.tau = (T)tau; A.rho = (T)rho; A.omr = (T)(1 - rho);
    A.ga = (T)g_a; A.gb = (T)g_b;
    A.partials = (double*)workspace;
    A.ticket = (unsigned int*)((double*)workspace + 4 * (size_t)PG_MAX_BLOCKS);
    A.diag_out = diag_out;
    constexpr int VEC = 16 / (int)sizeof(T);
    if (t2_vec_ok<T>(W, {x_out, u_out, x, b, z})) {
        PCS_KIND_SWITCH(kind, (pds_grad_primal_kernel<T, VEC, KIND><<<pg_grid<VEC>((int)H, (int)W), T2_BLOCK, 0, s>>>(A)));
    } else {
        PCS_KIND_SWITCH(kind, (pds_grad_primal_kernel<T, 1, KIND><<<pg_grid<1>((int)H, (int)W), T2_BLOCK, 0, s>>>(A)));
    }
    return check_launch(what);
}
PCS_EXPORT_REAL(pcs_pds_grad_primal_step,
                (void* x_out, void* u_out, const void* x, const void* b, const void* z, int64_t H, int64_t W, int kind,
                 int edge, int data_mode, int g_mode, double g_a, double g_b, double tau, double rho, double step0,
                 double step1, void* diag_out_f64, void* workspace, void* stream),
                ((T*)x_out, (T*)u_out, (const T*)x, (const T*)b, (const T*)z, H, W, kind, edge, data_mode, g_mode, g_a, g_b,
                 tau, rho, step0, step1, (double*)diag_out_f64, workspace, (cudaStream_t)stream))

template <typename T>
static int pcs_pds_grad_dual_step_impl(T* z_out, const T* u, const T* z, int64_t H, int64_t W, int kind, int edge,
                                       int h_mode, double lam, double sigma, double rho, double step0, double step1,
                                       double* diag_out, void* workspace, cudaStream_t s) {
    const char* what = "pcs_pds_grad_dual_step";
    const int rc = pg_check(what, H, W, kind, edge, step0, step1);
    if (rc) return rc;
    PCS_REQUIRE(z_out && u && z && diag_out && workspace, "%s: null pointer", what);
    PCS_REQUIRE(z_out != z && (const T*)z_out != u, "%s: the output must not alias an input", what);
    PCS_REQUIRE((h_mode == HM_L21 || h_mode == HM_L1) && lam > 0 && sigma > 0, "%s: bad prox mode / step", what);
    DualArgs<T> A;
    A.z_out = z_out; A.u = u; A.z = z;
    A.H = (int)H; A.W = (int)W; A.kind = kind; A.edge = edge; A.h_mode = h_mode;
    A.ih0 = (T)(1.0 / step0); A.ih1 = (T)(1.0 / step1); A.sigma = (T)sigma; A.rho = (T)rho; A.omr = (T)(1 - rho);
    A.lam = (T)lam;
    A.partials = (double*)workspace;
    A.ticket = (unsigned int*)((double*)workspace + 4 * (size_t)PG_MAX_BLOCKS);
    A.diag_out = diag_out + 2;
    constexpr int VEC = 16 / (int)sizeof(T);
    if (t2_vec_ok<T>(W, {z_out, u, z})) {
        PCS_KIND_SWITCH(kind, (pds_grad_dual_kernel<T, VEC, KIND><<<pg_grid<VEC>((int)H, (int)W), T2_BLOCK, 0, s>>>(A)));
    } else {
        PCS_KIND_SWITCH(kind, (pds_grad_dual_kernel<T, 1, KIND><<<pg_grid<1>((int)H, (int)W), T2_BLOCK, 0, s>>>(A)));
    }
    return check_launch(what);
}
PCS_EXPORT_REAL(pcs_pds_grad_dual_step,
                (void* z_out, const void* u, const void* z, int64_t H, int64_t W, int kind, int edge, int h_mode, double lam,
                 double sigma, double rho, double step0, double step1, void* diag_out_f64, void* workspace, void* stream),
                ((T*)z_out, (const T*)u, (const T*)z, H, W, kind, edge, h_mode, lam, sigma, rho, step0, step1,
                 (double*)diag_out_f64, workspace, (cudaStream_t)stream))
