// libpcsb200: fused Chambolle-Pock iteration for compressed-sensing MRI with a TV penalty (complex data, sm_100a).
//
// BASELINE.json configs[3] / SURVEY.md 8(d) "C4":  ChambollePockSplitting (= PrimalDualSplitting with F = None, rho = 1;
// pycsou/opt/proxalgs.py:628-713, update_iterand :343-355, update_diagnostics :357-394) on
//     K = LinOpVStack(Masking * DFT, Gradient(shape, kind='forward'))            (linop/base.py:159-314, sampling.py:125-196,
//                                                                                 linop/diff.py:777-882)
//     H = ProxFuncHStack(L2BallLoss(m, y, eps), lam * L21Norm(2N, TV groups))    (func/base.py:21-89, func/loss.py:69,
//                                                                                 func/penalty.py:134-191, 480-560)
//     G = None,  dual variable z = [u (m samples) ; w0 ; w1 (N pixels each)],  all complex.
// One iteration is
//     S  = Mask^T u                 (pcs_cps_mri_scatter: zero fill + scatter; only for the first update of a solve:
//                                    the ball step of one update leaves Mask^T u' behind for the next)  [cuFFT: g = DFT^H S]
//     xt = x - tau (g + Grad^T w),  v = 2 xt - x,  x' = rho xt + (1-rho) x
//     w' = rho fenchel_prox_{sigma lam L21}(w + sigma Grad v) + (1-rho) w                 (pcs_cps_mri_tv_step: one kernel,
//                                                                                          v through shared memory)
//                                                                                         [cuFFT: V = DFT v]
//     p  = u + sigma V[mask],  u' = rho (p - sigma (y + proj_ball_eps(p/sigma - y))) + (1-rho) u   (pcs_cps_mri_ball_step:
//                                                                                          norm pass + update pass over m)
// and the four sums ||x'-x||^2, ||x||^2, ||z'-z||^2, ||z||^2 of update_diagnostics (deterministic reductions).
// The two DFTs are library calls made by the host (cuFFT through torch.fft), exactly where the reference calls pocketfft.
// HBM traffic per iteration outside the FFTs: read x, g, w (4 N), write x', w', v (4 N), S (N + m) and 6 m for the ball step.
#include "pcs_common.cuh"
#include <math.h>
#include <stdlib.h>
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

using namespace pcs;

namespace {

template <typename R> struct Cx;
template <> struct Cx<float> { typedef float2 type; };
template <> struct Cx<double> { typedef double2 type; };

template <typename V, typename R> __device__ __forceinline__ V cmake(R a, R b) { V v; v.x = a; v.y = b; return v; }
template <typename V> __device__ __forceinline__ V cadd(V a, V b) { a.x += b.x; a.y += b.y; return a; }
template <typename V> __device__ __forceinline__ V csub(V a, V b) { a.x -= b.x; a.y -= b.y; return a; }
template <typename V, typename R> __device__ __forceinline__ V cscale(V a, R s) { a.x *= s; a.y *= s; return a; }
// a + s b
template <typename V, typename R> __device__ __forceinline__ V caxpy(V a, R s, V b) {
    a.x = fma(s, b.x, a.x); a.y = fma(s, b.y, a.y); return a;
}
template <typename V> __device__ __forceinline__ double cabs2(V a) { return (double)a.x * a.x + (double)a.y * a.y; }

template <typename R> __device__ __forceinline__ R rsq(R v);
template <> __device__ __forceinline__ float rsq<float>(float v) { return rsqrtf(v); }
template <> __device__ __forceinline__ double rsq<double>(double v) { return rsqrt(v); }
template <typename R> __device__ __forceinline__ R tiny();
template <> __device__ __forceinline__ float tiny<float>() { return 1e-30f; }
template <> __device__ __forceinline__ double tiny<double>() { return 1e-300; }

constexpr int CPS_TH = 32;                 // tile rows
constexpr int CPS_NT = 256;
constexpr int CPS_MAX_BLOCKS = 4096;       // partial-sum slots in the workspace
template <typename R> struct TileW { static constexpr int value = sizeof(R) == 4 ? 128 : 64; };   // tile columns

template <typename R> struct CpsTvArgs {
    typedef typename Cx<R>::type V;
    V* x_out; V* w_out; V* v_out;
    const V* x; const V* w; const V* g;
    int H, W;
    R lam, tau, sigma, rho, omr, is0, is1, gscale;
    double* partials; unsigned int* ticket; double* diag_out;
};

// ---- row-marching form of the TV step (the default; the tile kernel below stays for PCS_CPS_TV_TILE=1) -------------------
// The tile kernel reads w0 and w1 three times per pixel (twice for Grad^T w, once for the dual update), stages v through
// shared memory with two block barriers per tile and computes point indices by division: 65 us for 268 MB at 2048^2
// complex64.  Here a thread owns NC = 16 / sizeof(complex) adjacent columns and MARCHES down a segment of rows: x, g, w0,
// w1 of a row are loaded once (16-byte coalesced loads, the next row's in flight while this one is worked on), w0 of the
// row above and v of the row above stay in registers, the left neighbour's w1 and the right neighbour's v come from the
// adjacent lanes by warp shuffles.  A warp covers 32 NC columns and owns 31 NC of them: its last lane only computes v
// (needed as "right neighbour" by lane 30) -- no shared memory, no block barrier.  Per pixel the operations are those of
// the tile kernel, in the same order.
constexpr int CPM_NT = 128;                 // threads per block (4 warps: 124 NC owned columns)
#ifndef CPM_MINB
#define CPM_MINB 6
#endif
template <typename R, int NC> struct CpmRow { typename Cx<R>::type x[NC], g[NC], w0[NC], w1[NC]; };

template <typename R, int NC>
__device__ __forceinline__ void cpm_load_row(CpmRow<R, NC>& r, const CpsTvArgs<R>& A, int gi, int gj0, bool colok, int64_t N) {
    typedef typename Cx<R>::type V;
    const V zero = cmake<V, R>(R(0), R(0));
    if (colok && gi < A.H) {
        const int64_t idx = (int64_t)gi * A.W + gj0;
        if (NC == 2) {
            const float4 a = *reinterpret_cast<const float4*>(A.x + idx), b = *reinterpret_cast<const float4*>(A.g + idx);
            const float4 c = *reinterpret_cast<const float4*>(A.w + idx), d = *reinterpret_cast<const float4*>(A.w + N + idx);
            r.x[0] = cmake<V, R>((R)a.x, (R)a.y); r.x[NC - 1] = cmake<V, R>((R)a.z, (R)a.w);
            r.g[0] = cmake<V, R>((R)b.x, (R)b.y); r.g[NC - 1] = cmake<V, R>((R)b.z, (R)b.w);
            r.w0[0] = cmake<V, R>((R)c.x, (R)c.y); r.w0[NC - 1] = cmake<V, R>((R)c.z, (R)c.w);
            r.w1[0] = cmake<V, R>((R)d.x, (R)d.y); r.w1[NC - 1] = cmake<V, R>((R)d.z, (R)d.w);
        } else {
            r.x[0] = A.x[idx]; r.g[0] = A.g[idx]; r.w0[0] = A.w[idx]; r.w1[0] = A.w[N + idx];
        }
    } else {
#pragma unroll
        for (int e = 0; e < NC; ++e) r.x[e] = r.g[e] = r.w0[e] = r.w1[e] = zero;
    }
}

template <typename R>
__global__ void __launch_bounds__(CPM_NT, CPM_MINB) cps_tv_march_kernel(const CpsTvArgs<R> A, int seg_rows) {
    typedef typename Cx<R>::type V;
    constexpr int NC = 16 / (int)sizeof(V);
    __shared__ double red[4 * (CPM_NT / 32)];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int H = A.H, W = A.W;
    const int64_t N = (int64_t)H * W;
    const V zero = cmake<V, R>(R(0), R(0));
    // columns: warp `gw` of the grid row covers [31 NC gw, + 32 NC); lane 31 is the redundant one
    const int gw = (int)blockIdx.x * (CPM_NT / 32) + warp;
    const int gj0 = gw * 31 * NC + lane * NC;
    const bool colok = gj0 < W;                         // (W is a multiple of NC: all NC columns or none)
    const bool owner = colok && lane < 31;
    const int i0 = (int)blockIdx.y * seg_rows, i1 = min(H, i0 + seg_rows);
    double s_dx = 0.0, s_x = 0.0, s_dw = 0.0, s_w = 0.0;
    if (i0 < i1) {
        V w0_up[NC], a0p[NC], a1p[NC], v_prev[NC];      // row i - 1: w0 (for Grad^T w of row i and the dual update), w1, v
#pragma unroll
        for (int e = 0; e < NC; ++e) {
            w0_up[e] = (colok && i0 > 0) ? A.w[(int64_t)(i0 - 1) * W + gj0 + e] : zero;
            a0p[e] = a1p[e] = v_prev[e] = zero;
        }
        CpmRow<R, NC> cur, nxt;
        cpm_load_row<R, NC>(cur, A, i0, gj0, colok, N);
        for (int i = i0; i <= i1; ++i) {                 // row i1 (if it exists): v only, for the dual update of row i1 - 1
            if (i < i1) cpm_load_row<R, NC>(nxt, A, i + 1, gj0, colok && i + 1 <= i1, N);
            // left neighbour's w1 (column gj0 - 1)
            V w1l;
            w1l.x = __shfl_up_sync(0xffffffffu, cur.w1[NC - 1].x, 1);
            w1l.y = __shfl_up_sync(0xffffffffu, cur.w1[NC - 1].y, 1);
            if (lane == 0) w1l = (colok && gj0 > 0 && i < H) ? A.w[N + (int64_t)i * W + gj0 - 1] : zero;
            V v[NC];
            const bool rowok = i < H;
#pragma unroll
            for (int e = 0; e < NC; ++e) {
                const int gj = gj0 + e;
                v[e] = zero;
                if (colok && rowok) {
                    const V x = cur.x[e], g = cscale(cur.g[e], A.gscale);
                    const V w0c = (i < H - 1) ? cur.w0[e] : zero, w0u = (i > 0) ? w0_up[e] : zero;
                    const V w1c = (gj < W - 1) ? cur.w1[e] : zero;
                    const V w1lv = (gj > 0) ? (e == 0 ? w1l : cur.w1[e == 0 ? 0 : e - 1]) : zero;
                    V a = cscale(csub(w0u, w0c), A.is0);
                    a = caxpy(a, A.is1, csub(w1lv, w1c));
                    const V xt = caxpy(x, -A.tau, cadd(g, a));
                    v[e] = caxpy(cscale(x, R(-1)), R(2), xt);
                    if (owner && i < i1) {
                        const V xn = caxpy(cscale(x, A.omr), A.rho, xt);
                        cur.x[e] = xn;                      // (kept for the vector store below)
                        s_dx += cabs2(csub(xn, x));
                        s_x += cabs2(x);
                    }
                }
            }
            if (owner && i < i1) {
                const int64_t idx = (int64_t)i * W + gj0;
                if (NC == 2) {
                    *reinterpret_cast<float4*>(A.x_out + idx) = make_float4((float)cur.x[0].x, (float)cur.x[0].y, (float)cur.x[NC - 1].x, (float)cur.x[NC - 1].y);
                    *reinterpret_cast<float4*>(A.v_out + idx) = make_float4((float)v[0].x, (float)v[0].y, (float)v[NC - 1].x, (float)v[NC - 1].y);
                } else {
                    A.x_out[idx] = cur.x[0];
                    A.v_out[idx] = v[0];
                }
            }
            // dual update of row i - 1: needs v of rows i - 1 (v_prev) and i (v), and v_prev one column to the right
            V vr_last;
            vr_last.x = __shfl_down_sync(0xffffffffu, v_prev[0].x, 1);
            vr_last.y = __shfl_down_sync(0xffffffffu, v_prev[0].y, 1);
            if (i > i0 && owner) {
                const int gi = i - 1;
                V n0[NC], n1[NC];
#pragma unroll
                for (int e = 0; e < NC; ++e) {
                    const int gj = gj0 + e;
                    const V vc = v_prev[e], vd = v[e], vr = (e == NC - 1) ? vr_last : v_prev[e == NC - 1 ? e : e + 1];
                    const V d0 = gi < H - 1 ? cscale(csub(vd, vc), A.is0) : zero;
                    const V d1 = gj < W - 1 ? cscale(csub(vr, vc), A.is1) : zero;
                    const V q0 = caxpy(a0p[e], A.sigma, d0), q1 = caxpy(a1p[e], A.sigma, d1);
                    const R ss = fma(q0.x, q0.x, fma(q0.y, q0.y, fma(q1.x, q1.x, fma(q1.y, q1.y, tiny<R>()))));
                    R f = A.lam * rsq<R>(ss);
                    f = f < R(1) ? f : R(1);
                    n0[e] = caxpy(cscale(a0p[e], A.omr), A.rho * f, q0);
                    n1[e] = caxpy(cscale(a1p[e], A.omr), A.rho * f, q1);
                    s_dw += cabs2(csub(n0[e], a0p[e])) + cabs2(csub(n1[e], a1p[e]));
                    s_w += cabs2(a0p[e]) + cabs2(a1p[e]);
                }
                const int64_t idx = (int64_t)gi * W + gj0;
                if (NC == 2) {
                    *reinterpret_cast<float4*>(A.w_out + idx) = make_float4((float)n0[0].x, (float)n0[0].y, (float)n0[NC - 1].x, (float)n0[NC - 1].y);
                    *reinterpret_cast<float4*>(A.w_out + N + idx) = make_float4((float)n1[0].x, (float)n1[0].y, (float)n1[NC - 1].x, (float)n1[NC - 1].y);
                } else {
                    A.w_out[idx] = n0[0];
                    A.w_out[N + idx] = n1[0];
                }
            }
#pragma unroll
            for (int e = 0; e < NC; ++e) { w0_up[e] = cur.w0[e]; a0p[e] = cur.w0[e]; a1p[e] = cur.w1[e]; v_prev[e] = v[e]; }
            cur = nxt;
        }
    }
    double v4[4] = {s_dx, s_x, s_dw, s_w};
    grid_sum_finish<4, CPM_NT>(v4, A.partials, A.ticket, A.diag_out, red);
}

template <typename R>
__global__ void __launch_bounds__(CPS_NT) cps_tv_kernel(const CpsTvArgs<R> A) {
    typedef typename Cx<R>::type V;
    constexpr int TH = CPS_TH, TW = TileW<R>::value, PW = TW + 1;
    extern __shared__ __align__(16) unsigned char smem[];
    V* const sv = reinterpret_cast<V*>(smem);                                   // v on the (TH+1) x (TW+1) points of the tile
    double* const red = reinterpret_cast<double*>(smem + (size_t)(TH + 1) * PW * sizeof(V));
    const int tid = threadIdx.x, H = A.H, W = A.W;
    const int tiles_w = (W + TW - 1) / TW, tiles_h = (H + TH - 1) / TH;
    const int64_t N = (int64_t)H * W;
    const V* __restrict__ w0 = A.w;
    const V* __restrict__ w1 = A.w + N;
    double s_dx = 0.0, s_x = 0.0, s_dw = 0.0, s_w = 0.0;
    const V zero = cmake<V, R>(R(0), R(0));

    const V* __restrict__ xin = A.x;
    const V* __restrict__ gin = A.g;
    V* __restrict__ xout = A.x_out;
    V* __restrict__ vout = A.v_out;
    V* __restrict__ w0out = A.w_out;
    V* __restrict__ w1out = A.w_out + N;
    constexpr int UN = sizeof(R) == 4 ? 4 : 2;   // points per thread and batch: all loads of a batch are issued before its arithmetic

    for (int tile = blockIdx.x; tile < tiles_w * tiles_h; tile += gridDim.x) {
        const int ti = tile / tiles_w, tj = tile - ti * tiles_w;
        const int i0 = ti * TH, j0 = tj * TW;
        // primal update on the tile plus one row below / one column to the right (the forward differences of v need them)
        for (int pb = tid; pb < (TH + 1) * PW; pb += UN * CPS_NT) {
            V xv[UN], gv[UN], w0c[UN], w0u[UN], w1c[UN], w1l[UN];
            bool in[UN];
#pragma unroll
            for (int k = 0; k < UN; ++k) {
                const int p = pb + k * CPS_NT;
                const int r = p / PW, c = p - r * PW;
                const int gi = i0 + r, gj = j0 + c;
                in[k] = p < (TH + 1) * PW && gi < H && gj < W;
                const int64_t idx = in[k] ? (int64_t)gi * W + gj : 0;
                xv[k] = in[k] ? xin[idx] : zero;
                gv[k] = in[k] ? gin[idx] : zero;
                // Gradient^T w (pylops FirstDerivative forward, adjoint: y[:-1] -= z[:-1]/h; y[1:] += z[:-1]/h)
                w0c[k] = (in[k] && gi < H - 1) ? w0[idx] : zero;
                w0u[k] = (in[k] && gi > 0) ? w0[idx - W] : zero;
                w1c[k] = (in[k] && gj < W - 1) ? w1[idx] : zero;
                w1l[k] = (in[k] && gj > 0) ? w1[idx - 1] : zero;
            }
#pragma unroll
            for (int k = 0; k < UN; ++k) {
                const int p = pb + k * CPS_NT;
                if (p < (TH + 1) * PW) {
                    const int r = p / PW, c = p - r * PW;
                    V v = zero;
                    if (in[k]) {
                        const V x = xv[k], g = cscale(gv[k], A.gscale);
                        V a = cscale(csub(w0u[k], w0c[k]), A.is0);
                        a = caxpy(a, A.is1, csub(w1l[k], w1c[k]));
                        const V xt = caxpy(x, -A.tau, cadd(g, a));
                        v = caxpy(cscale(x, R(-1)), R(2), xt);
                        if (r < TH && c < TW) {
                            const int64_t idx = (int64_t)(i0 + r) * W + j0 + c;
                            const V xn = caxpy(cscale(x, A.omr), A.rho, xt);
                            xout[idx] = xn;
                            vout[idx] = v;
                            s_dx += cabs2(csub(xn, x));
                            s_x += cabs2(x);
                        }
                    }
                    sv[p] = v;
                }
            }
        }
        __syncthreads();
        // dual update of the TV block: w' = rho P_{lam-ball}(w + sigma Grad v) + (1 - rho) w
        // (fenchel_prox of lam * L21Norm, core/functional.py:207 + func/penalty.py:551-557, in its projection form)
        static_assert((TH * TW) % (UN * CPS_NT) == 0, "tile size must be a whole number of batches");
        for (int pb = tid; pb < TH * TW; pb += UN * CPS_NT) {
            V a0[UN], a1[UN];
            bool in[UN];
#pragma unroll
            for (int k = 0; k < UN; ++k) {
                const int p = pb + k * CPS_NT;
                const int r = p / TW, c = p - r * TW;
                const int gi = i0 + r, gj = j0 + c;
                in[k] = gi < H && gj < W;
                const int64_t idx = in[k] ? (int64_t)gi * W + gj : 0;
                a0[k] = in[k] ? w0[idx] : zero;
                a1[k] = in[k] ? w1[idx] : zero;
            }
#pragma unroll
            for (int k = 0; k < UN; ++k) {
                const int p = pb + k * CPS_NT;
                const int r = p / TW, c = p - r * TW;
                const int gi = i0 + r, gj = j0 + c;
                if (in[k]) {
                    const int64_t idx = (int64_t)gi * W + gj;
                    const V vc = sv[r * PW + c], vd = sv[(r + 1) * PW + c], vr = sv[r * PW + c + 1];
                    const V d0 = gi < H - 1 ? cscale(csub(vd, vc), A.is0) : zero;
                    const V d1 = gj < W - 1 ? cscale(csub(vr, vc), A.is1) : zero;
                    const V q0 = caxpy(a0[k], A.sigma, d0), q1 = caxpy(a1[k], A.sigma, d1);
                    const R ss = fma(q0.x, q0.x, fma(q0.y, q0.y, fma(q1.x, q1.x, fma(q1.y, q1.y, tiny<R>()))));
                    R f = A.lam * rsq<R>(ss);
                    f = f < R(1) ? f : R(1);
                    const V n0 = caxpy(cscale(a0[k], A.omr), A.rho * f, q0), n1 = caxpy(cscale(a1[k], A.omr), A.rho * f, q1);
                    w0out[idx] = n0;
                    w1out[idx] = n1;
                    s_dw += cabs2(csub(n0, a0[k])) + cabs2(csub(n1, a1[k]));
                    s_w += cabs2(a0[k]) + cabs2(a1[k]);
                }
            }
        }
        __syncthreads();
    }
    double v4[4] = {s_dx, s_x, s_dw, s_w};
    grid_sum_finish<4, CPS_NT>(v4, A.partials, A.ticket, A.diag_out, red);
}

template <typename R>
__global__ void __launch_bounds__(CPS_NT) cps_scatter_kernel(typename Cx<R>::type* __restrict__ out,
                                                             const typename Cx<R>::type* __restrict__ u,
                                                             const int64_t* __restrict__ idx, int64_t m) {
    for (int64_t k = blockIdx.x * (int64_t)CPS_NT + threadIdx.x; k < m; k += (int64_t)gridDim.x * CPS_NT) out[idx[k]] = u[k];
}

template <typename R> struct CpsBallArgs {
    typedef typename Cx<R>::type V;
    V* u_out; V* spec_next; const V* u; const V* spec; const int64_t* idx; const V* y;
    int64_t m;
    R eps, sigma, inv_sigma, rho, omr, vscale;
    double* partials; unsigned int* ticket; double* sumsq; double* tmp; double* diag;
};

// d = (u + sigma V[mask]) / sigma - y for sample k (fenchel_prox evaluates prox at z / sigma, core/functional.py:207;
// the loss shifts by -y, func/loss.py:69)
template <typename R>
__device__ __forceinline__ void ball_point(const CpsBallArgs<R>& A, int64_t k, typename Cx<R>::type& p,
                                           typename Cx<R>::type& d) {
    p = caxpy(A.u[k], A.sigma * A.vscale, A.spec[A.idx[k]]);
    d = csub(cscale(p, A.inv_sigma), A.y[k]);
}

template <typename R> __global__ void __launch_bounds__(CPS_NT) cps_ball_norm_kernel(const CpsBallArgs<R> A) {
    typedef typename Cx<R>::type V;
    __shared__ double red[CPS_NT / 32];
    double acc = 0.0;
    for (int64_t k = blockIdx.x * (int64_t)CPS_NT + threadIdx.x; k < A.m; k += (int64_t)gridDim.x * CPS_NT) {
        V p, d;
        ball_point<R>(A, k, p, d);
        acc += cabs2(d);
    }
    double v1[1] = {acc};
    grid_sum_finish<1, CPS_NT>(v1, A.partials, A.ticket, A.sumsq, red);
}

// proj_l2_ball (math/prox.py:207-210): x if ||x|| <= r else r x / ||x||; then the Moreau identity and the momentum step
template <typename R> __global__ void __launch_bounds__(CPS_NT) cps_ball_apply_kernel(const CpsBallArgs<R> A) {
    typedef typename Cx<R>::type V;
    __shared__ double red[2 * (CPS_NT / 32)];
    const double nrm = sqrt(*A.sumsq);
    const bool inside = nrm <= (double)A.eps;
    const R f = inside ? R(1) : (R)((double)A.eps / nrm);
    double s_du = 0.0, s_u = 0.0;
    for (int64_t k = blockIdx.x * (int64_t)CPS_NT + threadIdx.x; k < A.m; k += (int64_t)gridDim.x * CPS_NT) {
        V p, d;
        ball_point<R>(A, k, p, d);
        const V prox = cadd(inside ? d : cscale(d, f), A.y[k]);
        const V ut = caxpy(p, -A.sigma, prox);
        const V uo = A.u[k];
        const V un = caxpy(cscale(uo, A.omr), A.rho, ut);
        A.u_out[k] = un;
        if (A.spec_next != nullptr) A.spec_next[A.idx[k]] = un;      // Mask^T u' for the next iteration's inverse DFT
        s_du += cabs2(csub(un, uo));
        s_u += cabs2(uo);
    }
    double v2[2] = {s_du, s_u};
    const bool last = grid_sum_finish<2, CPS_NT>(v2, A.partials, A.ticket, A.tmp, red);
    if (last && threadIdx.x == 0) {          // the TV kernel of this iteration wrote the w part of the dual sums
        A.diag[2] += A.tmp[0];
        A.diag[3] += A.tmp[1];
    }
}

// One-kernel form of the ball step (cooperative launch: every block resident, one grid-wide barrier).  A block keeps
// p = u + sigma V[mask] and the grid index of its samples in SHARED memory across the barrier (12 / 20 bytes per sample,
// up to spt samples per thread), so the gather V[mask] and the 8-byte index list are read ONCE instead of twice; u and
// y (coalesced) are read again.  The norm is summed in a fixed order (per-block partials, then every block adds up all of
// them in the same order): deterministic, and identical in all blocks.
template <typename R> __global__ void __launch_bounds__(CPS_NT) cps_ball_fused_kernel(const CpsBallArgs<R> A, int spt) {
    typedef typename Cx<R>::type V;
    extern __shared__ __align__(16) unsigned char smem[];
    V* const sp = reinterpret_cast<V*>(smem);                                        // [spt][CPS_NT]
    int* const sgi = reinterpret_cast<int*>(smem + (size_t)spt * CPS_NT * sizeof(V));  // [spt][CPS_NT]
    __shared__ double red[2 * (CPS_NT / 32)];
    __shared__ double s_total;
    cg::grid_group grid = cg::this_grid();
    const int tid = threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * CPS_NT, k0 = blockIdx.x * (int64_t)CPS_NT + tid;
    double acc = 0.0;
#pragma unroll 4
    for (int q = 0; q < spt; ++q) {
        const int64_t k = k0 + q * stride;
        if (k < A.m) {
            const int64_t g = A.idx[k];
            const V p = caxpy(A.u[k], A.sigma * A.vscale, A.spec[g]);
            sp[q * CPS_NT + tid] = p;
            sgi[q * CPS_NT + tid] = (int)g;
            acc += cabs2(csub(cscale(p, A.inv_sigma), A.y[k]));
        }
    }
    double v1[1] = {acc};
    block_sum<1, CPS_NT>(v1, red);
    if (tid == 0) A.partials[blockIdx.x] = v1[0];
    grid.sync();
    double t = 0.0;
    for (int i = tid; i < (int)gridDim.x; i += CPS_NT) t += __ldcg(&A.partials[i]);
    double v1b[1] = {t};
    __syncthreads();
    block_sum<1, CPS_NT>(v1b, red);
    if (tid == 0) s_total = v1b[0];
    __syncthreads();
    const double nrm = sqrt(s_total);
    const bool inside = nrm <= (double)A.eps;
    const R f = inside ? R(1) : (R)((double)A.eps / nrm);
    double s_du = 0.0, s_u = 0.0;
#pragma unroll 4
    for (int q = 0; q < spt; ++q) {
        const int64_t k = k0 + q * stride;
        if (k < A.m) {
            const V p = sp[q * CPS_NT + tid], yk = A.y[k], uo = A.u[k];
            const V d = csub(cscale(p, A.inv_sigma), yk);
            const V prox = cadd(inside ? d : cscale(d, f), yk);
            const V ut = caxpy(p, -A.sigma, prox);
            const V un = caxpy(cscale(uo, A.omr), A.rho, ut);
            A.u_out[k] = un;
            if (A.spec_next != nullptr) A.spec_next[sgi[q * CPS_NT + tid]] = un;
            s_du += cabs2(csub(un, uo));
            s_u += cabs2(uo);
        }
    }
    grid.sync();                               // every block has read the norm partials: the slots can be reused
    double v2[2] = {s_du, s_u};
    const bool last = grid_sum_finish<2, CPS_NT>(v2, A.partials, A.ticket, A.tmp, red);
    if (last && tid == 0) {
        A.diag[2] += A.tmp[0];
        A.diag[3] += A.tmp[1];
    }
}

// workspace (doubles): [0, 4 MAXB) partial sums | ticket | tmp[4] | sumsq
constexpr size_t CPS_WS_DOUBLES = 4 * (size_t)CPS_MAX_BLOCKS + 8;

template <typename R>
int cps_scatter_impl(void* spec_out, const void* u, const void* idx, int64_t N, int64_t m, cudaStream_t s) {
    typedef typename Cx<R>::type V;
    const char* what = "pcs_cps_mri_scatter";
    PCS_REQUIRE(N >= 1 && m >= 0 && m <= N, "%s: bad sizes", what);
    PCS_REQUIRE(spec_out && (m == 0 || (u && idx)), "%s: null pointer", what);
    if (cudaMemsetAsync(spec_out, 0, (size_t)N * sizeof(V), s) != cudaSuccess) return check_launch(what);
    if (m > 0) {
        const int64_t nb = (m + CPS_NT - 1) / CPS_NT;
        const int grid = (int)(nb < 8 * (int64_t)num_sms() ? nb : 8 * (int64_t)num_sms());
        cps_scatter_kernel<R><<<grid, CPS_NT, 0, s>>>((V*)spec_out, (const V*)u, (const int64_t*)idx, m);
    }
    return check_launch(what);
}

template <typename R>
int cps_tv_impl(void* x_out, void* w_out, void* v_out, const void* x, const void* w, const void* g, int64_t H, int64_t W,
                double gscale, double lam, double tau, double sigma, double rho, double step0, double step1,
                double* diag_out, void* workspace, cudaStream_t s) {
    typedef typename Cx<R>::type V;
    const char* what = "pcs_cps_mri_tv_step";
    PCS_REQUIRE(H >= 1 && W >= 1 && H < (1ll << 30) && W < (1ll << 30), "%s: bad dimensions", what);
    PCS_REQUIRE(x_out && w_out && v_out && x && w && g && diag_out && workspace, "%s: null pointer", what);
    PCS_REQUIRE(x_out != x && w_out != w, "%s: outputs must not alias inputs (ping-pong)", what);
    PCS_REQUIRE(lam > 0 && tau > 0 && sigma > 0 && step0 != 0 && step1 != 0, "%s: tau, sigma, lam must be positive", what);
    CpsTvArgs<R> A;
    A.x_out = (V*)x_out; A.w_out = (V*)w_out; A.v_out = (V*)v_out;
    A.x = (const V*)x; A.w = (const V*)w; A.g = (const V*)g;
    A.H = (int)H; A.W = (int)W;
    A.lam = (R)lam; A.tau = (R)tau; A.sigma = (R)sigma; A.rho = (R)rho; A.omr = (R)(1 - rho);
    A.is0 = (R)(1 / step0); A.is1 = (R)(1 / step1); A.gscale = (R)gscale;
    A.partials = (double*)workspace;
    A.ticket = (unsigned int*)((double*)workspace + 4 * (size_t)CPS_MAX_BLOCKS);
    A.diag_out = diag_out;
    {
        // row-marching kernel: needs 16-byte aligned rows and planes (PCS_CPS_TV_TILE=1 keeps the tile kernel)
        constexpr int NC = 16 / (int)sizeof(V);
        static const bool tile_only = [] { const char* e = getenv("PCS_CPS_TV_TILE"); return e && atoi(e) != 0; }();
        bool ok = !tile_only && W % NC == 0 && ((H * W) % NC) == 0;
        for (const void* q : {(const void*)x_out, (const void*)w_out, (const void*)v_out, x, w, g}) ok = ok && ((uintptr_t)q % 16) == 0;
        if (ok) {
            const int cols_per_block = (CPM_NT / 32) * 31 * NC;
            const int gx = (int)((W + cols_per_block - 1) / cols_per_block);
            // exactly ONE wave of resident blocks (a second, partial wave cost 1.3x), segments of at least 16 rows (a
            // segment re-reads one row of w0 and recomputes one row of v)
            int per_sm = 0;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, cps_tv_march_kernel<R>, CPM_NT, 0) != cudaSuccess || per_sm < 1) {
                (void)cudaGetLastError();
                per_sm = 4;
            }
            int64_t nseg = ((int64_t)per_sm * num_sms()) / gx;
            if (nseg < 1) nseg = 1;
            int64_t seg_rows = (H + nseg - 1) / nseg;
            if (seg_rows < 16) seg_rows = 16;
            nseg = (H + seg_rows - 1) / seg_rows;
            if ((int64_t)gx * nseg <= CPS_MAX_BLOCKS) {
                cps_tv_march_kernel<R><<<dim3((unsigned)gx, (unsigned)nseg), CPM_NT, 0, s>>>(A, (int)seg_rows);
                return check_launch(what);
            }
        }
    }
    constexpr int TW = TileW<R>::value;
    const int64_t ntiles = ((H + CPS_TH - 1) / CPS_TH) * ((W + TW - 1) / TW);
    // every tile its own block while the partial-sum slots last (6 blocks of 34 KB fit per SM), else a grid-stride loop
    const int64_t cap = CPS_MAX_BLOCKS;
    const int grid = (int)(ntiles < cap ? ntiles : cap);
    const size_t smem = (size_t)(CPS_TH + 1) * (TW + 1) * sizeof(V) + 4 * (CPS_NT / 32) * sizeof(double);
    cps_tv_kernel<R><<<grid, CPS_NT, smem, s>>>(A);
    return check_launch(what);
}

template <typename R>
int cps_ball_impl(void* u_out, void* spec_next, const void* u, const void* spec, const void* idx, const void* y, int64_t m,
                  double vscale, double eps, double sigma, double rho, double* diag_inout, void* workspace, cudaStream_t s) {
    typedef typename Cx<R>::type V;
    const char* what = "pcs_cps_mri_ball_step";
    PCS_REQUIRE(m >= 0 && sigma > 0 && eps >= 0, "%s: bad arguments", what);
    PCS_REQUIRE(diag_inout && workspace && (m == 0 || (u_out && u && spec && idx && y)), "%s: null pointer", what);
    if (m == 0) return 0;
    CpsBallArgs<R> A;
    A.u_out = (V*)u_out; A.spec_next = (V*)spec_next; A.u = (const V*)u; A.spec = (const V*)spec; A.idx = (const int64_t*)idx; A.y = (const V*)y;
    A.m = m;
    A.eps = (R)eps; A.sigma = (R)sigma; A.inv_sigma = (R)(1 / sigma); A.rho = (R)rho; A.omr = (R)(1 - rho);
    A.vscale = (R)vscale;
    PCS_REQUIRE(spec_next != spec, "%s: spec_next must not alias spec", what);
    double* ws = (double*)workspace;
    A.partials = ws;
    A.ticket = (unsigned int*)(ws + 4 * (size_t)CPS_MAX_BLOCKS);
    A.tmp = ws + 4 * (size_t)CPS_MAX_BLOCKS + 1;
    A.sumsq = ws + 4 * (size_t)CPS_MAX_BLOCKS + 5;
    A.diag = diag_inout;
    const int64_t nb = (m + CPS_NT - 1) / CPS_NT;
    const int64_t cap = 8 * (int64_t)num_sms() < CPS_MAX_BLOCKS ? 8 * (int64_t)num_sms() : CPS_MAX_BLOCKS;
    const int grid = (int)(nb < cap ? nb : cap);
    // one cooperative kernel when the samples fit in the shared memory of one resident wave, else norm pass + update pass
    static const bool no_coop = [] { const char* e = getenv("PCS_CPS_BALL_TWO_PASS"); return e && atoi(e) != 0; }();
    if (!no_coop) {                            // (grid indices are kept as int32: images below 2^31 pixels)
        const size_t per_sample = sizeof(V) + sizeof(int);
        for (int spt = 1; spt <= 16; ++spt) {
            const size_t smem = (size_t)spt * CPS_NT * per_sample;
            if (smem > 96 * 1024) break;
            int per_sm = 0;
            if (smem > 48 * 1024)
                cudaFuncSetAttribute(cps_ball_fused_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, cps_ball_fused_kernel<R>, CPS_NT, smem) != cudaSuccess ||
                per_sm < 1) {
                (void)cudaGetLastError();
                break;
            }
            int64_t cgrid = (int64_t)per_sm * num_sms();
            if (cgrid > CPS_MAX_BLOCKS) cgrid = CPS_MAX_BLOCKS;
            if (nb < cgrid) cgrid = nb;
            if (cgrid * CPS_NT * spt < m) continue;          // more samples per thread needed
            void* args[] = {(void*)&A, (void*)&spt};
            if (cudaLaunchCooperativeKernel((void*)cps_ball_fused_kernel<R>, dim3((unsigned)cgrid), dim3(CPS_NT), args, smem, s) ==
                cudaSuccess)
                return check_launch(what);
            (void)cudaGetLastError();          // cooperative launch not possible here: fall through to the two passes
            break;
        }
    }
    cps_ball_norm_kernel<R><<<grid, CPS_NT, 0, s>>>(A);
    cps_ball_apply_kernel<R><<<grid, CPS_NT, 0, s>>>(A);
    return check_launch(what);
}

}  // namespace

extern "C" int64_t pcs_cps_mri_workspace_bytes(void) { return (int64_t)(CPS_WS_DOUBLES * sizeof(double)); }

#define PCS_EXPORT_CPLX(name, impl, params, args)                          \
    extern "C" int name##_c64 params { return impl<float> args; }          \
    extern "C" int name##_c128 params { return impl<double> args; }

PCS_EXPORT_CPLX(pcs_cps_mri_scatter, cps_scatter_impl,
                (void* spec_out, const void* u, const void* idx_i64, int64_t N, int64_t m, void* stream),
                (spec_out, u, idx_i64, N, m, (cudaStream_t)stream))
PCS_EXPORT_CPLX(pcs_cps_mri_tv_step, cps_tv_impl,
                (void* x_out, void* w_out, void* v_out, const void* x, const void* w, const void* g, int64_t H, int64_t W,
                 double gscale, double lam, double tau, double sigma, double rho, double step0, double step1,
                 void* diag_out_f64, void* workspace, void* stream),
                (x_out, w_out, v_out, x, w, g, H, W, gscale, lam, tau, sigma, rho, step0, step1, (double*)diag_out_f64,
                 workspace, (cudaStream_t)stream))
PCS_EXPORT_CPLX(pcs_cps_mri_ball_step, cps_ball_impl,
                (void* u_out, void* spec_next, const void* u, const void* spec, const void* idx_i64, const void* y,
                 int64_t m, double vscale, double eps, double sigma, double rho, void* diag_inout_f64, void* workspace,
                 void* stream),
                (u_out, spec_next, u, spec, idx_i64, y, m, vscale, eps, sigma, rho, (double*)diag_inout_f64, workspace,
                 (cudaStream_t)stream))
