// libpcsb200: a whole AcceleratedProximalGradientDescent solve in ONE kernel launch (sm_100a).
//
// Replaces AcceleratedProximalGradientDescent.iterate() -- the driver loop of core/solver.py:55-76 with
// update_iterand / update_diagnostics / stopping_metric of opt/proxalgs.py:586-622 -- for the sparse-deconvolution
// problem of BASELINE.json configs[0]:
//     F = 1/2 ||h * x - y||^2   (Convolve1D on a flat vector, zero "same" boundary; linop/conv.py:20-164)
//     G in {none, lam * L1Norm, NonNegativeOrthant}
// At N = 4096 the op-by-op path is pure launch latency (ten launches of a few microseconds of work per update, and a
// host round trip for the stop test).  Here one block keeps x, the residual and the filter in shared memory (its own elements of x, y and the
// previous prox output in registers) and runs every update -- convolution, correlation, prox, momentum, the two norms of the
// relative improvement AND the reference's stop test -- without leaving the SM:
//     r = h * x - y          g = h (corr) r          x~ = prox_{tau G}(x - tau g)
//     t' = next(t, it)       a = (t - 1) / t'        x' = (1 + a) x~ - a x~_prev
//     metric = ||x' - x|| / ||x||   (inf if x = 0);  loop while (it <= max_iter and metric > thr) or it <= min_iter
// The momentum schedule and the stop test are evaluated in fp64 with the reference's operation order (sqrt and
// division are correctly rounded on both sides), so the number of updates is the reference's.
#include "pcs_common.cuh"
#include <math.h>

using namespace pcs;

namespace {

constexpr int APGD_MAX_TAPS = 256;
constexpr int APGD_PAD = 272;                 // zeros kept on either side of x and of the residual (>= padded taps + chunk)
constexpr int APGD_C = 8;                     // taps per register tile
constexpr int APGD_MAX_N = 8192;

enum { AG_NONE = 0, AG_L1 = 1, AG_NONNEG = 2 };
enum { ACC_NONE = 0, ACC_BT = 1, ACC_CD = 2 };

template <typename T> struct ApgdArgs {
    T* x;                 // in: x0, out: final iterate
    T* aux;               // in: past_aux, out: last prox output (past_aux of the reference's iterand)
    const T* y;
    const T* h;
    int nh, offset, n, g_mode, accel;
    T g_param, tau;
    double d, t0, threshold;
    long long it0, max_iter, min_iter;
    double* diag;         // rows of 4 doubles: ||x'-x||^2, ||x||^2, 0, 0 (row index = update index - it0)
    long long* n_iter;    // [0]: update index after the loop (the reference's self.iter); [1]: t as raw double bits
};

// deterministic block sum of two doubles (fixed tree), result broadcast to every thread
template <int NT> __device__ __forceinline__ void block_sum2(double& a, double& b, double* scratch) {
    constexpr int NW = NT / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    a = warp_sum(a);
    b = warp_sum(b);
    if (lane == 0) {
        scratch[warp] = a;
        scratch[32 + warp] = b;
    }
    __syncthreads();
    if (warp == 0) {
        double ta = lane < NW ? scratch[lane] : 0.0, tb = lane < NW ? scratch[32 + lane] : 0.0;
        ta = warp_sum(ta);
        tb = warp_sum(tb);
        if (lane == 0) {
            scratch[64] = ta;
            scratch[65] = tb;
        }
    }
    __syncthreads();
    a = scratch[64];
    b = scratch[65];
    __syncthreads();
}

// Thread t owns the EPT consecutive elements [t EPT, (t+1) EPT).  x and the residual live in shared memory with zeros
// on either side (the zero "same" boundary, no bounds checks).  Both filter passes are register tiled: for a chunk of
// C taps a thread loads the EPT + C - 1 samples its outputs need once and issues EPT x C multiply-adds on static
// registers -- (EPT + 2 C - 1) shared loads per EPT x C FMAs, EPT independent accumulation chains.
template <typename T, int EPT, int NT>
__global__ void __launch_bounds__(NT, 1) apgd_run_kernel(const ApgdArgs<T> A) {
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int C = APGD_C;
    const int n = A.n, off = A.offset;
    const int nhp = (A.nh + C - 1) / C * C;                    // taps padded with zeros to whole chunks
    const int np = n + 2 * APGD_PAD;                           // padded length
    T* sx = reinterpret_cast<T*>(smem) + APGD_PAD;             // current iterate, zero padded
    T* sr = sx + np;                                           // residual h * x - y, zero padded
    T* sh = sr + np - APGD_PAD;                                // taps, zero padded to nhp
    double* scratch = reinterpret_cast<double*>(smem + ((size_t)(2 * np + APGD_MAX_TAPS) * sizeof(T) + 15) / 16 * 16);
    const int tid = threadIdx.x, i0 = tid * EPT;
    const bool active = i0 < n;                  // threads past the end only take part in the barriers and sums

    for (int i = tid; i < 2 * np + APGD_MAX_TAPS; i += NT) reinterpret_cast<T*>(smem)[i] = T(0);
    __syncthreads();
    T xa[EPT], pa[EPT];                          // this thread's elements: iterate, previous prox output
#pragma unroll
    for (int e = 0; e < EPT; ++e) {
        const int i = i0 + e;
        xa[e] = i < n ? A.x[i] : T(0);
        pa[e] = i < n ? A.aux[i] : T(0);
        if (i < n) sx[i] = xa[e];
    }
    for (int k = tid; k < A.nh; k += NT) sh[k] = A.h[k];
    __syncthreads();

    const T tau = A.tau;
    double t_old = A.t0, metric = INFINITY;
    long long it = A.it0;
    while ((it <= A.max_iter && metric > A.threshold) || it <= A.min_iter) {      // core/solver.py:65-66
        // residual r = h * x - y : (h * x)[i0 + e] = sum_k h[k] x[i0 + e + off - k]
        if (active) {
            T acc[EPT];
#pragma unroll
            for (int e = 0; e < EPT; ++e) acc[e] = T(0);
            const T* px = sx + i0 + off;
            for (int K = 0; K < nhp; K += C) {
                T hk[C], wv[EPT + C - 1];
#pragma unroll
                for (int c = 0; c < C; ++c) hk[c] = sh[K + c];
#pragma unroll
                for (int m = 0; m < EPT + C - 1; ++m) wv[m] = px[m - K - (C - 1)];
#pragma unroll
                for (int c = 0; c < C; ++c)
#pragma unroll
                    for (int e = 0; e < EPT; ++e) acc[e] = fma(hk[c], wv[e - c + C - 1], acc[e]);
            }
#pragma unroll
            for (int e = 0; e < EPT; ++e)
                if (i0 + e < n) sr[i0 + e] = acc[e] - A.y[i0 + e];
        }
        __syncthreads();
        double t, a;
        if (A.accel == ACC_BT) t = (1.0 + sqrt(1.0 + 4.0 * t_old * t_old)) / 2.0;     // opt/proxalgs.py:592-598
        else if (A.accel == ACC_CD) t = ((double)it + A.d) / A.d;
        else { t = 1.0; t_old = 1.0; }
        a = (t_old - 1.0) / t;
        const T ta = (T)a, t1a = (T)(1.0 + a);
        double s_d = 0.0, s_x = 0.0;
        if (active) {
            // gradient g = h^T r : g[i0 + e] = sum_k h[k] r[i0 + e - off + k]; then prox, momentum and the two norms
            T g[EPT];
#pragma unroll
            for (int e = 0; e < EPT; ++e) g[e] = T(0);
            const T* pr = sr + i0 - off;
            for (int K = 0; K < nhp; K += C) {
                T hk[C], wv[EPT + C - 1];
#pragma unroll
                for (int c = 0; c < C; ++c) hk[c] = sh[K + c];
#pragma unroll
                for (int m = 0; m < EPT + C - 1; ++m) wv[m] = pr[K + m];
#pragma unroll
                for (int c = 0; c < C; ++c)
#pragma unroll
                    for (int e = 0; e < EPT; ++e) g[e] = fma(hk[c], wv[e + c], g[e]);
            }
#pragma unroll
            for (int e = 0; e < EPT; ++e) {
                if (i0 + e < n) {
                    const T v = xa[e] - mul_rn(tau, g[e]);      // x - tau * grad, product rounded on its own
                    T p;
                    if (A.g_mode == AG_L1) p = prox_l1<T>(v, A.g_param);
                    else if (A.g_mode == AG_NONNEG) p = v < T(0) ? T(0) : v;
                    else p = v;
                    const T xn = mul_rn(t1a, p) - mul_rn(ta, pa[e]);
                    const double dd = (double)xn - (double)xa[e];
                    s_d += dd * dd;
                    s_x += (double)xa[e] * (double)xa[e];
                    pa[e] = p;
                    xa[e] = xn;
                }
            }
        }
        __syncthreads();                                         // everybody is done reading sx / sr
#pragma unroll
        for (int e = 0; e < EPT; ++e)
            if (i0 + e < n) sx[i0 + e] = xa[e];
        block_sum2<NT>(s_d, s_x, scratch);                       // also publishes the new sx
        if (tid == 0) {
            double* row = A.diag + 4 * (it - A.it0);
            row[0] = s_d; row[1] = s_x; row[2] = 0.0; row[3] = 0.0;
        }
        metric = s_x == 0.0 ? INFINITY : sqrt(s_d) / sqrt(s_x);  // opt/proxalgs.py:607-622 (np.linalg.norm ratio)
        t_old = t;
        ++it;
    }
#pragma unroll
    for (int e = 0; e < EPT; ++e) {
        if (i0 + e < n) {
            A.x[i0 + e] = xa[e];
            A.aux[i0 + e] = pa[e];
        }
    }
    if (tid == 0) {
        A.n_iter[0] = it;
        A.n_iter[1] = __double_as_longlong(t_old);
    }
}

template <typename T, int EPT, int NT> int apgd_launch(const ApgdArgs<T>& A, size_t smem, cudaStream_t s) {
    cudaFuncSetAttribute(apgd_run_kernel<T, EPT, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    apgd_run_kernel<T, EPT, NT><<<1, NT, smem, s>>>(A);
    return check_launch("pcs_apgd_conv1d_run");
}

template <typename T>
int pcs_apgd_conv1d_run_impl(T* x, T* aux, const T* y, const T* h, int nh, int offset, int64_t n, int g_mode,
                             double g_param, double tau, int accel, double d, double t0, int64_t it0, int64_t max_iter,
                             int64_t min_iter, double threshold, double* diag, long long* n_iter, cudaStream_t s) {
    const char* what = "pcs_apgd_conv1d_run";
    PCS_REQUIRE(x && aux && y && h && diag && n_iter, "%s: null pointer", what);
    PCS_REQUIRE(n >= 1 && n <= APGD_MAX_N, "%s: n must be in [1, %d]", what, APGD_MAX_N);
    PCS_REQUIRE(nh >= 1 && nh <= APGD_MAX_TAPS && offset >= 0 && offset < nh, "%s: bad filter", what);
    PCS_REQUIRE(g_mode >= AG_NONE && g_mode <= AG_NONNEG && accel >= ACC_NONE && accel <= ACC_CD, "%s: bad mode", what);
    PCS_REQUIRE(tau > 0 && (g_mode != AG_L1 || g_param > 0) && (accel != ACC_CD || d > 0), "%s: bad parameters", what);
    PCS_REQUIRE(it0 >= 0 && max_iter >= -1 && min_iter >= -1, "%s: bad iteration bounds", what);
    ApgdArgs<T> A;
    A.x = x; A.aux = aux; A.y = y; A.h = h; A.nh = nh; A.offset = offset; A.n = (int)n; A.g_mode = g_mode;
    A.accel = accel; A.g_param = (T)g_param; A.tau = (T)tau; A.d = d; A.t0 = t0; A.threshold = threshold;
    A.it0 = it0; A.max_iter = max_iter; A.min_iter = min_iter; A.diag = diag; A.n_iter = n_iter;
    const size_t smem = ((size_t)(2 * (n + 2 * APGD_PAD) + APGD_MAX_TAPS) * sizeof(T) + 15) / 16 * 16 + 66 * sizeof(double);
    PCS_REQUIRE(smem <= 227 * 1024, "%s: problem does not fit in shared memory", what);
    // Up to 17 consecutive elements per thread (17 independent FMA chains, ~1.4 shared loads per 8 FMAs).  The count
    // is ODD on purpose: thread t's window starts EPT words after thread t-1's, and an odd stride spreads a warp's
    // loads over all 32 banks (16 elements per thread would be a 16-way bank conflict on every load).
    if (n <= 5 * 128) return apgd_launch<T, 5, 128>(A, smem, s);
    if (n <= 9 * 256) return apgd_launch<T, 9, 256>(A, smem, s);
    if (n <= 17 * 256) return apgd_launch<T, 17, 256>(A, smem, s);
    return apgd_launch<T, 17, 512>(A, smem, s);
}

}  // namespace

extern "C" int64_t pcs_apgd_conv1d_run_max_n(void) { return (int64_t)APGD_MAX_N; }

PCS_EXPORT_REAL(pcs_apgd_conv1d_run,
                (void* x_inout, void* aux_inout, const void* y, const void* h, int nh, int offset, int64_t n, int g_mode,
                 double g_param, double tau, int accel, double d, double t0, int64_t it0, int64_t max_iter,
                 int64_t min_iter, double threshold, void* diag_out_f64, void* n_iter_out_i64, void* stream),
                ((T*)x_inout, (T*)aux_inout, (const T*)y, (const T*)h, nh, offset, n, g_mode, g_param, tau, accel, d, t0,
                 it0, max_iter, min_iter, threshold, (double*)diag_out_f64, (long long*)n_iter_out_i64,
                 (cudaStream_t)stream))
