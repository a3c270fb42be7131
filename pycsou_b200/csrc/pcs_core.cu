// libpcsb200: error plumbing, elementwise / proximal kernels and deterministic reductions (sm_100a).
//
// Reference rows: a19-a25.  Formulas follow the reference literally where rounding
// could differ (see pcs_common.cuh: prox_l1, l21_factor).
#include "pcs_common.cuh"
#include <stdarg.h>
#include <string.h>

namespace pcs {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return PCS_ERR_CUDA;
    }
    return PCS_OK;
}

// planning override (pcs_set_num_sms): lets the tests drive the one-wave / uneven-split launch geometries of the fused
// kernels at small image sizes; 0 = the device's own count
static int g_num_sms_override = 0;

int num_sms() {
    if (g_num_sms_override > 0) return g_num_sms_override;
    static thread_local int cached_dev = -1, cached = 148;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev != cached_dev) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) cached = n;
        cached_dev = dev;
    }
    return cached;
}

}  // namespace pcs

using namespace pcs;

extern "C" int pcs_version(void) { return 100; }
extern "C" const char* pcs_last_error(void) { return pcs::g_err; }
extern "C" int pcs_num_sms(void) { return pcs::num_sms(); }
extern "C" int pcs_set_num_sms(int n) {
    const int prev = pcs::g_num_sms_override;
    pcs::g_num_sms_override = n > 0 ? n : 0;
    return prev;
}

// ------------------------------------------------------------------------------------------------ elementwise
constexpr int EW_BLOCK = 256;
constexpr int EW_CTAS_PER_SM = 8;

enum EwOp {
    EW_LINCOMB2,    // out = a*x + b*y
    EW_LINCOMB3,    // out = a*x + b*y + c*z
    EW_SCALE,       // out = a*x
    EW_MUL,         // out = x*y
    EW_MULCONJ,     // (real: same as MUL)
    EW_PROX_L1,     // out = x - a*clip(x/a,-1,1)
    EW_NONNEG,      // out = max(x,0)
    EW_SEGMENT,     // out = clip(x,a,b)
    EW_SCALE_DEV,   // out = x * s[0]  (device scalar; used by l2 prox / l2-ball projection)
    EW_PROX_LOGBARRIER,   // out = (x + sqrt(x^2 + 4a)) / 2
    EW_PROX_KL,           // out = (x - a + sqrt((x - a)^2 + 4 a y)) / 2
    EW_PROX_SHANNON,      // out = a * W(exp(x/a - 1) / a), W = principal branch of the Lambert function
};

// a * W(exp(x/a - 1)/a) evaluated in the log domain (no overflow for large x/a): w solves w + log w = L with
// L = x/a - 1 - log a; Newton on that equation, w <- w (1 + L - log w) / (1 + w), converges quadratically from
// w0 = L - log L (L > 1) or exp(L) / (1 + exp(L)) (else); 12 steps reach double precision from either start.
template <typename T> __device__ __forceinline__ T lambert_scaled(T x, T a) {
    const double L = (double)x / (double)a - 1.0 - log((double)a);
    double w;
    if (L > 1.0) w = L - log(L);
    else { const double e = exp(L); w = e / (1.0 + e); }
    if (!(w > 0.0)) return T(0);                       // exp(L) underflowed: W(z) ~ z = 0
#pragma unroll 1
    for (int it = 0; it < 12; ++it) {
        const double wn = w * (1.0 + L - log(w)) / (1.0 + w);
        if (!(wn > 0.0)) break;
        const bool done = fabs(wn - w) <= 1e-16 * wn;
        w = wn;
        if (done) break;
    }
    return (T)((double)a * w);
}

template <typename T, int OP>
__device__ __forceinline__ T ew_apply(T x, T y, T z, T a, T b, T c) {
    if (OP == EW_LINCOMB2) return a * x + b * y;
    if (OP == EW_LINCOMB3) return a * x + b * y + c * z;
    if (OP == EW_SCALE) return a * x;
    if (OP == EW_MUL || OP == EW_MULCONJ) return x * y;
    if (OP == EW_PROX_L1) return prox_l1<T>(x, a);
    if (OP == EW_NONNEG) return x < T(0) ? T(0) : x;
    if (OP == EW_SEGMENT) { T v = x < a ? a : x; return v > b ? b : v; }
    if (OP == EW_PROX_LOGBARRIER) return (x + tsqrt<T>(x * x + T(4) * a)) / T(2);
    if (OP == EW_PROX_KL) { const T d = x - a; return (d + tsqrt<T>(d * d + T(4) * a * y)) / T(2); }
    if (OP == EW_PROX_SHANNON) return lambert_scaled<T>(x, a);
    return x;
}

template <typename T, int OP, int NIN>
__global__ void __launch_bounds__(EW_BLOCK) ew_kernel(T* __restrict__ out, const T* __restrict__ x,
                                                      const T* __restrict__ y, const T* __restrict__ z, T a, T b, T c,
                                                      int64_t n, int vec_ok) {
    typedef typename Vec<T>::type V;
    constexpr int VN = Vec<T>::N;
    const int64_t tid = (int64_t)blockIdx.x * EW_BLOCK + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * EW_BLOCK;
    if (vec_ok) {
        const int64_t nv = n / VN;
        int64_t i = tid;
        if (NIN == 1) {
            // one-operand kernels: two independent 16-byte loads in flight per thread (one kept them at 4.6-5.4 TB/s)
            for (; i + stride < nv; i += 2 * stride) {
                const V v0 = reinterpret_cast<const V*>(x)[i], v1 = reinterpret_cast<const V*>(x)[i + stride];
                V o0, o1;
                const T* p0 = reinterpret_cast<const T*>(&v0);
                const T* p1 = reinterpret_cast<const T*>(&v1);
                T* q0 = reinterpret_cast<T*>(&o0);
                T* q1 = reinterpret_cast<T*>(&o1);
#pragma unroll
                for (int k = 0; k < VN; ++k) {
                    q0[k] = ew_apply<T, OP>(p0[k], p0[k], p0[k], a, b, c);
                    q1[k] = ew_apply<T, OP>(p1[k], p1[k], p1[k], a, b, c);
                }
                reinterpret_cast<V*>(out)[i] = o0;
                reinterpret_cast<V*>(out)[i + stride] = o1;
            }
        }
        for (; i < nv; i += stride) {
            V vx = reinterpret_cast<const V*>(x)[i], vy = vx, vz = vx, vo;
            if (NIN > 1) vy = reinterpret_cast<const V*>(y)[i];
            if (NIN > 2) vz = reinterpret_cast<const V*>(z)[i];
            const T* px = reinterpret_cast<const T*>(&vx);
            const T* py = reinterpret_cast<const T*>(&vy);
            const T* pz = reinterpret_cast<const T*>(&vz);
            T* po = reinterpret_cast<T*>(&vo);
#pragma unroll
            for (int k = 0; k < VN; ++k) po[k] = ew_apply<T, OP>(px[k], py[k], pz[k], a, b, c);
            reinterpret_cast<V*>(out)[i] = vo;
        }
        for (int64_t r = nv * VN + tid; r < n; r += stride)
            out[r] = ew_apply<T, OP>(x[r], NIN > 1 ? y[r] : T(0), NIN > 2 ? z[r] : T(0), a, b, c);
    } else {
        for (int64_t i = tid; i < n; i += stride)
            out[i] = ew_apply<T, OP>(x[i], NIN > 1 ? y[i] : T(0), NIN > 2 ? z[i] : T(0), a, b, c);
    }
}

static inline int aligned16(const void* p) { return p == nullptr || (((uintptr_t)p) & 15u) == 0; }

template <typename T, int OP, int NIN>
static int ew_launch(T* out, const T* x, const T* y, const T* z, T a, T b, T c, int64_t n, cudaStream_t s,
                     const char* what) {
    PCS_REQUIRE(n >= 0, "%s: negative size", what);
    if (n == 0) return PCS_OK;
    PCS_REQUIRE(out && x && (NIN < 2 || y) && (NIN < 3 || z), "%s: null pointer", what);
    int vec_ok = aligned16(out) && aligned16(x) && aligned16(y) && aligned16(z);
    int grid = grid_for((n + Vec<T>::N - 1) / Vec<T>::N, EW_BLOCK, EW_CTAS_PER_SM);
    ew_kernel<T, OP, NIN><<<grid, EW_BLOCK, 0, s>>>(out, x, y, z, a, b, c, n, vec_ok);
    return check_launch(what);
}

template <typename T> static int pcs_lincomb2_impl(T* out, T a, const T* x, T b, const T* y, int64_t n, cudaStream_t s) {
    return ew_launch<T, EW_LINCOMB2, 2>(out, x, y, nullptr, a, b, T(0), n, s, "pcs_lincomb2");
}
PCS_EXPORT_REAL(pcs_lincomb2, (void* out, double a, const void* x, double b, const void* y, int64_t n, void* stream),
                ((T*)out, (T)a, (const T*)x, (T)b, (const T*)y, n, (cudaStream_t)stream))

template <typename T>
static int pcs_lincomb3_impl(T* out, T a, const T* x, T b, const T* y, T c, const T* z, int64_t n, cudaStream_t s) {
    return ew_launch<T, EW_LINCOMB3, 3>(out, x, y, z, a, b, c, n, s, "pcs_lincomb3");
}
PCS_EXPORT_REAL(pcs_lincomb3,
                (void* out, double a, const void* x, double b, const void* y, double c, const void* z, int64_t n,
                 void* stream),
                ((T*)out, (T)a, (const T*)x, (T)b, (const T*)y, (T)c, (const T*)z, n, (cudaStream_t)stream))

template <typename T> static int pcs_scale_impl(T* out, T a, const T* x, int64_t n, cudaStream_t s) {
    return ew_launch<T, EW_SCALE, 1>(out, x, nullptr, nullptr, a, T(0), T(0), n, s, "pcs_scale");
}
PCS_EXPORT_REAL(pcs_scale, (void* out, double a, const void* x, int64_t n, void* stream),
                ((T*)out, (T)a, (const T*)x, n, (cudaStream_t)stream))

template <typename T> static int pcs_mul_impl(T* out, const T* x, const T* y, int64_t n, cudaStream_t s) {
    return ew_launch<T, EW_MUL, 2>(out, x, y, nullptr, T(0), T(0), T(0), n, s, "pcs_mul");
}
PCS_EXPORT_REAL(pcs_mul, (void* out, const void* x, const void* y, int64_t n, void* stream),
                ((T*)out, (const T*)x, (const T*)y, n, (cudaStream_t)stream))

// LogBarrier.prox (func/penalty.py:824-840), ShannonEntropy.prox (:905-922), KLDivergence.prox (func/loss.py:666-682)
template <typename T> static int pcs_prox_logbarrier_impl(T* out, const T* x, T tau, int64_t n, cudaStream_t s) {
    PCS_REQUIRE(tau > T(0), "pcs_prox_logbarrier: tau must be positive");
    return ew_launch<T, EW_PROX_LOGBARRIER, 1>(out, x, nullptr, nullptr, tau, T(0), T(0), n, s, "pcs_prox_logbarrier");
}
PCS_EXPORT_REAL(pcs_prox_logbarrier, (void* out, const void* x, double tau, int64_t n, void* stream),
                ((T*)out, (const T*)x, (T)tau, n, (cudaStream_t)stream))
template <typename T> static int pcs_prox_shannon_impl(T* out, const T* x, T tau, int64_t n, cudaStream_t s) {
    PCS_REQUIRE(tau > T(0), "pcs_prox_shannon: tau must be positive");
    return ew_launch<T, EW_PROX_SHANNON, 1>(out, x, nullptr, nullptr, tau, T(0), T(0), n, s, "pcs_prox_shannon");
}
PCS_EXPORT_REAL(pcs_prox_shannon, (void* out, const void* x, double tau, int64_t n, void* stream),
                ((T*)out, (const T*)x, (T)tau, n, (cudaStream_t)stream))
template <typename T> static int pcs_prox_kl_impl(T* out, const T* x, const T* data, T tau, int64_t n, cudaStream_t s) {
    PCS_REQUIRE(tau > T(0), "pcs_prox_kl: tau must be positive");
    return ew_launch<T, EW_PROX_KL, 2>(out, x, data, nullptr, tau, T(0), T(0), n, s, "pcs_prox_kl");
}
PCS_EXPORT_REAL(pcs_prox_kl, (void* out, const void* x, const void* data, double tau, int64_t n, void* stream),
                ((T*)out, (const T*)x, (const T*)data, (T)tau, n, (cudaStream_t)stream))

// L1Norm.prox (func/penalty.py:194-245 via func/base.py:239-240)
template <typename T> static int pcs_prox_l1_impl(T* out, const T* x, T tau, int64_t n, cudaStream_t s) {
    return ew_launch<T, EW_PROX_L1, 1>(out, x, nullptr, nullptr, tau, T(0), T(0), n, s, "pcs_prox_l1");
}
PCS_EXPORT_REAL(pcs_prox_l1, (void* out, const void* x, double tau, int64_t n, void* stream),
                ((T*)out, (const T*)x, (T)tau, n, (cudaStream_t)stream))

// proj_nonnegative_orthant (math/prox.py:295-297)
template <typename T> static int pcs_proj_nonneg_impl(T* out, const T* x, int64_t n, cudaStream_t s) {
    return ew_launch<T, EW_NONNEG, 1>(out, x, nullptr, nullptr, T(0), T(0), T(0), n, s, "pcs_proj_nonneg");
}
PCS_EXPORT_REAL(pcs_proj_nonneg, (void* out, const void* x, int64_t n, void* stream),
                ((T*)out, (const T*)x, n, (cudaStream_t)stream))

// proj_segment (math/prox.py:340-343)
template <typename T> static int pcs_proj_segment_impl(T* out, const T* x, T a, T b, int64_t n, cudaStream_t s) {
    return ew_launch<T, EW_SEGMENT, 1>(out, x, nullptr, nullptr, a, b, T(0), n, s, "pcs_proj_segment");
}
PCS_EXPORT_REAL(pcs_proj_segment, (void* out, const void* x, double a, double b, int64_t n, void* stream),
                ((T*)out, (const T*)x, (T)a, (T)b, n, (cudaStream_t)stream))

// ------------------------------------------------------------------------------------------------ L21 prox
// L21Norm.prox for the "planes" layout groups = tile(arange(G), C): component c of group g is x[c*G + g]
// (func/penalty.py:551-557).  `cplx` = 1 treats each element as (re,im) pair of T.
template <typename T, int C>
__global__ void __launch_bounds__(EW_BLOCK) prox_l21_kernel(T* __restrict__ out, const T* __restrict__ x, T tau,
                                                            int64_t G, int cplx) {
    const int64_t stride = (int64_t)gridDim.x * EW_BLOCK;
    const int64_t e = cplx ? 2 : 1;
    for (int64_t g = (int64_t)blockIdx.x * EW_BLOCK + threadIdx.x; g < G; g += stride) {
        T v[C * 2];
        T ss = T(0);
#pragma unroll
        for (int c = 0; c < C; ++c) {
            for (int k = 0; k < e; ++k) {
                v[c * 2 + k] = x[((int64_t)c * G + g) * e + k];
                ss += v[c * 2 + k] * v[c * 2 + k];
            }
        }
        const T f = l21_factor<T>(tsqrt<T>(ss), tau);
#pragma unroll
        for (int c = 0; c < C; ++c)
            for (int k = 0; k < e; ++k) out[((int64_t)c * G + g) * e + k] = f * v[c * 2 + k];
    }
}

// real planes, every plane 16-byte aligned: a thread owns Vec<T>::N consecutive groups (16-byte loads and stores)
template <typename T, int C>
__global__ void __launch_bounds__(EW_BLOCK) prox_l21_vec_kernel(T* __restrict__ out, const T* __restrict__ x, T tau,
                                                                int64_t G) {
    typedef typename Vec<T>::type VT;
    constexpr int V = Vec<T>::N;
    const int64_t stride = (int64_t)gridDim.x * EW_BLOCK, Gv = G / V;
    for (int64_t p = (int64_t)blockIdx.x * EW_BLOCK + threadIdx.x; p < Gv; p += stride) {
        T v[C][V], ss[V];
#pragma unroll
        for (int c = 0; c < C; ++c) *reinterpret_cast<VT*>(v[c]) = reinterpret_cast<const VT*>(x + (int64_t)c * G)[p];
#pragma unroll
        for (int k = 0; k < V; ++k) {
            ss[k] = T(0);
#pragma unroll
            for (int c = 0; c < C; ++c) ss[k] += v[c][k] * v[c][k];
        }
#pragma unroll
        for (int k = 0; k < V; ++k) {
            const T f = l21_factor<T>(tsqrt<T>(ss[k]), tau);
#pragma unroll
            for (int c = 0; c < C; ++c) v[c][k] = f * v[c][k];
        }
#pragma unroll
        for (int c = 0; c < C; ++c) reinterpret_cast<VT*>(out + (int64_t)c * G)[p] = *reinterpret_cast<const VT*>(v[c]);
    }
}

template <typename T>
static int pcs_prox_l21_impl(T* out, const T* x, T tau, int64_t G, int C, int cplx, cudaStream_t s) {
    PCS_REQUIRE(G >= 0 && C >= 1 && C <= 4, "pcs_prox_l21: need 1 <= n_comp <= 4 (got %d)", C);
    if (G == 0) return PCS_OK;
    PCS_REQUIRE(out && x, "pcs_prox_l21: null pointer");
    if (!cplx && G % Vec<T>::N == 0 && aligned16(out) && aligned16(x)) {
        const int gridv = grid_for(G / Vec<T>::N, EW_BLOCK, EW_CTAS_PER_SM);
        switch (C) {
            case 1: prox_l21_vec_kernel<T, 1><<<gridv, EW_BLOCK, 0, s>>>(out, x, tau, G); break;
            case 2: prox_l21_vec_kernel<T, 2><<<gridv, EW_BLOCK, 0, s>>>(out, x, tau, G); break;
            case 3: prox_l21_vec_kernel<T, 3><<<gridv, EW_BLOCK, 0, s>>>(out, x, tau, G); break;
            default: prox_l21_vec_kernel<T, 4><<<gridv, EW_BLOCK, 0, s>>>(out, x, tau, G); break;
        }
        return check_launch("pcs_prox_l21");
    }
    int grid = grid_for(G, EW_BLOCK, EW_CTAS_PER_SM);
    switch (C) {
        case 1: prox_l21_kernel<T, 1><<<grid, EW_BLOCK, 0, s>>>(out, x, tau, G, cplx); break;
        case 2: prox_l21_kernel<T, 2><<<grid, EW_BLOCK, 0, s>>>(out, x, tau, G, cplx); break;
        case 3: prox_l21_kernel<T, 3><<<grid, EW_BLOCK, 0, s>>>(out, x, tau, G, cplx); break;
        default: prox_l21_kernel<T, 4><<<grid, EW_BLOCK, 0, s>>>(out, x, tau, G, cplx); break;
    }
    return check_launch("pcs_prox_l21");
}
PCS_EXPORT_REAL(pcs_prox_l21,
                (void* out, const void* x, double tau, int64_t n_groups, int n_comp, int is_complex, void* stream),
                ((T*)out, (const T*)x, (T)tau, n_groups, n_comp, is_complex, (cudaStream_t)stream))

// ------------------------------------------------------------------------------------------------ reductions
constexpr int RD_BLOCK = 256;
constexpr int RD_CTAS_PER_SM = 4;

enum RdOp { RD_SUMSQ, RD_DIFFSUMSQ, RD_SUMABS, RD_DOT, RD_MIN, RD_MAX, RD_NEGLOG, RD_XLOGX, RD_KL };

// Workspace layout (doubles): [0 .. 4*max_blocks) partials, then one 8-byte ticket slot.
extern "C" int64_t pcs_reduce_workspace_bytes(void) {
    return (int64_t)(4 * (int64_t)pcs::num_sms() * 8 + 16) * 8;
}

// one term of the sum (a = x[i], d = y[i] where the reduction has a second operand)
template <int OP> __device__ __forceinline__ double rd_term(double a, double d) {
    if (OP == RD_SUMSQ) return a * a;
    if (OP == RD_DIFFSUMSQ) { const double t = a - d; return t * t; }
    if (OP == RD_SUMABS) return fabs(a);
    if (OP == RD_DOT) return a * d;
    if (OP == RD_NEGLOG) return a > 0.0 ? -log(a) : INFINITY;                       // func/penalty.py:819-822
    if (OP == RD_XLOGX) return a > 0.0 ? a * log(a) : (a == 0.0 ? 0.0 : INFINITY);  // func/penalty.py:899-903
    if (OP == RD_KL) {                                                              // func/loss.py:659-664
        const double z = (a > 0.0 && d > 0.0) ? d * log(d / a) : ((a == 0.0 && d >= 0.0) ? 0.0 : INFINITY);
        return z - d + a;
    }
    return 0.0;
}
template <int OP> constexpr bool rd_binary() { return OP == RD_DIFFSUMSQ || OP == RD_DOT || OP == RD_KL; }

// Grid-stride sum in double.  With 16-byte aligned operands a thread loads RD_U x 16 bytes per operand and trip (all
// loads issued before the first use) and keeps RD_U independent accumulation chains: with one 4-byte load and one chain
// per trip the fp32 sum of squares ran at 3.4 TB/s (profiles/r02_probe_ops.txt).
constexpr int RD_U = 4;
template <typename T, int OP>
__global__ void __launch_bounds__(RD_BLOCK) reduce_kernel(const T* __restrict__ x, const T* __restrict__ y, int64_t n,
                                                          int vec_ok, double* __restrict__ out, double* partials,
                                                          unsigned int* ticket) {
    __shared__ double sm[RD_BLOCK / 32];
    typedef typename Vec<T>::type VT;
    constexpr int V = Vec<T>::N;
    const int64_t stride = (int64_t)gridDim.x * RD_BLOCK;
    const int64_t gtid = (int64_t)blockIdx.x * RD_BLOCK + threadIdx.x;
    double acc[RD_U];
#pragma unroll
    for (int u = 0; u < RD_U; ++u) acc[u] = 0.0;
    int64_t done = 0;                      // elements [0, done) are covered by the vector loop
    if (vec_ok) {
        const int64_t nv = n / V;
        const VT* __restrict__ xv = reinterpret_cast<const VT*>(x);
        const VT* __restrict__ yv = reinterpret_cast<const VT*>(y);
        for (int64_t i0 = gtid; i0 < nv; i0 += RD_U * stride) {
            T a[RD_U][V], d[RD_U][V];
#pragma unroll
            for (int u = 0; u < RD_U; ++u) {
                const int64_t i = i0 + u * stride;
                if (i < nv) {
                    *reinterpret_cast<VT*>(a[u]) = xv[i];
                    if (rd_binary<OP>()) *reinterpret_cast<VT*>(d[u]) = yv[i];
                }
            }
#pragma unroll
            for (int u = 0; u < RD_U; ++u) {
                if (i0 + u * stride < nv) {
#pragma unroll
                    for (int e = 0; e < V; ++e)
                        acc[u] += rd_term<OP>((double)a[u][e], rd_binary<OP>() ? (double)d[u][e] : 0.0);
                }
            }
        }
        done = nv * V;
    }
    for (int64_t i = done + gtid; i < n; i += stride)
        acc[0] += rd_term<OP>((double)x[i], rd_binary<OP>() ? (double)y[i] : 0.0);
    static_assert(RD_U == 4, "the final sum below is written for four chains");
    double tot[1] = {(acc[0] + acc[1]) + (acc[2] + acc[3])};
    grid_sum_finish<1, RD_BLOCK>(tot, partials, ticket, out, sm);
}

template <typename T, int OP>
static int reduce_launch(const T* x, const T* y, int64_t n, double* out, void* ws, cudaStream_t s, const char* what) {
    PCS_REQUIRE(n >= 0 && out && ws, "%s: bad arguments", what);
    PCS_REQUIRE(n == 0 || (x && ((OP != RD_DIFFSUMSQ && OP != RD_DOT && OP != RD_KL) || y)), "%s: null pointer", what);
    const int vec_ok = ((uintptr_t)x % 16 == 0) && (y == nullptr || (uintptr_t)y % 16 == 0);
    int grid = grid_for(vec_ok ? (n + Vec<T>::N - 1) / Vec<T>::N : n, RD_BLOCK, RD_CTAS_PER_SM);
    double* partials = (double*)ws;
    unsigned int* ticket = (unsigned int*)(partials + 4 * (int64_t)pcs::num_sms() * 8);
    reduce_kernel<T, OP><<<grid, RD_BLOCK, 0, s>>>(x, y, n, vec_ok, out, partials, ticket);
    return check_launch(what);
}

#define PCS_REDUCE1(name, OP)                                                                                  \
    template <typename T> static int name##_impl(const T* x, int64_t n, double* out, void* ws, cudaStream_t s) { \
        return reduce_launch<T, OP>(x, nullptr, n, out, ws, s, #name);                                         \
    }                                                                                                          \
    PCS_EXPORT_REAL(name, (const void* x, int64_t n, void* out_f64, void* workspace, void* stream),             \
                    ((const T*)x, n, (double*)out_f64, workspace, (cudaStream_t)stream))
#define PCS_REDUCE2(name, OP)                                                                                  \
    template <typename T>                                                                                      \
    static int name##_impl(const T* x, const T* y, int64_t n, double* out, void* ws, cudaStream_t s) {         \
        return reduce_launch<T, OP>(x, y, n, out, ws, s, #name);                                               \
    }                                                                                                          \
    PCS_EXPORT_REAL(name, (const void* x, const void* y, int64_t n, void* out_f64, void* workspace, void* stream), \
                    ((const T*)x, (const T*)y, n, (double*)out_f64, workspace, (cudaStream_t)stream))

PCS_REDUCE1(pcs_reduce_sumsq, RD_SUMSQ)          // ||x||^2           (np.linalg.norm, opt/proxalgs.py:372-383)
PCS_REDUCE1(pcs_reduce_sumabs, RD_SUMABS)        // sum |x|           (L1Norm.__call__, func/penalty.py:240-241)
PCS_REDUCE2(pcs_reduce_diffsumsq, RD_DIFFSUMSQ)  // ||x - y||^2
PCS_REDUCE2(pcs_reduce_dot, RD_DOT)              // <x, y>
PCS_REDUCE1(pcs_reduce_neglog, RD_NEGLOG)        // -sum log x (inf unless x > 0)            LogBarrier.__call__
PCS_REDUCE1(pcs_reduce_xlogx, RD_XLOGX)          // sum x log x (0 at 0, inf below)          ShannonEntropy.__call__
PCS_REDUCE2(pcs_reduce_kl, RD_KL)                // sum y log(y/x) - y + x                   KLDivergence.__call__

// min and max in one pass (IndicatorFunctional.__call__ conditions, func/penalty.py:610,666)
template <typename T>
__global__ void __launch_bounds__(RD_BLOCK) minmax_kernel(const T* __restrict__ x, int64_t n, double* out,
                                                          double* partials, unsigned int* ticket) {
    __shared__ double smn[RD_BLOCK / 32], smx[RD_BLOCK / 32];
    __shared__ bool is_last;
    const int64_t stride = (int64_t)gridDim.x * RD_BLOCK;
    double mn = INFINITY, mx = -INFINITY;
    for (int64_t i = (int64_t)blockIdx.x * RD_BLOCK + threadIdx.x; i < n; i += stride) {
        double a = (double)x[i];
        mn = fmin(mn, a);
        mx = fmax(mx, a);
    }
    auto blockminmax = [&]() {
        for (int o = 16; o > 0; o >>= 1) {
            mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
            mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        }
        if ((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = mn; smx[threadIdx.x >> 5] = mx; }
        __syncthreads();
        if (threadIdx.x < 32) {
            mn = threadIdx.x < RD_BLOCK / 32 ? smn[threadIdx.x] : INFINITY;
            mx = threadIdx.x < RD_BLOCK / 32 ? smx[threadIdx.x] : -INFINITY;
            for (int o = 16; o > 0; o >>= 1) {
                mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
                mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            }
        }
    };
    blockminmax();
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = mn;
        partials[gridDim.x + blockIdx.x] = mx;
        __threadfence();
        is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        mn = INFINITY; mx = -INFINITY;
        for (int i = threadIdx.x; i < gridDim.x; i += RD_BLOCK) {
            mn = fmin(mn, __ldcg(&partials[i]));
            mx = fmax(mx, __ldcg(&partials[gridDim.x + i]));
        }
        __syncthreads();
        blockminmax();
        if (threadIdx.x == 0) { out[0] = mn; out[1] = mx; *ticket = 0u; }
    }
}

template <typename T> static int pcs_reduce_minmax_impl(const T* x, int64_t n, double* out, void* ws, cudaStream_t s) {
    PCS_REQUIRE(n >= 0 && out && ws && (n == 0 || x), "pcs_reduce_minmax: bad arguments");
    int grid = grid_for(n, RD_BLOCK, RD_CTAS_PER_SM);
    double* partials = (double*)ws;
    unsigned int* ticket = (unsigned int*)(partials + 4 * (int64_t)pcs::num_sms() * 8);
    minmax_kernel<T><<<grid, RD_BLOCK, 0, s>>>(x, n, out, partials, ticket);
    return check_launch("pcs_reduce_minmax");
}
PCS_EXPORT_REAL(pcs_reduce_minmax, (const void* x, int64_t n, void* out2_f64, void* workspace, void* stream),
                ((const T*)x, n, (double*)out2_f64, workspace, (cudaStream_t)stream))

// One Newton step on g(t) = sum_i max(|x_i| - t, 0) - radius - t*c0 (convex, piecewise linear, decreasing), started at
// t = 0: t' = (S(t) - radius) / (C(t) + c0) with S = sum of |x_i| over {|x_i| > t}, C = the size of that set.  The steps
// increase monotonically and reach the EXACT root after finitely many (the active set can only shrink; it is Michelot's
// algorithm).  radius > 0, c0 = 0: the threshold of the l1-ball projection, which the reference brackets with brentq
// (math/prox.py:160-164).  radius = 0, c0 = 1/(2 tau): the threshold of the squared-l1 prox, which the reference finds
// by a sort and a cumulative sum (func/penalty.py:310-316).  The step reads the previous (S, C) from `state[2*src]`
// on the device (src < 0: t = 0), so a batch of steps needs no host round trip.
template <typename T>
__global__ void __launch_bounds__(RD_BLOCK) threshold_newton_kernel(const T* __restrict__ x, int64_t n, double radius,
                                                                    double c0, double* state, int src, int dst,
                                                                    double* partials, unsigned int* ticket) {
    __shared__ double sm[RD_BLOCK / 32];
    double t = 0.0;
    if (src >= 0) t = fmax((__ldcg(&state[2 * src]) - radius) / (__ldcg(&state[2 * src + 1]) + c0), 0.0);
    const int64_t stride = (int64_t)gridDim.x * RD_BLOCK;
    double acc[2] = {0.0, 0.0};
    for (int64_t i = (int64_t)blockIdx.x * RD_BLOCK + threadIdx.x; i < n; i += stride) {
        const double a = fabs((double)x[i]);
        if (a > t) { acc[0] += a; acc[1] += 1.0; }
    }
    grid_sum_finish<2, RD_BLOCK>(acc, partials, ticket, state + 2 * dst, sm);
}

template <typename T>
static int pcs_threshold_newton_step_impl(const T* x, int64_t n, double radius, double c0, double* state, int src,
                                          int dst, void* ws, cudaStream_t s) {
    PCS_REQUIRE(n >= 0 && state && ws && (n == 0 || x), "pcs_threshold_newton_step: bad arguments");
    PCS_REQUIRE(src >= -1 && src <= 1 && dst >= 0 && dst <= 1 && src != dst, "pcs_threshold_newton_step: slots are 0/1");
    int grid = grid_for(n, RD_BLOCK, RD_CTAS_PER_SM);
    double* partials = (double*)ws;
    unsigned int* ticket = (unsigned int*)(partials + 4 * (int64_t)pcs::num_sms() * 8);
    threshold_newton_kernel<T><<<grid, RD_BLOCK, 0, s>>>(x, n, radius, c0, state, src, dst, partials, ticket);
    return check_launch("pcs_threshold_newton_step");
}
PCS_EXPORT_REAL(pcs_threshold_newton_step,
                (const void* x, int64_t n, double radius, double c0, void* state4_f64, int src_slot, int dst_slot,
                 void* workspace, void* stream),
                ((const T*)x, n, radius, c0, (double*)state4_f64, src_slot, dst_slot, workspace, (cudaStream_t)stream))

// sum over groups of the group 2-norm, planes layout (L21Norm.__call__, func/penalty.py:548-549)
template <typename T>
__global__ void __launch_bounds__(RD_BLOCK) sumgroupnorm_kernel(const T* __restrict__ x, int64_t G, int C, int cplx,
                                                                int vec_ok, double* out, double* partials,
                                                                unsigned int* ticket) {
    __shared__ double sm[RD_BLOCK / 32];
    const int64_t stride = (int64_t)gridDim.x * RD_BLOCK;
    const int64_t e = cplx ? 2 : 1;
    double acc[1] = {0.0};
    // group norm in the working precision (as np.linalg.norm of a float32 group is), sum over groups in double; the fp64
    // square root per group of the all-double version made this kernel compute bound (1.07 TB/s in fp32)
    constexpr int U = 4;                   // groups per thread and trip: U x C independent loads in flight
    if (vec_ok) {
        // real planes with 16-byte aligned rows: a thread owns V consecutive groups (one 16-byte load per plane) and
        // two such packs per trip
        typedef typename Vec<T>::type VT;
        constexpr int V = Vec<T>::N, UV = 2;
        const int64_t Gv = G / V;
        for (int64_t p0 = (int64_t)blockIdx.x * RD_BLOCK + threadIdx.x; p0 < Gv; p0 += UV * stride) {
            T ss[UV][V];
#pragma unroll
            for (int u = 0; u < UV; ++u)
#pragma unroll
                for (int k = 0; k < V; ++k) ss[u][k] = T(0);
            for (int c = 0; c < C; ++c) {
                T v[UV][V];
#pragma unroll
                for (int u = 0; u < UV; ++u) {
                    const int64_t pp = p0 + u * stride;
                    if (pp < Gv) *reinterpret_cast<VT*>(v[u]) = reinterpret_cast<const VT*>(x + (int64_t)c * G)[pp];
                    else
#pragma unroll
                        for (int k = 0; k < V; ++k) v[u][k] = T(0);
                }
#pragma unroll
                for (int u = 0; u < UV; ++u)
#pragma unroll
                    for (int k = 0; k < V; ++k) ss[u][k] = fma(v[u][k], v[u][k], ss[u][k]);
            }
#pragma unroll
            for (int u = 0; u < UV; ++u)
#pragma unroll
                for (int k = 0; k < V; ++k) acc[0] += (double)tsqrt<T>(ss[u][k]);
        }
        grid_sum_finish<1, RD_BLOCK>(acc, partials, ticket, out, sm);
        return;
    }
    for (int64_t g0 = (int64_t)blockIdx.x * RD_BLOCK + threadIdx.x; g0 < G; g0 += U * stride) {
        T ss[U];
#pragma unroll
        for (int u = 0; u < U; ++u) ss[u] = T(0);
        for (int c = 0; c < C; ++c)
            for (int k = 0; k < e; ++k) {
                T v[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int64_t g = g0 + u * stride;
                    v[u] = g < G ? x[((int64_t)c * G + g) * e + k] : T(0);
                }
#pragma unroll
                for (int u = 0; u < U; ++u) ss[u] = fma(v[u], v[u], ss[u]);
            }
#pragma unroll
        for (int u = 0; u < U; ++u) acc[0] += (double)tsqrt<T>(ss[u]);
    }
    grid_sum_finish<1, RD_BLOCK>(acc, partials, ticket, out, sm);
}

template <typename T>
static int pcs_reduce_sumgroupnorm_impl(const T* x, int64_t G, int C, int cplx, double* out, void* ws, cudaStream_t s) {
    PCS_REQUIRE(G >= 0 && C >= 1 && out && ws && (G == 0 || x), "pcs_reduce_sumgroupnorm: bad arguments");
    // vector path: real data, every plane 16-byte aligned and a whole number of vectors long
    const int vec_ok = !cplx && G > 0 && ((uintptr_t)x % 16 == 0) && (G % Vec<T>::N == 0);
    int grid = grid_for(vec_ok ? G / Vec<T>::N : G, RD_BLOCK, RD_CTAS_PER_SM);
    double* partials = (double*)ws;
    unsigned int* ticket = (unsigned int*)(partials + 4 * (int64_t)pcs::num_sms() * 8);
    sumgroupnorm_kernel<T><<<grid, RD_BLOCK, 0, s>>>(x, G, C, cplx, vec_ok, out, partials, ticket);
    return check_launch("pcs_reduce_sumgroupnorm");
}
PCS_EXPORT_REAL(pcs_reduce_sumgroupnorm,
                (const void* x, int64_t n_groups, int n_comp, int is_complex, void* out_f64, void* workspace,
                 void* stream),
                ((const T*)x, n_groups, n_comp, is_complex, (double*)out_f64, workspace, (cudaStream_t)stream))

// ---- L21Norm for ARBITRARY label arrays (func/penalty.py:532-560): groups in CSR form ----------------------------------
// `members` lists the coordinates of group 0, then of group 1, ... (stable order); group g owns members[offsets[g] ..
// offsets[g+1]).  Groups may be ragged and of any size.  One warp per group: lanes stride over the members, the sum of
// squares is accumulated in double and warp-reduced (deterministic), then
//   * prox: out[i] = clip(1 - tau / ||x_g||, 0) * x[i] for every member i            (penalty.py:551-557)
//   * value: sum_g ||x_g||_2                                                         (penalty.py:548-549)
template <typename T>
__device__ __forceinline__ double l21_csr_group_sumsq(const T* __restrict__ x, const int64_t* __restrict__ members,
                                                      int64_t a, int64_t b, int e, int lane) {
    double ss = 0.0;
    for (int64_t k = a + lane; k < b; k += 32) {
        const int64_t i = members[k] * e;
        for (int c = 0; c < e; ++c) {
            const double v = (double)x[i + c];
            ss = fma(v, v, ss);
        }
    }
    return warp_sum(ss);
}

template <typename T>
__global__ void __launch_bounds__(EW_BLOCK) prox_l21_csr_kernel(T* __restrict__ out, const T* __restrict__ x, T tau,
                                                                const int64_t* __restrict__ members,
                                                                const int64_t* __restrict__ offsets, int64_t G, int cplx) {
    const int lane = threadIdx.x & 31, e = cplx ? 2 : 1;
    const int64_t nwarps = (int64_t)gridDim.x * (EW_BLOCK / 32);
    for (int64_t g = (int64_t)blockIdx.x * (EW_BLOCK / 32) + (threadIdx.x >> 5); g < G; g += nwarps) {
        const int64_t a = offsets[g], b = offsets[g + 1];
        const double ss = l21_csr_group_sumsq<T>(x, members, a, b, e, lane);
        const T f = l21_factor<T>((T)sqrt(ss), tau);
        for (int64_t k = a + lane; k < b; k += 32) {
            const int64_t i = members[k] * e;
            for (int c = 0; c < e; ++c) out[i + c] = f * x[i + c];
        }
    }
}

template <typename T>
static int pcs_prox_l21_csr_impl(T* out, const T* x, T tau, const int64_t* members, const int64_t* offsets, int64_t G,
                                 int cplx, cudaStream_t s) {
    PCS_REQUIRE(G >= 0, "pcs_prox_l21_csr: negative group count");
    if (G == 0) return PCS_OK;
    PCS_REQUIRE(out && x && members && offsets, "pcs_prox_l21_csr: null pointer");
    prox_l21_csr_kernel<T><<<grid_for(G * 32, EW_BLOCK, EW_CTAS_PER_SM), EW_BLOCK, 0, s>>>(out, x, tau, members, offsets,
                                                                                          G, cplx);
    return check_launch("pcs_prox_l21_csr");
}
PCS_EXPORT_REAL(pcs_prox_l21_csr,
                (void* out, const void* x, double tau, const void* members_i64, const void* offsets_i64,
                 int64_t n_groups, int is_complex, void* stream),
                ((T*)out, (const T*)x, (T)tau, (const int64_t*)members_i64, (const int64_t*)offsets_i64, n_groups,
                 is_complex, (cudaStream_t)stream))

template <typename T>
__global__ void __launch_bounds__(RD_BLOCK) sumgroupnorm_csr_kernel(const T* __restrict__ x,
                                                                    const int64_t* __restrict__ members,
                                                                    const int64_t* __restrict__ offsets, int64_t G,
                                                                    int cplx, double* out, double* partials,
                                                                    unsigned int* ticket) {
    __shared__ double sm[RD_BLOCK / 32];
    const int lane = threadIdx.x & 31, e = cplx ? 2 : 1;
    const int64_t nwarps = (int64_t)gridDim.x * (RD_BLOCK / 32);
    double acc[1] = {0.0};
    for (int64_t g = (int64_t)blockIdx.x * (RD_BLOCK / 32) + (threadIdx.x >> 5); g < G; g += nwarps) {
        const double ss = l21_csr_group_sumsq<T>(x, members, offsets[g], offsets[g + 1], e, lane);
        if (lane == 0) acc[0] += sqrt(ss);
    }
    grid_sum_finish<1, RD_BLOCK>(acc, partials, ticket, out, sm);
}

template <typename T>
static int pcs_reduce_sumgroupnorm_csr_impl(const T* x, const int64_t* members, const int64_t* offsets, int64_t G,
                                            int cplx, double* out, void* ws, cudaStream_t s) {
    PCS_REQUIRE(G >= 0 && out && ws && (G == 0 || (x && members && offsets)), "pcs_reduce_sumgroupnorm_csr: bad arguments");
    double* partials = (double*)ws;
    unsigned int* ticket = (unsigned int*)(partials + 4 * (int64_t)pcs::num_sms() * 8);
    sumgroupnorm_csr_kernel<T><<<grid_for(G * 32, RD_BLOCK, RD_CTAS_PER_SM), RD_BLOCK, 0, s>>>(x, members, offsets, G, cplx,
                                                                                              out, partials, ticket);
    return check_launch("pcs_reduce_sumgroupnorm_csr");
}
PCS_EXPORT_REAL(pcs_reduce_sumgroupnorm_csr,
                (const void* x, const void* members_i64, const void* offsets_i64, int64_t n_groups, int is_complex,
                 void* out_f64, void* workspace, void* stream),
                ((const T*)x, (const int64_t*)members_i64, (const int64_t*)offsets_i64, n_groups, is_complex,
                 (double*)out_f64, workspace, (cudaStream_t)stream))

// out = x * f(s) where s is a device-resident double: mode 0: f = clip(1 - tau/sqrt(s), 0) (L2Norm.prox,
// func/penalty.py:23-70 via proj_l2_ball math/prox.py:207-210 -> x - tau*proj(x/tau) = x*max(1 - tau/||x||, 0));
// mode 1: f = (sqrt(s) <= radius ? 1 : radius/sqrt(s)) (proj_l2_ball); mode 2: f = p/sqrt(s).  s holds ||x||^2.
template <typename T>
__global__ void __launch_bounds__(EW_BLOCK) scale_by_norm_kernel(T* __restrict__ out, const T* __restrict__ x,
                                                                 const double* __restrict__ s, double p, int mode,
                                                                 int64_t n) {
    const double nrm = sqrt(s[0]);
    double f;
    if (mode == 0) {
        // reference: x - tau * proj_l2_ball(x / tau, 1): ||x/tau|| <= 1 -> x - x = 0, else x - tau*(x/tau)/||x/tau||
        f = (nrm / p <= 1.0) ? 0.0 : 1.0 - p / nrm;
    } else if (mode == 1) {
        f = (nrm <= p) ? 1.0 : p / nrm;
    } else {
        f = nrm > 0.0 ? p / nrm : 0.0;   // mode 2: rescale to norm p (power iteration)
    }
    const T ft = (T)f;
    const int64_t stride = (int64_t)gridDim.x * EW_BLOCK;
    for (int64_t i = (int64_t)blockIdx.x * EW_BLOCK + threadIdx.x; i < n; i += stride) out[i] = ft * x[i];
}

template <typename T>
static int pcs_scale_by_norm_impl(T* out, const T* x, const double* sumsq, double p, int mode, int64_t n,
                                  cudaStream_t s) {
    PCS_REQUIRE(n >= 0 && (n == 0 || (out && x && sumsq)) && (mode >= 0 && mode <= 2), "pcs_scale_by_norm: bad arguments");
    if (n == 0) return PCS_OK;
    scale_by_norm_kernel<T><<<grid_for(n, EW_BLOCK, EW_CTAS_PER_SM), EW_BLOCK, 0, s>>>(out, x, sumsq, p, mode, n);
    return check_launch("pcs_scale_by_norm");
}
PCS_EXPORT_REAL(pcs_scale_by_norm,
                (void* out, const void* x, const void* sumsq_f64, double param, int mode, int64_t n, void* stream),
                ((T*)out, (const T*)x, (const double*)sumsq_f64, param, mode, n, (cudaStream_t)stream))

// L2Norm.prox (func/penalty.py:23-70 through LpNorm.prox, func/base.py:215-240) and proj_l2_ball (math/prox.py:167-210)
// under the names SURVEY.md 8(b) lists: thin wrappers of pcs_scale_by_norm (sumsq_f64 = ||x||^2 from pcs_reduce_sumsq).
template <typename T>
static int pcs_prox_l2_impl(T* out, const T* x, const double* sumsq, double tau, int64_t n, cudaStream_t s) {
    return pcs_scale_by_norm_impl<T>(out, x, sumsq, tau, 0, n, s);
}
PCS_EXPORT_REAL(pcs_prox_l2, (void* out, const void* x, const void* sumsq_f64, double tau, int64_t n, void* stream),
                ((T*)out, (const T*)x, (const double*)sumsq_f64, tau, n, (cudaStream_t)stream))
template <typename T>
static int pcs_proj_l2_ball_impl(T* out, const T* x, const double* sumsq, double radius, int64_t n, cudaStream_t s) {
    return pcs_scale_by_norm_impl<T>(out, x, sumsq, radius, 1, n, s);
}
PCS_EXPORT_REAL(pcs_proj_l2_ball, (void* out, const void* x, const void* sumsq_f64, double radius, int64_t n, void* stream),
                ((T*)out, (const T*)x, (const double*)sumsq_f64, radius, n, (cudaStream_t)stream))
