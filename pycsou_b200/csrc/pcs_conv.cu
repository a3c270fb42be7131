// libpcsb200: compact-filter convolution / correlation with zero "same" boundaries (sm_100a).
//
// Replaces pylops.signalprocessing.Convolve1D / Convolve2D -> scipy.signal.convolve / correlate reached from
// pycsou/linop/conv.py:163, 294.  Semantics (oracle/pylops_standin.py):
//   forward  y[i] = sum_k h[k] x[i + off - k]        (filter centred at `off`, zero extension)
//   adjoint  y[i] = sum_k h[k] x[i + k - off]        (correlation = exact transpose)
// 2-D: the same along both axes.  Three kernels:
//   conv1d       one axis of an (outer, n, inner) view; taps in shared memory, inputs through L1.
//   conv2d dir.  dense kh x kw filter, shared-memory tile with halo, 4 outputs per thread.
//   conv2d sep.  rank-1 filter (h = hcol (x) hrow): row pass and column pass through one shared-memory tile;
//                optional fused "- y" (residual r = A x - y of the data term).
#include "pcs_common.cuh"
#include <type_traits>
#include <stdlib.h>

using namespace pcs;

// ------------------------------------------------------------------------------------------------ 1-D
constexpr int C1_BLOCK = 256;

template <typename T>
__global__ void __launch_bounds__(C1_BLOCK) conv1d_kernel(T* __restrict__ out, const T* __restrict__ x,
                                                          const T* __restrict__ h, int nh, int off, int sgn,
                                                          int64_t outer, int64_t n, int64_t inner) {
    extern __shared__ unsigned char smem_raw[];
    T* taps = reinterpret_cast<T*>(smem_raw);
    for (int k = threadIdx.x; k < nh; k += C1_BLOCK) taps[k] = h[k];
    __syncthreads();
    const int64_t total = outer * n * inner, stride = (int64_t)gridDim.x * C1_BLOCK;
    for (int64_t f = (int64_t)blockIdx.x * C1_BLOCK + threadIdx.x; f < total; f += stride) {
        const int64_t i = (f / inner) % n;
        // position of tap k: i + sgn*(k - off); valid k range so that 0 <= pos < n
        int k_lo, k_hi;
        if (sgn > 0) {
            k_lo = (int)tmax<int64_t>(0, off - i);
            k_hi = (int)tmin<int64_t>(nh - 1, n - 1 - i + off);
        } else {
            k_lo = (int)tmax<int64_t>(0, i + off - (n - 1));
            k_hi = (int)tmin<int64_t>(nh - 1, i + off);
        }
        T acc = T(0);
        const T* base = x + f + (int64_t)sgn * (k_lo - off) * inner;
        const int64_t step = (int64_t)sgn * inner;
        for (int k = k_lo; k <= k_hi; ++k, base += step) acc += taps[k] * base[0];
        out[f] = acc;
    }
}

template <typename T>
static int conv1d_launch(T* out, const T* x, const T* h, int nh, int off, int sgn, int64_t outer, int64_t n,
                         int64_t inner, cudaStream_t s, const char* what) {
    PCS_REQUIRE(outer >= 0 && n >= 0 && inner >= 0 && nh >= 1, "%s: bad dimensions", what);
    PCS_REQUIRE(off >= 0 && off < nh, "%s: offset must satisfy 0 <= offset < filter size", what);
    PCS_REQUIRE((size_t)nh * sizeof(T) <= 48 * 1024, "%s: filter too long for the direct kernel (%d taps)", what, nh);
    const int64_t total = outer * n * inner;
    if (total == 0) return PCS_OK;
    PCS_REQUIRE(out && x && h && out != x, "%s: null or aliased pointer", what);
    conv1d_kernel<T><<<grid_for(total, C1_BLOCK, 8), C1_BLOCK, nh * sizeof(T), s>>>(out, x, h, nh, off, sgn, outer, n,
                                                                                  inner);
    return check_launch(what);
}
#define C1_PARAMS                                                                                               \
    (void* out, const void* x, const void* h, int nh, int offset, int64_t outer, int64_t n, int64_t inner, void* stream)
template <typename T>
static int pcs_conv1d_fwd_impl(T* o, const T* x, const T* h, int nh, int off, int64_t a, int64_t n, int64_t b,
                               cudaStream_t s) {
    return conv1d_launch<T>(o, x, h, nh, off, -1, a, n, b, s, "pcs_conv1d_fwd");
}
PCS_EXPORT_REAL(pcs_conv1d_fwd, C1_PARAMS,
                ((T*)out, (const T*)x, (const T*)h, nh, offset, outer, n, inner, (cudaStream_t)stream))
template <typename T>
static int pcs_conv1d_adj_impl(T* o, const T* x, const T* h, int nh, int off, int64_t a, int64_t n, int64_t b,
                               cudaStream_t s) {
    return conv1d_launch<T>(o, x, h, nh, off, +1, a, n, b, s, "pcs_conv1d_adj");
}
PCS_EXPORT_REAL(pcs_conv1d_adj, C1_PARAMS,
                ((T*)out, (const T*)x, (const T*)h, nh, offset, outer, n, inner, (cudaStream_t)stream))

// ------------------------------------------------------------------------------------------------ 2-D tiles
constexpr int C2_TW = 64;       // tile width  (columns)
constexpr int C2_TH = 32;       // tile height (rows)
constexpr int C2_BX = 64, C2_BY = 4;   // 256 threads; each thread owns C2_TH / C2_BY = 8 rows of one column
constexpr int C2_RPT = C2_TH / C2_BY;
constexpr size_t C2_SMEM_LIMIT = 200 * 1024;

// Load the input tile with its halo; (r0,c0) = image coordinates of smem (0,0).  Outside the image -> 0.
template <typename T>
__device__ __forceinline__ void load_tile(T* sm, int sh, int sw, const T* __restrict__ x, int64_t H, int64_t W,
                                          int64_t r0, int64_t c0) {
    const int tid = threadIdx.y * C2_BX + threadIdx.x;
    for (int e = tid; e < sh * sw; e += C2_BX * C2_BY) {
        const int lr = e / sw, lc = e - lr * sw;
        const int64_t r = r0 + lr, c = c0 + lc;
        sm[e] = (r >= 0 && r < H && c >= 0 && c < W) ? x[r * W + c] : T(0);
    }
}

// dense filter.  sgn = -1: convolution (forward), +1: correlation (adjoint).
template <typename T>
__global__ void __launch_bounds__(C2_BX * C2_BY) conv2d_direct_kernel(T* __restrict__ out, const T* __restrict__ x,
                                                                      const T* __restrict__ h, int kh, int kw,
                                                                      int off0, int off1, int sgn, int64_t H,
                                                                      int64_t W) {
    extern __shared__ unsigned char smem_raw[];
    T* taps = reinterpret_cast<T*>(smem_raw);
    T* tile = taps + kh * kw;
    const int sh = C2_TH + kh - 1, sw = C2_TW + kw - 1;
    const int tid = threadIdx.y * C2_BX + threadIdx.x;
    // store taps in "correlation order": out[r,c] = sum_{a,b} g[a,b] tile[r+a, c+b]
    for (int e = tid; e < kh * kw; e += C2_BX * C2_BY) {
        const int a = e / kw, b = e - a * kw;
        taps[e] = (sgn > 0) ? h[e] : h[(kh - 1 - a) * kw + (kw - 1 - b)];
    }
    const int ca = (sgn > 0) ? off0 : kh - 1 - off0, cb = (sgn > 0) ? off1 : kw - 1 - off1;
    const int64_t R0 = (int64_t)blockIdx.y * C2_TH, C0 = (int64_t)blockIdx.x * C2_TW;
    load_tile<T>(tile, sh, sw, x, H, W, R0 - ca, C0 - cb);
    __syncthreads();
    T acc[C2_RPT];
#pragma unroll
    for (int j = 0; j < C2_RPT; ++j) acc[j] = T(0);
    const int lc = threadIdx.x;
    for (int a = 0; a < kh; ++a) {
        for (int b = 0; b < kw; ++b) {
            const T g = taps[a * kw + b];
#pragma unroll
            for (int j = 0; j < C2_RPT; ++j) acc[j] += g * tile[(threadIdx.y + j * C2_BY + a) * sw + lc + b];
        }
    }
    const int64_t c = C0 + lc;
    if (c < W) {
#pragma unroll
        for (int j = 0; j < C2_RPT; ++j) {
            const int64_t r = R0 + threadIdx.y + j * C2_BY;
            if (r < H) out[r * W + c] = acc[j];
        }
    }
}

// Register-tiled form for filters up to KM x KM taps (KM = 7, 15): a thread owns 8 rows x 2 adjacent columns.  For every
// pair of sample columns it loads a 22-sample run of each (one 64/128-bit shared load per row) and feeds 4 FMAs per
// sample and filter row: ~0.15 shared loads per FMA instead of 1.1 in the kernel above (which stays for larger filters).
// The filter sits in shared memory in correlation order, embedded in a KM x (KM + 2) array with a zero column on either
// side, so the two outputs of a thread read taps b-1, b, b+1 without bounds checks.
constexpr int C2R_TW = 64, C2R_TH = 64, C2R_RP = 8;
#ifndef C2R_PACKED_F32
#define C2R_PACKED_F32 1
#endif
__device__ __forceinline__ unsigned long long c2_pack2(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void c2_unpack2(unsigned long long v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ unsigned long long c2_fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
#ifndef C2R_MINB_F32
#define C2R_MINB_F32 3
#endif
template <typename T, int KM>
__global__ void __launch_bounds__(256, sizeof(T) == 4 ? C2R_MINB_F32 : 1) conv2d_reg_kernel(T* __restrict__ out, const T* __restrict__ x,
                                                         const T* __restrict__ h, int kh, int kw, int off0, int off1,
                                                         int sgn, const T* __restrict__ sub, int64_t H, int64_t W) {
    constexpr int KP = KM + 2, SH = C2R_TH + KM - 1, SW = C2R_TW + KM + 1, NS = C2R_RP + KM - 1;
    typedef typename std::conditional<sizeof(T) == 4, float2, double2>::type V2;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* taps = reinterpret_cast<T*>(smem_raw);                 // [KM][KP], column index b + 1
    T* tile = taps + ((KM * KP + 3) / 4) * 4;                 // [SH][SW], 16-byte aligned
    const int tid = threadIdx.x;
    for (int e = tid; e < KM * KP; e += 256) {
        const int a = e / KP, b = e - a * KP - 1;
        T g = T(0);
        if (a < kh && b >= 0 && b < kw) g = (sgn > 0) ? h[a * kw + b] : h[(kh - 1 - a) * kw + (kw - 1 - b)];
        taps[e] = g;
    }
    const int ca = (sgn > 0) ? off0 : kh - 1 - off0, cb = (sgn > 0) ? off1 : kw - 1 - off1;
    const int64_t R0 = (int64_t)blockIdx.y * C2R_TH, C0 = (int64_t)blockIdx.x * C2R_TW;
    {   // tile load: warp <-> rows, lane <-> columns (row pointer and row validity computed once per row)
        const int lane = tid & 31, wrp = tid >> 5;
        for (int lr = wrp; lr < SH; lr += 8) {
            const int64_t r = R0 - ca + lr;
            const bool rv = r >= 0 && r < H;
            const T* xr = x + (rv ? r : 0) * W;
            T* trw = tile + lr * SW;
#pragma unroll
            for (int lc = lane; lc < SW; lc += 32) {
                const int64_t c = C0 - cb + lc;
                trw[lc] = (rv && c >= 0 && c < W) ? xr[c] : T(0);
            }
        }
    }
    __syncthreads();
    const int tx = tid & 31, ty = tid >> 5;
    T acc0[C2R_RP], acc1[C2R_RP];
#pragma unroll
    for (int j = 0; j < C2R_RP; ++j) acc0[j] = acc1[j] = T(0);
    const T* trow = tile + (ty * C2R_RP) * SW + 2 * tx;
    if constexpr (sizeof(T) == 4 && C2R_PACKED_F32) {
        // packed fp32 (fma.rn.f32x2 = SASS FFMA2): one instruction updates the two outputs (col, col+1) of a row:
        // acc2[j] += (s0, s0) * (g[b], g[b-1]) + (s1, s1) * (g[b+1], g[b]); half the FMA issue slots of the scalar form
        unsigned long long acc2[C2R_RP];
#pragma unroll
        for (int j = 0; j < C2R_RP; ++j) acc2[j] = 0ull;
#pragma unroll 1
        for (int beta = 0; beta < KM + 1; beta += 2) {
            unsigned long long p0[NS], p1[NS];
#pragma unroll
            for (int j = 0; j < NS; ++j) {
                const float2 v = *reinterpret_cast<const float2*>(trow + j * SW + beta);
                p0[j] = c2_pack2(v.x, v.x);
                p1[j] = c2_pack2(v.y, v.y);
            }
            const T* tp = taps + beta;
#pragma unroll
            for (int a = 0; a < KM; ++a) {
                const float gm = tp[a * KP], g0 = tp[a * KP + 1], gp = tp[a * KP + 2];
                const unsigned long long t0 = c2_pack2(g0, gm), t1 = c2_pack2(gp, g0);
#pragma unroll
                for (int j = 0; j < C2R_RP; ++j) {
                    acc2[j] = c2_fma2(p0[j + a], t0, acc2[j]);
                    acc2[j] = c2_fma2(p1[j + a], t1, acc2[j]);
                }
            }
        }
#pragma unroll
        for (int j = 0; j < C2R_RP; ++j) {
            float lo, hi;
            c2_unpack2(acc2[j], lo, hi);
            acc0[j] = lo; acc1[j] = hi;
        }
    } else
#pragma unroll 1
    for (int beta = 0; beta < KM + 1; beta += 2) {
        T s0[NS], s1[NS];
#pragma unroll
        for (int j = 0; j < NS; ++j) {
            const V2 v = *reinterpret_cast<const V2*>(trow + j * SW + beta);
            s0[j] = v.x; s1[j] = v.y;
        }
        const T* tp = taps + beta;                            // tp[a * KP + 0 / 1 / 2] = taps b = beta-1 / beta / beta+1
#pragma unroll
        for (int a = 0; a < KM; ++a) {
            const T gm = tp[a * KP], g0 = tp[a * KP + 1], gp = (beta + 1 < KM) ? tp[a * KP + 2] : T(0);
#pragma unroll
            for (int j = 0; j < C2R_RP; ++j) {
                acc0[j] += g0 * s0[j + a];
                acc0[j] += gp * s1[j + a];
                acc1[j] += gm * s0[j + a];
                acc1[j] += g0 * s1[j + a];
            }
        }
    }
    const int64_t c = C0 + 2 * tx;
    const bool pair = (W % 2 == 0) && c + 1 < W;
#pragma unroll
    for (int j = 0; j < C2R_RP; ++j) {
        const int64_t r = R0 + ty * C2R_RP + j;
        if (r >= H || c >= W) continue;
        T v0 = acc0[j], v1 = acc1[j];
        const int64_t idx = r * W + c;
        if (sub != nullptr) { v0 -= sub[idx]; if (c + 1 < W) v1 -= sub[idx + 1]; }
        if (pair) {
            V2 o; o.x = v0; o.y = v1;
            *reinterpret_cast<V2*>(out + idx) = o;
        } else {
            out[idx] = v0;
            if (c + 1 < W) out[idx + 1] = v1;
        }
    }
}
template <typename T, int KM>
static int conv2d_reg_launch(T* out, const T* x, const T* h, int kh, int kw, int off0, int off1, int sgn, const T* sub,
                             int64_t H, int64_t W, cudaStream_t s, const char* what) {
    constexpr int KP = KM + 2, SH = C2R_TH + KM - 1, SW = C2R_TW + KM + 1;
    const size_t smem = ((size_t)((KM * KP + 3) / 4) * 4 + (size_t)SH * SW) * sizeof(T);
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(conv2d_reg_kernel<T, KM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    dim3 grid((unsigned)((W + C2R_TW - 1) / C2R_TW), (unsigned)((H + C2R_TH - 1) / C2R_TH));
    conv2d_reg_kernel<T, KM><<<grid, 256, smem, s>>>(out, x, h, kh, kw, off0, off1, sgn, sub, H, W);
    return check_launch(what);
}

template <typename T>
static int conv2d_direct_launch(T* out, const T* x, const T* h, int kh, int kw, int off0, int off1, int sgn, int64_t H,
                                int64_t W, cudaStream_t s, const char* what) {
    PCS_REQUIRE(H >= 0 && W >= 0 && kh >= 1 && kw >= 1, "%s: bad dimensions", what);
    PCS_REQUIRE(off0 >= 0 && off0 < kh && off1 >= 0 && off1 < kw, "%s: offset outside the filter", what);
    if (H * W == 0) return PCS_OK;
    PCS_REQUIRE(out && x && h && out != x, "%s: null or aliased pointer", what);
    static const bool no_reg = [] { const char* e = getenv("PCS_CONV2D_NO_REGTILE"); return e && atoi(e) != 0; }();
    if (!no_reg && ((uintptr_t)out % (2 * sizeof(T))) == 0) {
        if (kh <= 7 && kw <= 7) return conv2d_reg_launch<T, 7>(out, x, h, kh, kw, off0, off1, sgn, nullptr, H, W, s, what);
        if (kh <= 15 && kw <= 15) return conv2d_reg_launch<T, 15>(out, x, h, kh, kw, off0, off1, sgn, nullptr, H, W, s, what);
    }
    const size_t smem = ((size_t)kh * kw + (size_t)(C2_TH + kh - 1) * (C2_TW + kw - 1)) * sizeof(T);
    PCS_REQUIRE(smem <= C2_SMEM_LIMIT, "%s: filter %dx%d too large for the direct kernel", what, kh, kw);
    cudaFuncSetAttribute(conv2d_direct_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C2_SMEM_LIMIT);
    dim3 grid((unsigned)((W + C2_TW - 1) / C2_TW), (unsigned)((H + C2_TH - 1) / C2_TH));
    conv2d_direct_kernel<T><<<grid, dim3(C2_BX, C2_BY), smem, s>>>(out, x, h, kh, kw, off0, off1, sgn, H, W);
    return check_launch(what);
}
#define C2D_PARAMS                                                                                             \
    (void* out, const void* x, const void* h, int kh, int kw, int offset0, int offset1, int64_t H, int64_t W,  \
     void* stream)
template <typename T>
static int pcs_conv2d_direct_fwd_impl(T* o, const T* x, const T* h, int kh, int kw, int o0, int o1, int64_t H,
                                      int64_t W, cudaStream_t s) {
    return conv2d_direct_launch<T>(o, x, h, kh, kw, o0, o1, -1, H, W, s, "pcs_conv2d_direct_fwd");
}
PCS_EXPORT_REAL(pcs_conv2d_direct_fwd, C2D_PARAMS,
                ((T*)out, (const T*)x, (const T*)h, kh, kw, offset0, offset1, H, W, (cudaStream_t)stream))
template <typename T>
static int pcs_conv2d_direct_adj_impl(T* o, const T* x, const T* h, int kh, int kw, int o0, int o1, int64_t H,
                                      int64_t W, cudaStream_t s) {
    return conv2d_direct_launch<T>(o, x, h, kh, kw, o0, o1, +1, H, W, s, "pcs_conv2d_direct_adj");
}
PCS_EXPORT_REAL(pcs_conv2d_direct_adj, C2D_PARAMS,
                ((T*)out, (const T*)x, (const T*)h, kh, kw, offset0, offset1, H, W, (cudaStream_t)stream))

// rank-1 filter h[a,b] = hcol[a] * hrow[b]; out = conv/corr(x) (- sub if sub != nullptr)
template <typename T>
__global__ void __launch_bounds__(C2_BX * C2_BY) conv2d_sep_kernel(T* __restrict__ out, const T* __restrict__ x,
                                                                   const T* __restrict__ hcol,
                                                                   const T* __restrict__ hrow, int kh, int kw,
                                                                   int off0, int off1, int sgn,
                                                                   const T* __restrict__ sub, int64_t H, int64_t W) {
    extern __shared__ unsigned char smem_raw[];
    T* tcol = reinterpret_cast<T*>(smem_raw);
    T* trow = tcol + kh;
    const int sh = C2_TH + kh - 1, sw = C2_TW + kw - 1;
    T* tile = trow + kw;              // sh x sw
    T* mid = tile + sh * sw;          // sh x C2_TW (row pass result)
    const int tid = threadIdx.y * C2_BX + threadIdx.x;
    for (int e = tid; e < kh; e += C2_BX * C2_BY) tcol[e] = (sgn > 0) ? hcol[e] : hcol[kh - 1 - e];
    for (int e = tid; e < kw; e += C2_BX * C2_BY) trow[e] = (sgn > 0) ? hrow[e] : hrow[kw - 1 - e];
    const int ca = (sgn > 0) ? off0 : kh - 1 - off0, cb = (sgn > 0) ? off1 : kw - 1 - off1;
    const int64_t R0 = (int64_t)blockIdx.y * C2_TH, C0 = (int64_t)blockIdx.x * C2_TW;
    load_tile<T>(tile, sh, sw, x, H, W, R0 - ca, C0 - cb);
    __syncthreads();
    // row pass: mid[lr, lc] = sum_b trow[b] tile[lr, lc + b]
    for (int e = tid; e < sh * C2_TW; e += C2_BX * C2_BY) {
        const int lr = e / C2_TW, lc = e - lr * C2_TW;
        T acc = T(0);
        const T* p = tile + lr * sw + lc;
        for (int b = 0; b < kw; ++b) acc += trow[b] * p[b];
        mid[e] = acc;
    }
    __syncthreads();
    // column pass with a register window: thread owns C2_RPT consecutive rows of one column
    const int lc = threadIdx.x, lr0 = threadIdx.y * C2_RPT;
    T acc[C2_RPT];
#pragma unroll
    for (int j = 0; j < C2_RPT; ++j) acc[j] = T(0);
    for (int a = 0; a < kh + C2_RPT - 1; ++a) {
        const T v = mid[(lr0 + a) * C2_TW + lc];
#pragma unroll
        for (int j = 0; j < C2_RPT; ++j) {
            const int t = a - j;
            if (t >= 0 && t < kh) acc[j] += tcol[t] * v;
        }
    }
    const int64_t c = C0 + lc;
    if (c < W) {
#pragma unroll
        for (int j = 0; j < C2_RPT; ++j) {
            const int64_t r = R0 + lr0 + j;
            if (r < H) out[r * W + c] = sub ? acc[j] - sub[r * W + c] : acc[j];
        }
    }
}

// ---- streaming form of the separable kernel (odd square-ish filters up to 15 taps per axis) -------------------------
// The tile kernel above re-reads a (32 + kh - 1) x (64 + kw - 1) halo tile per 32 x 64 outputs, stages it with one integer
// division per element and spends one shared-memory load per multiply-add in its row pass: ~1 TB/s at 4096^2.  Here a
// block owns a strip of 128 output columns and MARCHES down a segment of rows, 8 input rows per step:
//   global rows -> registers (the loads of step t + 1 are in flight while step t is filtered) -> shared stage ->
//   row pass (a thread filters 4 adjacent outputs of one row from an 18-sample register window, 128-bit conflict-free
//   shared loads) -> ring of row-filtered lines -> column pass (thread <-> column, 4 rows) -> global.
// Every input row is read once per strip (+ kh - 1 rows per segment), nothing is recomputed vertically.  KH / KW are the
// exact tap counts (compile time), so an output accumulates exactly its taps, in ascending order, with one fma each --
// the same operations in the same order as the tile kernel.
constexpr int CS_TC = 128;              // output columns per strip
constexpr int CS_BR = 8;                // input rows per step
constexpr int CS_RUN = 4;               // row-pass outputs per thread
constexpr int CS_RG = 4;                // column-pass outputs per thread
constexpr int CS_NT = 256;
template <typename T> constexpr int cs_pitch(int n) {                  // smallest pitch >= n with pitch bytes = 16 mod 128
    const int per16 = 16 / (int)sizeof(T), per128 = 128 / (int)sizeof(T);
    int p = (n + per16 - 1) / per16 * per16;
    while (p % per128 != per16) p += per16;
    return p;
}
template <typename T, int K> struct CsGeo {
    static constexpr int SW = CS_TC + K - 1;                      // staged columns
    static constexpr int PX = cs_pitch<T>(SW + 3);                // (+3: the row-pass window is read in whole vectors)
    static constexpr int PM = cs_pitch<T>(CS_TC);
    static constexpr int NB = (CS_BR + K - 1 + CS_BR - 1) / CS_BR;  // ring boxes of CS_BR rows
    static constexpr int RING = NB * CS_BR;
    static constexpr int LAG = NB - 1;                            // output box s is complete after load step s + LAG
    static constexpr int NLD = (CS_BR * SW + CS_NT - 1) / CS_NT;  // staged elements per thread and step
    static constexpr int KT = (K + 3) / 4 * 4;                    // taps per axis in shared memory (padded)
    static constexpr size_t SMEM = ((size_t)2 * CS_BR * PX + (size_t)RING * PM + 2 * KT) * sizeof(T);
};

template <typename T, int K>
__global__ void __launch_bounds__(CS_NT) conv2d_sep_stream_kernel(T* __restrict__ out, const T* __restrict__ x,
                                                                   const T* __restrict__ hcol,
                                                                   const T* __restrict__ hrow, int sgn, int ca, int cb,
                                                                   const T* __restrict__ sub, int H, int W, int seg_rows) {
    using G = CsGeo<T, K>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* const stage = reinterpret_cast<T*>(smem_raw);               // 2 x CS_BR x PX
    T* const ring = stage + 2 * CS_BR * G::PX;                     // RING x PM
    T* const s_col = ring + G::RING * G::PM;                       // taps in correlation order (the device arrays hold them
    T* const s_row = s_col + G::KT;                                // in convolution order): read back into registers per pass
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < K) {
        s_col[tid] = (sgn > 0) ? hcol[tid] : hcol[K - 1 - tid];
        s_row[tid] = (sgn > 0) ? hrow[tid] : hrow[K - 1 - tid];
    }
    const int c0 = (int)blockIdx.x * CS_TC;
    const int ra = (int)blockIdx.y * seg_rows, rb = min(H, ra + seg_rows);
    if (ra >= rb) return;
    const int nout = (rb - ra + CS_BR - 1) / CS_BR;               // output boxes
    const int nloop = nout + G::LAG;

    // staging: element e of a step = (row e / SW, column e % SW) of the box whose corner is (ra - ca + 8 t, c0 - cb)
    T regs[G::NLD];
    auto gload = [&](int t) {
        const int r0 = ra - ca + CS_BR * t;
#pragma unroll
        for (int k = 0; k < G::NLD; ++k) {
            const int e = tid + k * CS_NT;
            const int lr = e / G::SW, lc = e - lr * G::SW;
            const int r = r0 + lr, c = c0 - cb + lc;
            regs[k] = (e < CS_BR * G::SW && r >= 0 && r < H && c >= 0 && c < W) ? x[(int64_t)r * W + c] : T(0);
        }
    };
    auto sstore = [&](int t) {
        T* st = stage + (t & 1) * CS_BR * G::PX;
#pragma unroll
        for (int k = 0; k < G::NLD; ++k) {
            const int e = tid + k * CS_NT;
            const int lr = e / G::SW, lc = e - lr * G::SW;
            if (e < CS_BR * G::SW) st[lr * G::PX + lc] = regs[k];
        }
    };
    // row pass: lanes 0-7 of a quarter warp are the 8 rows of the step (same columns): conflict-free 128-bit loads
    const int prow = lane & (CS_BR - 1), jc = (warp * (32 / CS_BR) + lane / CS_BR) * CS_RUN;
    constexpr int VN = Vec<T>::N;
    constexpr int WIN = (CS_RUN + K - 1 + VN - 1) / VN * VN;
    typedef typename Vec<T>::type VT;
    // column pass: thread <-> (column j, rows [4 rg, 4 rg + 4) of the box)
    const int j = tid & (CS_TC - 1), rg = tid >> 7;
    const bool vcol = c0 + j < W;

    gload(0);
    for (int t = 0; t < nloop; ++t) {
        sstore(t);
        __syncthreads();                       // stage t complete; ring box t % NB free (its column pass ended last step)
        if (t + 1 < nloop) gload(t + 1);
        {
            const T* xs = stage + (t & 1) * CS_BR * G::PX + prow * G::PX + jc;
            T acc[CS_RUN], tr[K];
#pragma unroll
            for (int b = 0; b < K; ++b) tr[b] = s_row[b];
#pragma unroll
            for (int o = 0; o < CS_RUN; ++o) acc[o] = T(0);
#pragma unroll
            for (int k = 0; k < WIN; k += VN) {
                T v[VN];
                *reinterpret_cast<VT*>(v) = *reinterpret_cast<const VT*>(xs + k);
#pragma unroll
                for (int e = 0; e < VN; ++e) {
#pragma unroll
                    for (int o = 0; o < CS_RUN; ++o) {
                        const int b = k + e - o;
                        if (b >= 0 && b < K) acc[o] = fma(tr[b], v[e], acc[o]);
                    }
                }
            }
            T* dst = ring + ((t % G::NB) * CS_BR + prow) * G::PM + jc;
#pragma unroll
            for (int o = 0; o < CS_RUN; o += VN) *reinterpret_cast<VT*>(dst + o) = *reinterpret_cast<const VT*>(acc + o);
        }
        __syncthreads();                       // ring box t % NB complete
        const int sbox = t - G::LAG;
        if (sbox >= 0 && vcol) {
            // output rows ra + 8 sbox + 4 rg + o need ring rows (relative to the segment's first staged row)
            // 8 sbox + 4 rg + o + a, a < K
            int rr = (sbox * CS_BR + CS_RG * rg) % G::RING;
            T acc[CS_RG], tc[K];
#pragma unroll
            for (int a = 0; a < K; ++a) tc[a] = s_col[a];
#pragma unroll
            for (int o = 0; o < CS_RG; ++o) acc[o] = T(0);
#pragma unroll
            for (int k = 0; k < CS_RG + K - 1; ++k) {
                const T m = ring[rr * G::PM + j];
                rr = (rr + 1 == G::RING) ? 0 : rr + 1;
#pragma unroll
                for (int o = 0; o < CS_RG; ++o) {
                    const int a = k - o;
                    if (a >= 0 && a < K) acc[o] = fma(tc[a], m, acc[o]);
                }
            }
            const int r0 = ra + sbox * CS_BR + CS_RG * rg;
#pragma unroll
            for (int o = 0; o < CS_RG; ++o) {
                const int r = r0 + o;
                if (r < rb) {
                    const int64_t idx = (int64_t)r * W + c0 + j;
                    out[idx] = sub ? acc[o] - sub[idx] : acc[o];
                }
            }
        }
    }
}

template <typename T, int K>
static int conv2d_sep_stream_launch(T* out, const T* x, const T* hcol, const T* hrow, int off0, int off1, int sgn,
                                    const T* sub, int64_t H, int64_t W, cudaStream_t s, const char* what) {
    using G = CsGeo<T, K>;
    const int ca = (sgn > 0) ? off0 : K - 1 - off0, cb = (sgn > 0) ? off1 : K - 1 - off1;
    const int nstrips = (int)((W + CS_TC - 1) / CS_TC);
    // exactly ONE wave of resident blocks (a few blocks left over for a second wave nearly double the time); segments of at
    // least 64 rows (the ring fill costs K - 1 extra rows per segment)
    cudaFuncSetAttribute(conv2d_sep_stream_kernel<T, K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G::SMEM);
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, conv2d_sep_stream_kernel<T, K>, CS_NT, G::SMEM) != cudaSuccess ||
        per_sm < 1) {
        (void)cudaGetLastError();
        per_sm = 2;
    }
    int nseg = (per_sm * num_sms()) / nstrips;
    if (nseg < 1) nseg = 1;
    int seg_rows = (int)((H + nseg - 1) / nseg);
    seg_rows = (seg_rows + CS_BR - 1) / CS_BR * CS_BR;
    if (seg_rows < 64) seg_rows = 64;
    nseg = (int)((H + seg_rows - 1) / seg_rows);
    conv2d_sep_stream_kernel<T, K><<<dim3((unsigned)nstrips, (unsigned)nseg), CS_NT, G::SMEM, s>>>(
        out, x, hcol, hrow, sgn, ca, cb, sub, (int)H, (int)W, seg_rows);
    return check_launch(what);
}

template <typename T>
static int conv2d_sep_launch(T* out, const T* x, const T* hcol, const T* hrow, int kh, int kw, int off0, int off1,
                             int sgn, const T* sub, int64_t H, int64_t W, cudaStream_t s, const char* what) {
    PCS_REQUIRE(H >= 0 && W >= 0 && kh >= 1 && kw >= 1, "%s: bad dimensions", what);
    PCS_REQUIRE(off0 >= 0 && off0 < kh && off1 >= 0 && off1 < kw, "%s: offset outside the filter", what);
    if (H * W == 0) return PCS_OK;
    PCS_REQUIRE(out && x && hcol && hrow && out != x, "%s: null or aliased pointer", what);
    static const bool stream_off = [] { const char* e = getenv("PCS_CONV2D_SEP_TILE"); return e && atoi(e) != 0; }();
    if (kh == kw && (kh & 1) && kh >= 3 && kh <= 15 && H < (1ll << 31) - 64 && W < (1ll << 31) - 256 && !stream_off) {
        switch (kh) {      // exact tap counts at compile time (PCS_CONV2D_SEP_TILE=1 keeps the tile kernel: A/B runs)
            case 3: return conv2d_sep_stream_launch<T, 3>(out, x, hcol, hrow, off0, off1, sgn, sub, H, W, s, what);
            case 5: return conv2d_sep_stream_launch<T, 5>(out, x, hcol, hrow, off0, off1, sgn, sub, H, W, s, what);
            case 7: return conv2d_sep_stream_launch<T, 7>(out, x, hcol, hrow, off0, off1, sgn, sub, H, W, s, what);
            case 9: return conv2d_sep_stream_launch<T, 9>(out, x, hcol, hrow, off0, off1, sgn, sub, H, W, s, what);
            case 11: return conv2d_sep_stream_launch<T, 11>(out, x, hcol, hrow, off0, off1, sgn, sub, H, W, s, what);
            case 13: return conv2d_sep_stream_launch<T, 13>(out, x, hcol, hrow, off0, off1, sgn, sub, H, W, s, what);
            default: return conv2d_sep_stream_launch<T, 15>(out, x, hcol, hrow, off0, off1, sgn, sub, H, W, s, what);
        }
    }
    const size_t sh = C2_TH + kh - 1, sw = C2_TW + kw - 1;
    const size_t smem = ((size_t)kh + kw + sh * sw + sh * C2_TW) * sizeof(T);
    PCS_REQUIRE(smem <= C2_SMEM_LIMIT, "%s: filter %dx%d too large for the separable kernel", what, kh, kw);
    cudaFuncSetAttribute(conv2d_sep_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C2_SMEM_LIMIT);
    dim3 grid((unsigned)((W + C2_TW - 1) / C2_TW), (unsigned)((H + C2_TH - 1) / C2_TH));
    conv2d_sep_kernel<T><<<grid, dim3(C2_BX, C2_BY), smem, s>>>(out, x, hcol, hrow, kh, kw, off0, off1, sgn, sub, H, W);
    return check_launch(what);
}
#define C2S_PARAMS                                                                                              \
    (void* out, const void* x, const void* hcol, const void* hrow, int kh, int kw, int offset0, int offset1,    \
     const void* sub, int64_t H, int64_t W, void* stream)
template <typename T>
static int pcs_conv2d_sep_fwd_impl(T* o, const T* x, const T* hc, const T* hr, int kh, int kw, int o0, int o1,
                                   const T* sub, int64_t H, int64_t W, cudaStream_t s) {
    return conv2d_sep_launch<T>(o, x, hc, hr, kh, kw, o0, o1, -1, sub, H, W, s, "pcs_conv2d_sep_fwd");
}
PCS_EXPORT_REAL(pcs_conv2d_sep_fwd, C2S_PARAMS,
                ((T*)out, (const T*)x, (const T*)hcol, (const T*)hrow, kh, kw, offset0, offset1, (const T*)sub, H, W,
                 (cudaStream_t)stream))
template <typename T>
static int pcs_conv2d_sep_adj_impl(T* o, const T* x, const T* hc, const T* hr, int kh, int kw, int o0, int o1,
                                   const T* sub, int64_t H, int64_t W, cudaStream_t s) {
    return conv2d_sep_launch<T>(o, x, hc, hr, kh, kw, o0, o1, +1, sub, H, W, s, "pcs_conv2d_sep_adj");
}
PCS_EXPORT_REAL(pcs_conv2d_sep_adj, C2S_PARAMS,
                ((T*)out, (const T*)x, (const T*)hcol, (const T*)hrow, kh, kw, offset0, offset1, (const T*)sub, H, W,
                 (cudaStream_t)stream))

// ---- FFT path for large non-separable PSFs ---------------------------------------------------------------------------
// Convolve2D's default method in the reference is 'fft' (pycsou/linop/conv.py:168 -> scipy.signal.fftconvolve /
// correlate(method='fft')).  Here the transforms are library calls made by the host (cuFFT through torch.fft, where the
// reference calls pocketfft); the work around them is fused into two kernels of this library:
//   pcs_spectral_mul: X <- X * Hf or X * conj(Hf) in place on the half spectrum (forward / adjoint), and
//   pcs_crop_shift:   out[i][j] = scale * full[(i + s0) mod P0][(j + s1) mod P1] - sub[i][j], i.e. the "same" crop of the
//                     padded inverse transform, its 1/(P0 P1) normalisation and the data-term residual A x - y in one pass.
namespace {
template <typename R>
__global__ void __launch_bounds__(256) spectral_mul_kernel(R* __restrict__ a, const R* __restrict__ b, int conj_b, int64_t n) {
    typedef typename std::conditional<sizeof(R) == 4, float2, double2>::type V;
    V* av = reinterpret_cast<V*>(a);
    const V* bv = reinterpret_cast<const V*>(b);
    const R sg = conj_b ? R(-1) : R(1);
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) {
        const V x = av[k], h = bv[k];
        const R hi = sg * h.y;
        V r;
        r.x = x.x * h.x - x.y * hi;
        r.y = x.x * hi + x.y * h.x;
        av[k] = r;
    }
}
template <typename T>
__global__ void __launch_bounds__(256) crop_shift_kernel(T* __restrict__ out, const T* __restrict__ full,
                                                         const T* __restrict__ sub, int H, int W, int P0, int P1, int s0,
                                                         int s1, T scale) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= W) return;
    int fj = j + s1;
    fj = fj < 0 ? fj + P1 : (fj >= P1 ? fj - P1 : fj);
    for (int i = blockIdx.y; i < H; i += gridDim.y) {
        int fi = i + s0;
        fi = fi < 0 ? fi + P0 : (fi >= P0 ? fi - P0 : fi);
        T v = scale * full[(int64_t)fi * P1 + fj];
        if (sub != nullptr) v -= sub[(int64_t)i * W + j];
        out[(int64_t)i * W + j] = v;
    }
}
}  // namespace

template <typename T>
static int pcs_spectral_mul_impl(T* a, const T* b, int conj_b, int64_t n, cudaStream_t s) {
    const char* what = "pcs_spectral_mul";
    PCS_REQUIRE(n >= 0, "%s: bad size", what);
    if (n == 0) return PCS_OK;
    PCS_REQUIRE(a && b, "%s: null pointer", what);
    PCS_REQUIRE(((uintptr_t)a % (2 * sizeof(T))) == 0 && ((uintptr_t)b % (2 * sizeof(T))) == 0, "%s: unaligned complex array", what);
    const int64_t nb = (n + 255) / 256, cap = 8 * (int64_t)num_sms();
    spectral_mul_kernel<T><<<(unsigned)(nb < cap ? nb : cap), 256, 0, s>>>(a, b, conj_b, n);
    return check_launch(what);
}
PCS_EXPORT_REAL(pcs_spectral_mul, (void* a_inout, const void* b, int conj_b, int64_t n_complex, void* stream),
                ((T*)a_inout, (const T*)b, conj_b, n_complex, (cudaStream_t)stream))

template <typename T>
static int pcs_crop_shift_impl(T* out, const T* full, const T* sub, int64_t H, int64_t W, int64_t P0, int64_t P1,
                               int64_t s0, int64_t s1, double scale, cudaStream_t s) {
    const char* what = "pcs_crop_shift";
    PCS_REQUIRE(H >= 0 && W >= 0 && P0 >= H && P1 >= W && P0 < (1ll << 30) && P1 < (1ll << 30), "%s: bad dimensions", what);
    PCS_REQUIRE(s0 > -P0 && s0 < P0 && s1 > -P1 && s1 < P1 && H + (s0 > 0 ? s0 : 0) <= 2 * P0 && W + (s1 > 0 ? s1 : 0) <= 2 * P1,
                "%s: shift outside the padded array", what);
    if (H * W == 0) return PCS_OK;
    PCS_REQUIRE(out && full && out != full, "%s: null or aliased pointer", what);
    dim3 grid((unsigned)((W + 255) / 256), (unsigned)(H < 4096 ? H : 4096));
    crop_shift_kernel<T><<<grid, 256, 0, s>>>(out, full, sub, (int)H, (int)W, (int)P0, (int)P1, (int)s0, (int)s1, (T)scale);
    return check_launch(what);
}
PCS_EXPORT_REAL(pcs_crop_shift,
                (void* out, const void* full, const void* sub, int64_t H, int64_t W, int64_t P0, int64_t P1, int64_t shift0,
                 int64_t shift1, double scale, void* stream),
                ((T*)out, (const T*)full, (const T*)sub, H, W, P0, P1, shift0, shift1, scale, (cudaStream_t)stream))
