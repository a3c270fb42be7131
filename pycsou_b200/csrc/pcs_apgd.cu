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
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

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

// The reference's stop test, `metric > threshold` with metric = sqrt(s_d) / sqrt(s_x) (inf when s_x = 0; opt/proxalgs.py:607-622),
// decided without the two square roots and the division whenever the answer cannot depend on their rounding: the computed
// metric is within 4 ulp of sqrt(s_d / s_x), so s_d > thr^2 (1 + 1e-12) s_x means `above` and s_d < thr^2 (1 - 1e-12) s_x
// means `not above`; anything closer, non-finite or near the ends of the exponent range takes the reference's formula.
struct StopTest {
    double thr, lo, hi;
    bool fast;
    __device__ explicit StopTest(double threshold) : thr(threshold) {
        const double t2 = threshold * threshold;
        fast = t2 > 1e-290 && t2 < 1e290;
        lo = t2 * (1.0 - 1e-12);
        hi = t2 * (1.0 + 1e-12);
    }
    __device__ __forceinline__ bool above(double s_d, double s_x) const {
        if (fast) {
            const double ph = hi * s_x, pl = lo * s_x;
            if (pl > 1e-290 && ph < 1e290) {
                if (s_d > ph) return true;
                if (s_d < pl) return false;
            }
        }
        const double metric = s_x == 0.0 ? INFINITY : sqrt(s_d) / sqrt(s_x);
        return metric > thr;
    }
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
    const StopTest stop(A.threshold);
    double t_old = A.t0;
    bool above = INFINITY > A.threshold;                         // metric starts at inf
    long long it = A.it0;
    while ((it <= A.max_iter && above) || it <= A.min_iter) {      // core/solver.py:65-66
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
        above = stop.above(s_d, s_x);                            // opt/proxalgs.py:607-622 (np.linalg.norm ratio)
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

// ---- the same solve on a CLUSTER of thread blocks --------------------------------------------------------------------------
// One SM bounds the single-block kernel: at N = 4096 with a 65-tap filter an update is 532 k FP64 multiply-adds = 4.4 us of one
// SM's FP64 pipe (11.4 us measured).  Here a thread-block cluster (2 to 16 blocks, as many as the vector needs) shares the
// vector: block c owns M = 128 EPT consecutive elements and keeps them (and the residual) in ITS shared memory with a halo
// on either side.
// * Neighbours.  The thread that computes an element within the filter's reach of the edge of the segment also stores it
//   into the neighbour's halo with st.async -- a store through distributed shared memory that counts its bytes on an
//   mbarrier in the RECEIVING block -- so a block waits only for its own two neighbours and only for the bytes it is about
//   to read.  (A cluster barrier with release semantics is a MEMBAR.ALL.GPU per arrival: 0.9-1.5 k cycles per barrier, two
//   per update, were half the time of the first version of this kernel.)  No signal is needed in the other direction: a
//   block can only be sent new halo values by a neighbour that has received the values computed FROM the old ones, and
//   the threads that read a halo are the ones that send the edge values that neighbour is waiting for.
// * Passes.  Four warps run the two filter passes, 2 or 4 consecutive elements per thread.  What bounds a pass is shared-
//   memory wavefronts, not the FP64 pipe (16 blocks leave each SM 1/16 of the multiply-adds but every thread still reads
//   a window of samples per group of taps): the window is read as aligned pairs (apc_tap_pass), and each of the C tap
//   positions of a group has its own running sum, added at the end in a fixed tree.  So the taps of an element are not
//   added in the single-block kernel's order: iterates agree with it to rounding (1e-15 relative in fp64), not bit for bit.
// * Everything scalar is the FIFTH warp's: the momentum scalars of the next update (two FP64 divisions, a square root),
//   the block's sums of the last update (from the pass threads' partial sums in shared memory), their exchange -- the same
//   kind of store into slot `rank` of EVERY block; four sets of slots in turn, because a block may be up to two updates
//   ahead of another -- the sum of all blocks' slots in one order everywhere, the row of diagnostics and the stop test.
//   The pass warps do not wait for any of it: they start update `it` as soon as the neighbours' halos are in, and look at
//   the stop test of update it - 1 when update `it` is done; if it says stop the update is dropped (its input state is
//   still in registers).  The stop test is live and exact, but off the path of the updates; no global memory, no host.
__device__ __forceinline__ uint32_t apc_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t apc_mapa(uint32_t addr, int rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void apc_st_async(uint32_t raddr, double v, uint32_t rbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(raddr),
                 "l"(__double_as_longlong(v)), "r"(rbar) : "memory");
}
__device__ __forceinline__ void apc_st_async(uint32_t raddr, float v, uint32_t rbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(raddr),
                 "r"(__float_as_uint(v)), "r"(rbar) : "memory");
}
__device__ __forceinline__ void apc_mbar_init(uint32_t bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void apc_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// (try_wait suspends the thread for a while by itself.  Sixteen million failed tries in one wait can only be a bug: end the
// launch with an error rather than hang the GPU.)
__device__ __forceinline__ void apc_mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .u32 n;\n\t"
        "mov.u32 n, 0;\n\t"
        "APC_WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra APC_DONE_%=;\n\t"
        "add.u32 n, n, 1;\n\t"
        "setp.lt.u32 p, n, 0x1000000;\n\t"
        "@p bra APC_WAIT_%=;\n\t"
        "trap;\n\t"
        "APC_DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
constexpr int APC_MAXCL = 16;
template <typename T> struct ApcPair;
template <> struct ApcPair<double> { using type = double2; };
template <> struct ApcPair<float> { using type = float2; };
// One filter pass over a thread's EPT consecutive elements (EPT even, first element at an even index): for every group of C
// taps the EPT + C - 1 samples under them are read as aligned PAIRS (LDS.128 for doubles: 4 wavefronts for 64 values, the
// shared-memory bandwidth, where a 64-bit load at a stride of two doubles also costs 4) -- shared-memory wavefronts, not
// the FP64 pipe, bound this kernel.  `win` points at the first sample of the window of the first group, minus SH (0 or 1)
// so that it is pair-aligned; the window of group K starts K samples before it (DIR = -1, convolution) or after it
// (DIR = +1, correlation); tap c of the group multiplies sample e + C - 1 - c (DIR = -1) or e + c (DIR = +1) of element e.
template <typename T, int EPT, int SH, int DIR>
__device__ __forceinline__ void apc_tap_pass(const T* win, const T* sh, int nhp, T (&acc)[EPT][APGD_C]) {
    constexpr int C = APGD_C, NP = (EPT + C - 1 + SH + 1) / 2;
    using P = typename ApcPair<T>::type;
#pragma unroll 3
    for (int K = 0; K < nhp; K += C) {
        T hk[C], w[2 * NP];
#pragma unroll
        for (int c = 0; c < C; ++c) hk[c] = sh[K + c];
        const P* q = reinterpret_cast<const P*>(win + DIR * K);
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const P v = q[i];
            w[2 * i] = v.x;
            w[2 * i + 1] = v.y;
        }
#pragma unroll
        for (int c = 0; c < C; ++c)
#pragma unroll
            for (int e = 0; e < EPT; ++e) acc[e][c] = fma(hk[c], w[SH + (DIR < 0 ? e - c + C - 1 : e + c)], acc[e][c]);
    }
}
// The C running sums of an element (one per tap position within a group of C taps) are added in this fixed order.
template <typename T> __device__ __forceinline__ T tap_tree(const T (&a)[APGD_C]) {
    static_assert(APGD_C == 8, "tap_tree is written for groups of 8 taps");
    return ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
}
constexpr int APC_BLOCK = 160;     // four warps of filter passes + one warp for everything scalar
template <typename T, int EPT, int APC_NT>
__global__ void __launch_bounds__(APC_BLOCK, 1) apgd_cluster_kernel(const ApgdArgs<T> A) {
    static_assert(APC_NT == 128 && APC_BLOCK == APC_NT + 32 && EPT % 2 == 0, "four pass warps, one scalar warp, pair-aligned windows");
    extern __shared__ __align__(16) unsigned char smem[];
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank(), CL = (int)cluster.num_blocks();
    constexpr int C = APGD_C, M = APC_NT * EPT;
    const int n = A.n, off = A.offset;
    const int nhp = (A.nh + C - 1) / C * C;
    constexpr int HP = APGD_PAD;
    constexpr int LEN = M + 2 * HP;
    T* const bx = reinterpret_cast<T*>(smem);                  // iterate: halo | own | halo
    T* const br = bx + LEN;                                    // residual, same layout
    T* const sh = br + LEN;                                    // taps, zero padded to nhp
    double* const parts = reinterpret_cast<double*>(smem + ((size_t)(2 * LEN + APGD_MAX_TAPS) * sizeof(T) + 15) / 16 * 16);   // [4 sets][16 blocks][2]
    double2* const psum = reinterpret_cast<double2*>(parts + 4 * APC_MAXCL * 2);      // [128]: the pass threads' sums of one update
    double* const scal = reinterpret_cast<double*>(psum + APC_NT);                   // [2][2]: momentum scalars a, 1 + a of two updates
    unsigned long long* const bars = reinterpret_cast<unsigned long long*>(scal + 4); // x halo, r halo, sums[4]
    int* const dec = reinterpret_cast<int*>(bars + 6);                               // 1: the sums just looked at say stop
    T* const sx = bx + HP;                                     // own element j of this block at sx[j]
    T* const sr = br + HP;
    const int tid = threadIdx.x, lane = tid & 31;
    const bool scalar_warp = tid >= APC_NT;
    const int j0 = (scalar_warp ? 0 : tid) * EPT, g0 = rank * M + j0;      // local / global index of a pass thread's first element
    const uint32_t bar_x = apc_smem_u32(bars), bar_r = bar_x + 8, bar_p = bar_x + 16;
    const bool has_l = rank > 0, has_r = rank + 1 < CL;

    for (int i = tid; i < 2 * LEN + APGD_MAX_TAPS; i += APC_BLOCK) reinterpret_cast<T*>(smem)[i] = T(0);
    for (int i = tid; i < 4 * APC_MAXCL * 2; i += APC_BLOCK) parts[i] = 0.0;
    if (tid == 0) {
        for (int b = 0; b < 6; ++b) apc_mbar_init(bar_x + 8 * b, 1);
        *dec = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    cluster.sync();                                            // nobody stores into memory that is still being cleared; barriers exist
    T xa[EPT], pa[EPT], ya[EPT];
#pragma unroll
    for (int e = 0; e < EPT; ++e) {
        const int i = g0 + e;
        const bool in = !scalar_warp && i < n;
        xa[e] = in ? A.x[i] : T(0);
        pa[e] = in ? A.aux[i] : T(0);
        ya[e] = in ? A.y[i] : T(0);
        if (!scalar_warp) sx[j0 + e] = xa[e];
    }
    for (int k = tid; k < A.nh; k += APC_BLOCK) sh[k] = A.h[k];
    // Where this block's edge values go: the same offsets in the neighbours' copies of the two arrays, and their barriers.
    const uint32_t lbx = has_l ? apc_mapa(apc_smem_u32(bx), rank - 1) : 0u, rbx = has_r ? apc_mapa(apc_smem_u32(bx), rank + 1) : 0u;
    const uint32_t lbr = has_l ? apc_mapa(apc_smem_u32(br), rank - 1) : 0u, rbr = has_r ? apc_mapa(apc_smem_u32(br), rank + 1) : 0u;
    const uint32_t lbar = has_l ? apc_mapa(bar_x, rank - 1) : 0u, rbar = has_r ? apc_mapa(bar_x, rank + 1) : 0u;   // (+8: their bar_r)
    // Only what a neighbour's filter really reaches is sent: its residual pass reads x up to `off` elements to its right and
    // nh - 1 - off to its left, its gradient pass reads r nh - 1 - off to its right and `off` to its left.  (The passes also
    // read up to nhp + C samples away under the zero taps that pad the filter: those only have to be finite, and what
    // lies there -- zeros, or a value of some earlier update -- is.)  Local element j goes to the right halo of the block
    // on the left when j < wl, to the left halo of the block on the right when j >= M - wr.
    const int reach_a = off, reach_b = A.nh - 1 - off;
    auto push = [&](uint32_t left, uint32_t right, uint32_t bar_off, int wl, int wr, int j, T v) {
        if (has_l && j < wl) apc_st_async(left + (uint32_t)((HP + M + j) * sizeof(T)), v, lbar + bar_off);
        if (has_r && j >= M - wr) apc_st_async(right + (uint32_t)((HP + j - M) * sizeof(T)), v, rbar + bar_off);
    };
    // bytes a block receives per exchange: x from the left block's last reach_b and the right block's first reach_a
    // elements; r from the left block's last reach_a and the right block's first reach_b
    const uint32_t xbytes = (uint32_t)(((has_l ? reach_b : 0) + (has_r ? reach_a : 0)) * sizeof(T));
    const uint32_t rbytes = (uint32_t)(((has_l ? reach_a : 0) + (has_r ? reach_b : 0)) * sizeof(T));
    const uint32_t pbytes = (uint32_t)(CL * 2 * sizeof(double));
    const uint32_t my_parts = lane < CL ? apc_mapa(apc_smem_u32(parts), lane) : 0u, my_pbar = lane < CL ? apc_mapa(bar_p, lane) : 0u;
    if (tid == 0) apc_mbar_expect_tx(bar_x, xbytes);
    if (!scalar_warp) {
#pragma unroll
        for (int e = 0; e < EPT; ++e) push(lbx, rbx, 0u, reach_a, reach_b, j0 + e, xa[e]);
    }
    uint32_t ph_x = 0, ph_r = 0, ph_p = 0;                     // phase parities (ph_p: bit b for set b of the sums; scalar warp)
    const T tau = A.tau;
    const StopTest stop(A.threshold);
    long long it = A.it0;
    // momentum (scalar warp): t_m = t of the last update applied, t_c = t of update `it`; a = (t_m - 1) / t_c
    double t_m = A.t0, t_c = A.t0;
    auto momentum = [&](long long upd, double t_before, double* slot) -> double {
        double t;
        if (A.accel == ACC_BT) t = (1.0 + sqrt(1.0 + 4.0 * t_before * t_before)) / 2.0;
        else if (A.accel == ACC_CD) t = ((double)upd + A.d) / A.d;
        else { t = 1.0; t_before = 1.0; }
        const double a = (t_before - 1.0) / t;
        if (lane == 0) { slot[0] = a; slot[1] = 1.0 + a; }
        return t;
    };
    if (scalar_warp) t_c = momentum(it, t_m, scal + 2 * (int)((it - A.it0) & 1));
    __syncthreads();                                           // own part of sx, the taps and the first momentum scalars are in place
    // core/solver.py:65-66: update `it` runs when (it <= max_iter and metric(it - 1) > threshold) or it <= min_iter, the metric
    // before the first update being inf.  The sums behind metric(it - 1) come from all blocks; waiting for them would put an
    // all-to-all exchange and a chain of FP64 additions on the path of every update.  Instead the pass warps go straight on
    // to update `it`, on the assumption that the metric was above the threshold, while the scalar warp adds the block's sums
    // of update it - 1, exchanges them, adds all blocks' and sets `dec`; after update `it` everybody looks at `dec`, and if
    // it says stop the update is dropped (the pass threads keep the state before it in registers).  What the caller gets
    // -- iterates, number of updates, diagnostics -- is what the reference's loop gives.
    // Adds the pass threads' sums of update `upd` and sends the block's two sums to slot `rank` of set upd & 3 of every block.
    auto send_sums = [&](long long upd) {
        const unsigned set = (unsigned)(upd - A.it0) & 3u;
        if (lane == 0) apc_mbar_expect_tx(bar_p + 8 * set, pbytes);
        double s_d = 0.0, s_x = 0.0;
#pragma unroll
        for (int c = 0; c < APC_NT / 32; ++c) {
            const double2 v = psum[lane + 32 * c];
            s_d += v.x;
            s_x += v.y;
        }
        s_d = warp_sum(s_d);
        s_x = warp_sum(s_x);
        if (lane < CL) {
            const uint32_t slot = (uint32_t)((set * APC_MAXCL + rank) * 2 * sizeof(double));
            apc_st_async(my_parts + slot, s_d, my_pbar + 8 * set);
            apc_st_async(my_parts + slot + 8, s_x, my_pbar + 8 * set);
        }
    };
    // Adds all blocks' sums of update `upd` once they are here (same slots, same order in every block), writes the row of
    // diagnostics, returns the stop test.
    auto settle = [&](long long upd) -> bool {
        const unsigned set = (unsigned)(upd - A.it0) & 3u;
        apc_mbar_wait(bar_p + 8 * set, (ph_p >> set) & 1u);
        ph_p ^= 1u << set;
        const double* pp = parts + (size_t)set * APC_MAXCL * 2;
        double s_d = lane < APC_MAXCL ? pp[2 * lane] : 0.0, s_x = lane < APC_MAXCL ? pp[2 * lane + 1] : 0.0;   // (absent blocks: zero)
        s_d = warp_sum(s_d);
        s_x = warp_sum(s_x);
        if (rank == 0 && lane == 0) {
            double* row = A.diag + 4 * (upd - A.it0);
            row[0] = s_d; row[1] = s_x; row[2] = 0.0; row[3] = 0.0;
        }
        return stop.above(s_d, s_x);
    };
#ifdef APC_PHASE_TIMING                                          // -DAPC_PHASE_TIMING: cycles per phase of a pass thread, printed by two blocks
    long long ph[6] = {0, 0, 0, 0, 0, 0}, tk = clock64();
#define APC_TICK(i) { const long long now_ = clock64(); ph[i] += now_ - tk; tk = now_; }
#else
#define APC_TICK(i) {}
#endif
    T xk[EPT], pk[EPT];                                          // pass threads: the state before the update in flight
    bool pending = false;                                        // the sums of update it - 1 have not been looked at yet
    bool go = (it <= A.max_iter && INFINITY > A.threshold) || it <= A.min_iter;
    while (go) {
        double t_n = t_c;
        if (scalar_warp) {
            if (pending) send_sums(it - 1);
            t_n = momentum(it + 1, t_c, scal + 2 * (int)((it + 1 - A.it0) & 1));      // for the next update
        } else {
            if (tid == 0) apc_mbar_expect_tx(bar_r, rbytes);
            apc_mbar_wait(bar_x, ph_x);                          // the neighbours' edge values of x are in the halos
            ph_x ^= 1u;
            APC_TICK(0)
            // residual r = h * x - y on the own segment; tap c of every group of C has its own running sum
            T acc[EPT][C];
#pragma unroll
            for (int e = 0; e < EPT; ++e)
#pragma unroll
                for (int c = 0; c < C; ++c) acc[e][c] = T(0);
            // first sample of the first window: x[j0 + off - (C - 1)]; j0 is even, so its parity is that of off + 1
            if (off & 1) apc_tap_pass<T, EPT, 0, -1>(sx + j0 + off - (C - 1), sh, nhp, acc);
            else apc_tap_pass<T, EPT, 1, -1>(sx + j0 + off - (C - 1) - 1, sh, nhp, acc);
#pragma unroll
            for (int e = 0; e < EPT; ++e) {
                const T s = tap_tree(acc[e]);
                const T rv = (g0 + e < n) ? s - ya[e] : T(0);
                sr[j0 + e] = rv;
                push(lbr, rbr, 8u, reach_b, reach_a, j0 + e, rv);
            }
            APC_TICK(1)
        }
        __syncthreads();                                         // A: own part of sr complete; every pass thread is past the wait on bar_x
        if (scalar_warp) {
            if (pending) {
                const bool above = settle(it - 1);
                if (lane == 0) *dec = above ? 0 : 1;
            }
        } else {
            APC_TICK(2)
            if (tid == 0) apc_mbar_expect_tx(bar_x, xbytes);     // next update's x halos
            const double* my_scal = scal + 2 * (int)((it - A.it0) & 1);
            const T ta = (T)my_scal[0], t1a = (T)my_scal[1];
            apc_mbar_wait(bar_r, ph_r);                          // the neighbours' edge values of r are in the halos
            ph_r ^= 1u;
            APC_TICK(3)
            double s_d = 0.0, s_x = 0.0;
            T g[EPT][C];
#pragma unroll
            for (int e = 0; e < EPT; ++e)
#pragma unroll
                for (int c = 0; c < C; ++c) g[e][c] = T(0);
            if (off & 1) apc_tap_pass<T, EPT, 1, +1>(sr + j0 - off - 1, sh, nhp, g);
            else apc_tap_pass<T, EPT, 0, +1>(sr + j0 - off, sh, nhp, g);
#pragma unroll
            for (int e = 0; e < EPT; ++e) {
                xk[e] = xa[e];
                pk[e] = pa[e];
                if (g0 + e < n) {
                    const T v = xa[e] - mul_rn(tau, tap_tree(g[e]));
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
#pragma unroll
            for (int e = 0; e < EPT; ++e) {
                sx[j0 + e] = xa[e];
                push(lbx, rbx, 0u, reach_a, reach_b, j0 + e, xa[e]);
            }
            psum[tid] = make_double2(s_d, s_x);
            APC_TICK(4)
        }
        __syncthreads();                                         // B: own part of sx, the sums and `dec` complete; every pass thread is past the wait on bar_r
        if (pending && *dec != 0 && it > A.min_iter) {           // the loop had ended before this update: drop it
#pragma unroll
            for (int e = 0; e < EPT; ++e) { xa[e] = xk[e]; pa[e] = pk[e]; }
            pending = false;
            break;
        }
        pending = true;
        t_m = t_c;
        t_c = t_n;
        ++it;
        go = it <= A.max_iter || it <= A.min_iter;               // (and the metric above the threshold: known one update later)
        if (!scalar_warp) APC_TICK(5)
    }
#ifdef APC_PHASE_TIMING
    if ((tid == 0 || tid == 64) && (rank == 1 || rank == CL - 1) && it > A.it0)
        printf("apgd cluster block %d thread %d: %lld updates; cycles per update: wait x %lld, residual pass %lld, barrier A %lld, wait r %lld, "
               "gradient pass + prox + stores %lld, barrier B + loop %lld\n", rank, tid, it - A.it0, ph[0] / (it - A.it0), ph[1] / (it - A.it0),
               ph[2] / (it - A.it0), ph[3] / (it - A.it0), ph[4] / (it - A.it0), ph[5] / (it - A.it0));
#endif
    if (scalar_warp) {
        if (pending) {                                           // the last update's row of diagnostics
            send_sums(it - 1);
            settle(it - 1);
        }
        if (rank == 0 && lane == 0) {
            A.n_iter[0] = it;
            A.n_iter[1] = __double_as_longlong(t_m);
        }
    } else {
#pragma unroll
        for (int e = 0; e < EPT; ++e) {
            if (g0 + e < n) {
                A.x[g0 + e] = xa[e];
                A.aux[g0 + e] = pa[e];
            }
        }
        apc_mbar_wait(bar_x, ph_x);                              // everything sent to this block's halos has arrived ...
    }
    cluster.sync();                                              // ... and that holds for every block: nothing is in flight when one leaves
}

// what the last pcs_apgd_conv1d_run call of this host thread launched (pcs_apgd_last_launch; tests assert on it)
thread_local int64_t g_apgd_last[4] = {-1, 0, 0, 0};   // kernel (0 single block, 1 cluster), blocks, threads per block, elements per thread

// Returns -1 when the cluster kernel cannot run this problem (the caller falls back to the single-block kernel).
template <typename T, int EPT, int APC_NT> int apgd_cluster_launch(const ApgdArgs<T>& A, int CL, cudaStream_t s) {
    constexpr int M = APC_NT * EPT;
    const size_t smem = ((size_t)(2 * (M + 2 * APGD_PAD) + APGD_MAX_TAPS) * sizeof(T) + 15) / 16 * 16 + (4 * 2 * APC_MAXCL + 2 * APC_NT + 4) * sizeof(double) + 6 * sizeof(unsigned long long) + 16;
    auto kern = apgd_cluster_kernel<T, EPT, APC_NT>;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)CL); cfg.blockDim = dim3(APC_BLOCK); cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    // Function attributes and the "can one such cluster be resident" query: once per device, kernel and cluster size
    // (0 not asked yet, 1 yes, -1 no).  A race between host threads only repeats the calls.
    static signed char ready[64][4];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return -1;
    signed char& rdy = ready[dev][CL == 2 ? 0 : (CL == 4 ? 1 : (CL == 8 ? 2 : 3))];
    if (rdy == 0) {
        int nclusters = 0;
        const bool ok = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) == cudaSuccess &&
                        cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess &&
                        cudaOccupancyMaxActiveClusters(&nclusters, kern, &cfg) == cudaSuccess && nclusters >= 1;
        if (!ok) (void)cudaGetLastError();
        rdy = ok ? 1 : -1;
    }
    if (rdy < 0) return -1;
    if (cudaLaunchKernelEx(&cfg, kern, A) != cudaSuccess) {
        (void)cudaGetLastError();
        return -1;
    }
    g_apgd_last[0] = 1; g_apgd_last[1] = CL; g_apgd_last[2] = APC_BLOCK; g_apgd_last[3] = EPT;
    return check_launch("pcs_apgd_conv1d_run");
}
template <typename T> int apgd_cluster_try(const ApgdArgs<T>& A, cudaStream_t s) {
    const char* env = getenv("PCS_APGD_CLUSTER");               // PCS_APGD_CLUSTER=0 keeps the single-block kernel (read per call)
    const char* envn = getenv("PCS_APGD_CLUSTER_MIN_N");         // (experiments: smallest n that takes the cluster kernel)
    // An update of the cluster kernel costs about the same 1.9-2.3 us (fp64) whatever n: it wins from n = 512 on (single
    // block: 2.7 us there, 4.0 at 1024, 6.6 at 2047, 11.4 at 4096; profiles/r02_apgd_cluster.txt).
    const int min_n = envn ? atoi(envn) : 512;
    if ((env && atoi(env) == 0) || A.n < min_n) return -1;
    const int nhp = (A.nh + APGD_C - 1) / APGD_C * APGD_C, exw = nhp + APGD_C;
    // the smallest cluster that holds the vector at 2 elements per thread, else at 4
    for (int ept : {2, 4}) {
        for (int CL : {2, 4, 8, 16}) {
            if ((int64_t)CL * 128 * ept < A.n || 128 * ept < exw) continue;
            const int rc = ept == 2 ? apgd_cluster_launch<T, 2, 128>(A, CL, s) : apgd_cluster_launch<T, 4, 128>(A, CL, s);
            if (rc != -1) return rc;
        }
    }
    return -1;
}

template <typename T, int EPT, int NT> int apgd_launch(const ApgdArgs<T>& A, size_t smem, cudaStream_t s) {
    cudaFuncSetAttribute(apgd_run_kernel<T, EPT, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    apgd_run_kernel<T, EPT, NT><<<1, NT, smem, s>>>(A);
    g_apgd_last[0] = 0; g_apgd_last[1] = 1; g_apgd_last[2] = NT; g_apgd_last[3] = EPT;
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
    {
        const int rc = apgd_cluster_try<T>(A, s);              // a thread-block cluster shares the vector where it can
        if (rc != -1) return rc;
    }
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
extern "C" int pcs_apgd_last_launch(int64_t* out4) {
    if (!out4) return PCS_ERR_ARG;
    for (int i = 0; i < 4; ++i) out4[i] = g_apgd_last[i];
    return PCS_OK;
}

PCS_EXPORT_REAL(pcs_apgd_conv1d_run,
                (void* x_inout, void* aux_inout, const void* y, const void* h, int nh, int offset, int64_t n, int g_mode,
                 double g_param, double tau, int accel, double d, double t0, int64_t it0, int64_t max_iter,
                 int64_t min_iter, double threshold, void* diag_out_f64, void* n_iter_out_i64, void* stream),
                ((T*)x_inout, (T*)aux_inout, (const T*)y, (const T*)h, nh, offset, n, g_mode, g_param, tau, accel, d, t0,
                 it0, max_iter, min_iter, threshold, (double*)diag_out_f64, (long long*)n_iter_out_i64,
                 (cudaStream_t)stream))
