/* pcsb200.h -- C ABI of libpcsb200.so: B200 (sm_100a) kernels for pycsou's iterative proximal-solver hot path.
 *
 * Conventions
 *  - Plain C types only.  All array arguments are DEVICE pointers owned by the caller (PyTorch allocates them);
 *    the library never allocates, frees or retains pointers beyond the call.  Workspaces are passed in.
 *  - Every entry point named <op>_f32 / <op>_f64 exists for float and double; scalars are passed as double.
 *  - Calls are asynchronous on the cudaStream_t passed as `void* stream`; no synchronisation inside unless stated.
 *  - Return 0 on success, 1 = bad argument, 2 = CUDA failure; pcs_last_error() (thread-local) describes it.
 *  - Vectors are flat C-order arrays; operators get the n-D shape.  An n-D array processed along one axis is described
 *    by the triple (outer, n, inner) = (prod(shape[:axis]), shape[axis], prod(shape[axis+1:])).
 *  - Stencil / convolution outputs must not alias their inputs (caller ping-pongs); elementwise ops may alias.
 *
 * Each block cites the reference interface (file:line under /root/reference) the entry points replace.
 */
#ifndef PCSB200_H
#define PCSB200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

int pcs_version(void);
const char* pcs_last_error(void);
int pcs_num_sms(void);
/* Planning override of the SM count the launch geometries are sized for (n <= 0: the device's own); returns the
 * previous override.  Results never depend on it -- the parity tests use it to run the one-wave / uneven-split
 * geometries of the fused kernels on small images.  Process-wide; not for concurrent use. */
int pcs_set_num_sms(int n);

#define PCS_DECL_REAL(name, params) int name##_f32 params; int name##_f64 params;

/* ---- element-wise arithmetic of update_iterand (pycsou/opt/proxalgs.py:348-353, :591-599) ---------------------- */
PCS_DECL_REAL(pcs_lincomb2, (void* out, double a, const void* x, double b, const void* y, int64_t n, void* stream))
PCS_DECL_REAL(pcs_lincomb3, (void* out, double a, const void* x, double b, const void* y, double c, const void* z,
                             int64_t n, void* stream))
PCS_DECL_REAL(pcs_scale, (void* out, double a, const void* x, int64_t n, void* stream))
/* DiagonalOperator.__call__/adjoint (pycsou/linop/base.py:569-579) */
PCS_DECL_REAL(pcs_mul, (void* out, const void* x, const void* y, int64_t n, void* stream))

/* ---- proximal maps ------------------------------------------------------------------------------------------- */
/* L1Norm.prox = LpNorm.prox with proj_linfty_ball (func/penalty.py:194-245, func/base.py:239-240, math/prox.py:253-256) */
PCS_DECL_REAL(pcs_prox_l1, (void* out, const void* x, double tau, int64_t n, void* stream))
/* proj_nonnegative_orthant (math/prox.py:295-297), NonNegativeOrthant.prox (func/penalty.py:563-612) */
PCS_DECL_REAL(pcs_proj_nonneg, (void* out, const void* x, int64_t n, void* stream))
/* proj_segment (math/prox.py:340-343), Segment.prox (func/penalty.py:615-668); also proj_linfty_ball */
PCS_DECL_REAL(pcs_proj_segment, (void* out, const void* x, double a, double b, int64_t n, void* stream))
/* LogBarrier.prox (x + sqrt(x^2 + 4 tau))/2 (func/penalty.py:824-840); ShannonEntropy.prox tau W(exp(x/tau - 1)/tau)
 * with the Lambert function evaluated in the log domain (func/penalty.py:905-922 -> scipy.special.lambertw);
 * KLDivergence.prox (x - tau + sqrt((x - tau)^2 + 4 tau data))/2 (func/loss.py:666-682) */
PCS_DECL_REAL(pcs_prox_logbarrier, (void* out, const void* x, double tau, int64_t n, void* stream))
PCS_DECL_REAL(pcs_prox_shannon, (void* out, const void* x, double tau, int64_t n, void* stream))
PCS_DECL_REAL(pcs_prox_kl, (void* out, const void* x, const void* data, double tau, int64_t n, void* stream))
/* L21Norm.prox (func/penalty.py:551-557) for groups = tile(arange(n_groups), n_comp); is_complex: interleaved re/im */
PCS_DECL_REAL(pcs_prox_l21, (void* out, const void* x, double tau, int64_t n_groups, int n_comp, int is_complex,
                             void* stream))
/* L21Norm.prox (func/penalty.py:551-557) for an ARBITRARY label array (penalty.py:532-546): groups in CSR form --
 * members_i64[offsets_i64[g] .. offsets_i64[g+1]) are the coordinates of group g (ragged groups of any size) */
PCS_DECL_REAL(pcs_prox_l21_csr, (void* out, const void* x, double tau, const void* members_i64, const void* offsets_i64,
                                 int64_t n_groups, int is_complex, void* stream))
/* out = x * f(sumsq): mode 0 L2Norm.prox (func/penalty.py:23-70), mode 1 proj_l2_ball (math/prox.py:207-210),
 * mode 2 rescale to norm `param` (power iteration, core/linop.py:279-321).  sumsq_f64 = device double holding ||x||^2 */
PCS_DECL_REAL(pcs_scale_by_norm, (void* out, const void* x, const void* sumsq_f64, double param, int mode, int64_t n,
                                  void* stream))

/* the same under the reference's names: L2Norm.prox (mode 0) and proj_l2_ball (mode 1) */
PCS_DECL_REAL(pcs_prox_l2, (void* out, const void* x, const void* sumsq_f64, double tau, int64_t n, void* stream))
PCS_DECL_REAL(pcs_proj_l2_ball, (void* out, const void* x, const void* sumsq_f64, double radius, int64_t n, void* stream))

/* ---- deterministic reductions (np.linalg.norm in update_diagnostics, opt/proxalgs.py:366-394; functional values) -- */
int64_t pcs_reduce_workspace_bytes(void);   /* workspace must be zero-initialised once */
PCS_DECL_REAL(pcs_reduce_sumsq, (const void* x, int64_t n, void* out_f64, void* workspace, void* stream))
PCS_DECL_REAL(pcs_reduce_sumabs, (const void* x, int64_t n, void* out_f64, void* workspace, void* stream))
PCS_DECL_REAL(pcs_reduce_diffsumsq, (const void* x, const void* y, int64_t n, void* out_f64, void* workspace,
                                     void* stream))
PCS_DECL_REAL(pcs_reduce_dot, (const void* x, const void* y, int64_t n, void* out_f64, void* workspace, void* stream))
PCS_DECL_REAL(pcs_reduce_minmax, (const void* x, int64_t n, void* out2_f64, void* workspace, void* stream))
/* One Newton step t' = (S(t) - radius) / (C(t) + c0) on g(t) = sum max(|x_i| - t, 0) - radius - c0 t (S, C = sum and
 * count of the |x_i| above t).  state4_f64 = two (S, C) slots; the step reads slot src_slot (-1: t = 0) and writes
 * dst_slot.  Replaces the brentq bracket of proj_l1_ball (math/prox.py:160-164; radius > 0, c0 = 0) and the sort +
 * cumulative sum of SquaredL1Norm.prox (func/penalty.py:310-316; radius = 0, c0 = 1 / (2 tau)). */
PCS_DECL_REAL(pcs_threshold_newton_step, (const void* x, int64_t n, double radius, double c0, void* state4_f64,
                                          int src_slot, int dst_slot, void* workspace, void* stream))
/* LogBarrier / ShannonEntropy / KLDivergence values (func/penalty.py:819-822, 899-903; func/loss.py:659-664): +inf
 * outside the domain */
PCS_DECL_REAL(pcs_reduce_neglog, (const void* x, int64_t n, void* out_f64, void* workspace, void* stream))
PCS_DECL_REAL(pcs_reduce_xlogx, (const void* x, int64_t n, void* out_f64, void* workspace, void* stream))
PCS_DECL_REAL(pcs_reduce_kl, (const void* x, const void* data, int64_t n, void* out_f64, void* workspace, void* stream))
/* L21Norm.__call__ (func/penalty.py:548-549) */
PCS_DECL_REAL(pcs_reduce_sumgroupnorm, (const void* x, int64_t n_groups, int n_comp, int is_complex, void* out_f64,
                                        void* workspace, void* stream))

/* L21Norm.__call__ for an arbitrary label array (CSR groups, see pcs_prox_l21_csr) */
PCS_DECL_REAL(pcs_reduce_sumgroupnorm_csr, (const void* x, const void* members_i64, const void* offsets_i64,
                                            int64_t n_groups, int is_complex, void* out_f64, void* workspace,
                                            void* stream))

/* ---- finite differences: pylops FirstDerivative/SecondDerivative/Gradient/Laplacian behind
 *      pycsou/linop/diff.py:128-130, 218-219, 882, 957.  kind: 0 forward, 1 centered, 2 backward ------------------ */
PCS_DECL_REAL(pcs_firstderiv_fwd, (void* out, const void* x, int64_t outer, int64_t n, int64_t inner, int kind,
                                   int edge, double step, int accumulate, void* stream))
PCS_DECL_REAL(pcs_firstderiv_adj, (void* out, const void* x, int64_t outer, int64_t n, int64_t inner, int kind,
                                   int edge, double step, int accumulate, void* stream))
PCS_DECL_REAL(pcs_secondderiv_fwd, (void* out, const void* x, int64_t outer, int64_t n, int64_t inner, int kind,
                                    int edge, double step, int accumulate, void* stream))
PCS_DECL_REAL(pcs_secondderiv_adj, (void* out, const void* x, int64_t outer, int64_t n, int64_t inner, int kind,
                                    int edge, double step, int accumulate, void* stream))
/* Integration1D = pylops CausalIntegration(halfcurrent=False) (linop/diff.py:1071-1137): forward
 * y[i] = step * sum_{j<=i} x[j] along the axis, adjoint y[i] = step * sum_{j>=i} x[j] */
PCS_DECL_REAL(pcs_cumsum_fwd, (void* out, const void* x, int64_t outer, int64_t n, int64_t inner, double step, void* stream))
PCS_DECL_REAL(pcs_cumsum_adj, (void* out, const void* x, int64_t outer, int64_t n, int64_t inner, double step, void* stream))
/* 2-D Gradient: out = [d/daxis0 ; d/daxis1] (2*H*W); adjoint: out (H*W) from z (2*H*W) */
PCS_DECL_REAL(pcs_grad2d_fwd, (void* out, const void* x, int64_t H, int64_t W, int kind, int edge, double step0,
                               double step1, void* stream))
PCS_DECL_REAL(pcs_grad2d_adj, (void* out, const void* z, int64_t H, int64_t W, int kind, int edge, double step0,
                               double step1, void* stream))
PCS_DECL_REAL(pcs_laplacian2d_fwd, (void* out, const void* x, int64_t H, int64_t W, double w0, double w1, double step0,
                                    double step1, int edge, void* stream))
PCS_DECL_REAL(pcs_laplacian2d_adj, (void* out, const void* x, int64_t H, int64_t W, double w0, double w1, double step0,
                                    double step1, int edge, void* stream))

/* ---- convolution: pylops.signalprocessing.Convolve1D/Convolve2D behind pycsou/linop/conv.py:163, 294.
 *      forward y[i] = sum_k h[k] x[i+offset-k]; adjoint y[i] = sum_k h[k] x[i+k-offset]; zero "same" boundary ------ */
PCS_DECL_REAL(pcs_conv1d_fwd, (void* out, const void* x, const void* h, int nh, int offset, int64_t outer, int64_t n,
                               int64_t inner, void* stream))
PCS_DECL_REAL(pcs_conv1d_adj, (void* out, const void* x, const void* h, int nh, int offset, int64_t outer, int64_t n,
                               int64_t inner, void* stream))
PCS_DECL_REAL(pcs_conv2d_direct_fwd, (void* out, const void* x, const void* h, int kh, int kw, int offset0, int offset1,
                                      int64_t H, int64_t W, void* stream))
PCS_DECL_REAL(pcs_conv2d_direct_adj, (void* out, const void* x, const void* h, int kh, int kw, int offset0, int offset1,
                                      int64_t H, int64_t W, void* stream))
/* rank-1 filter h = hcol (x) hrow; `sub` (may be NULL) is subtracted from the result: r = A x - y (core/map.py:571-572) */
PCS_DECL_REAL(pcs_conv2d_sep_fwd, (void* out, const void* x, const void* hcol, const void* hrow, int kh, int kw,
                                   int offset0, int offset1, const void* sub, int64_t H, int64_t W, void* stream))
PCS_DECL_REAL(pcs_conv2d_sep_adj, (void* out, const void* x, const void* hcol, const void* hrow, int kh, int kw,
                                   int offset0, int offset1, const void* sub, int64_t H, int64_t W, void* stream))

/* FFT path of Convolve2D for large non-separable filters (the reference's default method='fft', linop/conv.py:168 ->
 * scipy.signal.fftconvolve / correlate): the host runs the padded real transforms (cuFFT), these two kernels do the
 * rest.  pcs_spectral_mul: a <- a * b (conj_b = 0) or a * conj(b) (conj_b = 1) on n_complex interleaved complex numbers,
 * in place.  pcs_crop_shift: out[i][j] = scale * full[(i + shift0) mod P0][(j + shift1) mod P1] - sub[i][j] (sub may be
 * NULL): the "same" crop of the P0 x P1 inverse transform, its normalisation and the residual A x - y in one pass. */
PCS_DECL_REAL(pcs_spectral_mul, (void* a_inout, const void* b, int conj_b, int64_t n_complex, void* stream))
PCS_DECL_REAL(pcs_crop_shift, (void* out, const void* full, const void* sub, int64_t H, int64_t W, int64_t P0,
                               int64_t P1, int64_t shift0, int64_t shift1, double scale, void* stream))

/* ---- sampling: Masking/DownSampling (pycsou/linop/sampling.py:190-196, 362-391), SubSampling = pylops.Restriction
 *      (sampling.py:121).  Elements are opaque 4/8/16-byte words; indices int64, unique for pcs_put ---------------- */
int pcs_take(void* out, const void* x, const void* idx_i64, int64_t outer, int64_t n, int64_t m, int64_t inner,
             int elem_bytes, void* stream);
int pcs_put(void* out, const void* y, const void* idx_i64, int64_t outer, int64_t n, int64_t m, int64_t inner,
            int elem_bytes, void* stream);
/* Pooling / unpooling (linop/sampling.py:525-536 -> skimage block_reduce with zero padding; np.repeat + crop): C-order
 * array of d0 x d1 x d2 samples (1 for unused leading axes), blocks b0 x b1 x b2; scale = 1 (sum) or 1/(b0 b1 b2). */
PCS_DECL_REAL(pcs_pool_fwd, (void* out, const void* x, int64_t d0, int64_t d1, int64_t d2, int64_t b0, int64_t b1,
                             int64_t b2, double scale, void* stream))
PCS_DECL_REAL(pcs_pool_adj, (void* out, const void* y, int64_t d0, int64_t d1, int64_t d2, int64_t b0, int64_t b1,
                             int64_t b2, void* stream))
int64_t pcs_mask_workspace_bytes(int64_t n);
int pcs_mask_count(const void* mask_u8, int64_t n, void* workspace, void* total_out_i64, void* stream);
int pcs_mask_indices(const void* mask_u8, int64_t n, const void* workspace, void* idx_out_i64, void* stream);
/* Allocation-free variant of the pair above.  NO counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * One call = count + scan + write; the chunk counts live in a file-scope `static __device__` array inside the library
 * (128 KiB per GPU that loads the library, held until exit) instead of a caller workspace, so the call allocates nothing.
 *   n <= pcs_mask_static_max_n() = 67 108 864 mask entries (an 8192^2 mask); larger masks are rejected (status 1).
 *   total_out_i64: device int64 <- popcount(mask).  idx_out_i64: device int64[idx_capacity] <- the first
 *   min(total, idx_capacity) sorted positions (never written beyond idx_capacity; size it n, or a known count).
 * All calls on a device share the one scratch array: calls that may overlap on the GPU (different streams) race --
 * serialise them.  Nothing in it has to be cleared between calls (every word read is first written by the same call). */
int64_t pcs_mask_static_max_n(void);
int pcs_mask_to_indices_static(const void* mask_u8, int64_t n, void* total_out_i64, void* idx_out_i64,
                               int64_t idx_capacity, void* stream);

/* ---- fused Condat-Vu (PDS) total-variation iteration: one launch = PrimalDualSplitting.update_iterand +
 *      update_diagnostics (pycsou/opt/proxalgs.py:343-394) for
 *        F = 1/2||A x - y||^2 (func/loss.py:165-219 through core/map.py:596-610), G, H = lam*L21Norm | lam*L1Norm
 *        (func/penalty.py:480-560, 194-245 through core/functional.py:207, 253-265), K = Gradient(kind='forward')
 *        (linop/diff.py:777-882).
 *   x, z -> x_out, z_out (ping-pong, no aliasing); z = [d/daxis0 plane ; d/daxis1 plane], each H_loc x W.
 *   Rows [row_begin, row_end) of the local H_loc x W array are updated; local row 0 is global row grow0 of an image
 *   with Hg rows (slab decomposition: halo rows live in the local array; single GPU: H_loc = Hg, grow0 = 0).
 *   data_mode 0: F = None; 1: g = x - b (b = y); 2: g = (A^T A) x - b with b = A^T y and A^T A given as the separable
 *   autocorrelation taps crow/ccol (HOST arrays of ntaps doubles, ntaps odd <= 29) plus the corner corrections `edge`
 *   (DEVICE, 4 blocks [left,right,top,bottom] of 14 x 14, sizes nl,nr,nt,nb) of the zero "same" boundary; 3: g = b.
 *   g_mode 0 none, 1 NonNegativeOrthant, 2 Segment[g_a,g_b], 3 L1 with threshold g_a.  h_mode 0 L21 (isotropic), 1 L1.
 *   strip_cols / nseg: launch geometry overrides (0 = automatic).  use_tma: 0 plain staging, 1 TMA when the image is
 *   16-byte aligned, 2 require TMA, 3 require TMA and the single-role kernel (deconvolution otherwise runs the
 *   warp-specialised one).  diag_out_f64[4] = ||x'-x||^2, ||x||^2, ||z'-z||^2, ||z||^2 (deterministic).
 *   workspace: pcs_pds_tv_workspace_bytes() bytes (about 23 MB: partial sums, counters and the 21 MB side buffer through
 *   which the blocks of the on-chip resident kernel hand their boundary rows to their neighbours), zero-initialised once. */
int64_t pcs_pds_tv_workspace_bytes(void);
/* What the last pcs_pds_tv_step / _run / _step_slab call of the calling host thread launched (HOST int64[8]):
 * {kernel (0 single-role, 1 warp-specialised with two roles, 2 persistent multi-update, 3 on-chip resident, 4 warp-specialised
 *  with three roles), blocks, threads per block, dynamic shared memory, strips,
 *  output columns per strip, long segments of the uneven one-wave split (-1: even split), TMA staging used}.
 * Diagnostic only: the parity tests assert that the benchmarked kernel and geometry are the ones under test. */
int pcs_pds_tv_last_launch(int64_t* out8_host);
PCS_DECL_REAL(pcs_pds_tv_step,
              (void* x_out, void* z_out, const void* x, const void* z, const void* b, int64_t H_loc, int64_t W,
               int64_t row_begin, int64_t row_end, int64_t grow0, int64_t Hg, int data_mode, int ntaps,
               const double* crow_host, const double* ccol_host, const void* edge, int nl, int nr, int nt, int nb,
               int g_mode, double g_a, double g_b, int h_mode, double lam, double tau, double sigma, double rho,
               double step0, double step1, int strip_cols, int nseg, int use_tma, void* diag_out_f64, void* workspace,
               void* stream))

/* n_iters consecutive updates enqueued by one call: update k reads buffer pair (k & 1 ? b : a), writes the other one
 * and puts its four sums in row k of diag_out_f64 (n_iters x 4 doubles).  The result is in pair a when n_iters is even.
 * For the part of GenericIterativeAlgorithm.iterate (core/solver.py:64-76) in which no stop test can fire (up to
 * min_iter).  Not for data_mode 3. */
/* (pcs_pds_tv_run: when the 7 N s working set stays in L2 -- <= 96 MB -- and there is no convolution, all n_iters updates
 * run in ONE cooperative launch of a persistent kernel with a grid-wide barrier between updates; results are bit-identical
 * to n_iters single launches.  PCS_PDS_PERSIST=0 / 1 in the environment forces it off / on.
 * Smaller still -- no convolution, W even and <= 1024, H <= 7 (f64) / 8 (f32) rows per SM, automatic geometry: x and z
 * stay in the SMs' shared memory for the whole call (pds_tv_resident_kernel), x and z bit-identical to single launches,
 * the sums of diag_out_f64 equal up to summation order; only the result pair is written in full, the other pair keeps
 * whatever it held except for some boundary rows.  PCS_PDS_RESIDENT=0 switches it off.) */
PCS_DECL_REAL(pcs_pds_tv_run,
              (void* x_a, void* z_a, void* x_b, void* z_b, const void* b, int64_t H, int64_t W, int data_mode, int ntaps,
               const double* crow_host, const double* ccol_host, const void* edge, int nl, int nr, int nt, int nb,
               int g_mode, double g_a, double g_b, int h_mode, double lam, double tau, double sigma, double rho,
               double step0, double step1, int strip_cols, int nseg, int use_tma, int64_t n_iters, void* diag_out_f64,
               void* workspace, void* stream))

/* ---- fused ChambollePockSplitting iteration for compressed-sensing MRI with a TV penalty (complex data):
 *      PrimalDualSplitting.update_iterand + update_diagnostics (pycsou/opt/proxalgs.py:343-394, CPS :628-713) for
 *        K = LinOpVStack(Masking * DFT, Gradient(kind='forward'))   (linop/base.py:159-314, sampling.py:125-196,
 *                                                                    diff.py:777-882)
 *        H = ProxFuncHStack(L2BallLoss(m, y, eps), lam * L21Norm)   (func/base.py:21-89, loss.py:69,
 *                                                                    penalty.py:134-191, 480-560),   G = F = None.
 *   _c64: complex64 (interleaved float pairs), _c128: complex128.  Dual variable z = [u (m) ; w0 (N) ; w1 (N)].
 *   One iteration = scatter -> [host: g = DFT^H spec] -> tv_step -> [host: V = DFT v] -> ball_step:
 *     scatter:   spec_out (N) = 0; spec_out[idx] = u                                       (Masking.adjoint)
 *     tv_step:   xt = x - tau (gscale g + Grad^T w); v = 2 xt - x; x' = rho xt + (1-rho) x;
 *                w' = rho fenchel_prox_{sigma lam L21}(w + sigma Grad v) + (1-rho) w;
 *                diag_out[4] = ||x'-x||^2, ||x||^2, ||w'-w||^2, ||w||^2
 *     ball_step: p = u + sigma vscale V[idx]; u' = rho (p - sigma (y + proj_{||.|| <= eps}(p/sigma - y))) + (1-rho) u;
 *                diag_inout[2] += ||u'-u||^2, diag_inout[3] += ||u||^2;  if spec_next != NULL also
 *                spec_next[idx] = u' (a buffer that a scatter has filled before keeps Mask^T u' up to date: the next
 *                iteration needs no scatter)
 *   gscale / vscale: normalisation of the DFTs (1/sqrt(N) each when the host runs unnormalised cuFFT transforms, 1
 *   for orthonormal ones).  workspace: pcs_cps_mri_workspace_bytes() bytes, zero-initialised once, shared by the
 *   three calls of a stream. */
int64_t pcs_cps_mri_workspace_bytes(void);
int pcs_cps_mri_scatter_c64(void* spec_out, const void* u, const void* idx_i64, int64_t N, int64_t m, void* stream);
int pcs_cps_mri_scatter_c128(void* spec_out, const void* u, const void* idx_i64, int64_t N, int64_t m, void* stream);
int pcs_cps_mri_tv_step_c64(void* x_out, void* w_out, void* v_out, const void* x, const void* w, const void* g,
                            int64_t H, int64_t W, double gscale, double lam, double tau, double sigma, double rho,
                            double step0, double step1, void* diag_out_f64, void* workspace, void* stream);
int pcs_cps_mri_tv_step_c128(void* x_out, void* w_out, void* v_out, const void* x, const void* w, const void* g,
                             int64_t H, int64_t W, double gscale, double lam, double tau, double sigma, double rho,
                             double step0, double step1, void* diag_out_f64, void* workspace, void* stream);
int pcs_cps_mri_ball_step_c64(void* u_out, void* spec_next, const void* u, const void* spec, const void* idx_i64,
                              const void* y, int64_t m, double vscale, double eps, double sigma, double rho,
                              void* diag_inout_f64, void* workspace, void* stream);
int pcs_cps_mri_ball_step_c128(void* u_out, void* spec_next, const void* u, const void* spec, const void* idx_i64,
                               const void* y, int64_t m, double vscale, double eps, double sigma, double rho,
                               void* diag_inout_f64, void* workspace, void* stream);

/* ---- a whole AcceleratedProximalGradientDescent solve in one launch: the loop of core/solver.py:55-76 with
 *      update_iterand / update_diagnostics / stopping_metric of opt/proxalgs.py:586-622, for
 *        F = 1/2||h * x - y||^2 (flat Convolve1D, linop/conv.py:20-164), G = none | g_param-threshold L1 | nonneg.
 *   The problem stays in shared memory (n <= pcs_apgd_conv1d_run_max_n(); at most 256 taps): of one block, or -- n >= 512 on a
 *   device that can place the cluster, unless PCS_APGD_CLUSTER=0 -- split over the 2 to 16 blocks of one thread-block
 *   cluster, which hand edge values to their neighbours through distributed shared memory.  Either way the launch runs
 *   updates it0, it0+1, ... while (it <= max_iter and ||x'-x||/||x|| > threshold) or it <= min_iter, exactly as the reference
 *   does (the cluster kernel starts an update before the sums behind the test are in and drops it if they say stop).
 *   g_mode 0 none, 1 L1 (g_param = tau * lambda), 2 NonNegativeOrthant.  accel 0 none, 1 'BT', 2 'CD' (parameter d).
 *   x_inout / aux_inout: iterate and past_aux (in: start, out: result).  diag_out_f64: one row of 4 doubles per update
 *   (||x'-x||^2, ||x||^2, 0, 0).  n_iter_out_i64[0] = update index after the loop, [1] = final t (raw double bits). */
int64_t pcs_apgd_conv1d_run_max_n(void);
/* What this host thread's last pcs_apgd_conv1d_run launched: out4_host = {kernel (0 single block, 1 cluster), blocks,
 * threads per block, elements per thread}.  Diagnostic only (the parity tests assert on it). */
int pcs_apgd_last_launch(int64_t* out4_host);
PCS_DECL_REAL(pcs_apgd_conv1d_run,
              (void* x_inout, void* aux_inout, const void* y, const void* h, int nh, int offset, int64_t n, int g_mode,
               double g_param, double tau, int accel, double d, double t0, int64_t it0, int64_t max_iter,
               int64_t min_iter, double threshold, void* diag_out_f64, void* n_iter_out_i64, void* stream))

/* ---- slab decomposition of one image over the GPUs of a box (the reference has no multi-device
 *      path: this is the partitioned form of PrimalDualSplitting.update_iterand, opt/proxalgs.py:343-355) ----------
 * pcs_pds_tv_step_slab = pcs_pds_tv_step on this rank's rows [row_begin, row_end) of its local H_loc x W arrays (halo
 * rows included), fused with the halo exchange: rows of x', z' that a row neighbour holds as halo are ALSO stored
 * into that neighbour's output arrays through NVLink.  Hand-over: every block that owns rows of the 16-row edge zone
 * at the top / bottom of the slab adds (its zone rows) x (its output columns) to a uint32 counter word in the neighbour's
 * memory when IT is done (system-scope release), so a side of step e is complete when that counter reads e * 16 * W
 * (mod 2^32; epochs count from 1 and the words start at zero).  Blocks that touch halo rows first acquire the value of
 * step epoch-1 on the counter words the neighbours add to.
 * peer_desc_host: HOST array of 20 int64
 *   [0] x_up [1] z_up [2] x_dn [3] z_dn   device pointers to the neighbours' OUTPUT arrays (pcs_peer_open), 0 = none
 *   [4] plane_up [5] plane_dn             elements per z plane of the neighbour's arrays (its H_loc * W)
 *   [6] drow_up  [7] drow_dn              neighbour's local row index minus mine, for the same image row
 *   [8] nx_up [9] nx_dn [10] nz_up [11] nz_dn   my first / last updated rows of x (of z0, z1) that are pushed (<= 16)
 *   [12] wait_up [13] wait_dn             my uint32 counter words, added to by the neighbours
 *   [14] sig_up  [15] sig_dn              uint32 counter words in the neighbours' memory
 *   [16] epoch (>= 1, +1 per step on every rank)   [17] device uint32 error word (set if a wait gave up)
 *   [18] wait timeout in milliseconds (0 = 20 s)
 *   [19] edge trim: rows by which the first / last row segment of every strip is shortened on the sides that have a
 *        neighbour (rounded down to a multiple of 16, capped at 1/32 of the slab): the blocks that hand rows over to a
 *        neighbour finish early and may start late, which hides the NVLink hand-over latency.  0 = equal segments.
 */
PCS_DECL_REAL(pcs_pds_tv_step_slab,
              (void* x_out, void* z_out, const void* x, const void* z, const void* b, int64_t H_loc, int64_t W,
               int64_t row_begin, int64_t row_end, int64_t grow0, int64_t Hg, int data_mode, int ntaps,
               const double* crow_host, const double* ccol_host, const void* edge, int nl, int nr, int nt, int nb,
               int g_mode, double g_a, double g_b, int h_mode, double lam, double tau, double sigma, double rho,
               double step0, double step1, int strip_cols, int nseg, int use_tma, void* diag_out_f64, void* workspace,
               void* stream, const int64_t* peer_desc_host))

/* ---- Condat-Vu total-variation iteration for ANY 2-D Gradient kind (forward / centered / backward, edge rows as in
 *      pylops; pycsou/linop/diff.py:777-882 -- `centered` is pycsou's default) in two kernels per update
 *      (pycsou/opt/proxalgs.py:343-394).  kind: 0 forward, 1 centered, 2 backward.  Arrays are H x W (z: two planes).
 *   primal: a = K^T z; xt = prox_{tau G}(x - tau g - tau a); u_out = 2 xt - x; x_out = rho xt + (1 - rho) x;
 *           diag_out_f64[0..1] = ||x_out - x||^2, ||x||^2.   data_mode 0: g = 0; 1: g = x - b (b = y); 3: g = b (given).
 *           g_mode 0 none, 1 NonNegativeOrthant, 2 Segment[g_a, g_b], 3 L1 with threshold g_a (= tau * lam_G).
 *   dual:   zt = fenchel_prox_{sigma H}(z + sigma K u); z_out = rho zt + (1 - rho) z;
 *           diag_out_f64[2..3] = ||z_out - z||^2, ||z||^2.   h_mode 0: lam * L21 (isotropic TV), 1: lam * L1.
 *   workspace: pcs_pds_tv_workspace_bytes() bytes, zero-initialised once (shared with pcs_pds_tv_step). */
PCS_DECL_REAL(pcs_pds_grad_primal_step,
              (void* x_out, void* u_out, const void* x, const void* b, const void* z, int64_t H, int64_t W, int kind,
               int edge, int data_mode, int g_mode, double g_a, double g_b, double tau, double rho, double step0,
               double step1, void* diag_out_f64, void* workspace, void* stream))
PCS_DECL_REAL(pcs_pds_grad_dual_step,
              (void* z_out, const void* u, const void* z, int64_t H, int64_t W, int kind, int edge, int h_mode, double lam,
               double sigma, double rho, double step0, double step1, void* diag_out_f64, void* workspace, void* stream))

/* Peer memory (CUDA IPC): the only entry points that own device memory.  A rank allocates its slab arena, publishes
 * the pcs_peer_handle_bytes()-byte handle, and maps its row neighbours' arenas. */
int pcs_peer_handle_bytes(void);
int pcs_peer_alloc(void** dptr_out, int64_t bytes, void* handle_out);   /* zero-initialised */
int pcs_peer_open(const void* handle, void** dptr_out);
int pcs_peer_close(void* dptr);
int pcs_peer_free(void* dptr);
int pcs_peer_can_access(int device, int peer);                          /* 1 / 0, negative on error */

#ifdef __cplusplus
}
#endif
#endif /* PCSB200_H */
