/* gpmpc_b200 -- C ABI of the B200-native Gaussian-process hot path of gp-mpc.
 *
 * One shared library (gpmpc_b200/libgpmpc_b200.so, sm_100a only).  Every compute entry point takes raw DEVICE
 * pointers, explicit sizes / leading dimensions (in elements), an explicit cudaStream_t (passed as
 * void*), and returns int:  0 = ok,  < 0 = bad argument (-1) or CUDA error (-(1000 + cudaError_t));
 * numerical failure of the Cholesky is reported LAPACK-style through the device-side `info` word
 * (first non-positive pivot, 1-based; 0 = ok).  Scratch is caller-provided and no compute entry allocates or
 * synchronises in steady state.  The exceptions, all one-time per device and all avoidable by calling gpk_init first:
 * the 8 KB exp table of the prediction front end (cudaMalloc + copy + one stream synchronisation), the library-owned
 * side stream + event pool of gpk_potrf_trtri (and the side stream + five events of gpk_predict_i8's optional
 * pipeline, created only when GPK_OZ_PIPELINE enables it), and -- not covered by gpk_init -- the item list of an int8 fit GEMM
 * the first time its SHAPE is seen on a device (cudaMalloc + copy + one stream synchronisation; a fit of one size
 * sees about ten shapes, all in its first iteration).  gpk_shutdown frees all of it.  Per-device state is not
 * thread-safe: one host thread (or process) per GPU, calls on one stream at a time (SURVEY.md 8b "Threading").
 * gpk_last_error() returns the message of the last failing call on this thread.
 *
 * The reference has no FFI of its own: its GP layer is Python over gpytorch/torch.  Each
 * entry below names the reference call site (file:line under the gp-mpc tree) whose library work it
 * replaces; the Python binding a maintainer would add is shown in INTEGRATION.md.
 *
 * Conventions
 *   - matrices are row-major; "npad" is N rounded up to a multiple of 128, padded matrices carry an
 *     identity block on the padding (K, L, W, U, K^-1) so that padded rows/columns are inert;
 *   - dtype: 0 = float32, 1 = float64 (prediction entry points only; the fit is always float64, as in
 *     the reference where float64 data promotes the kernel matrix, gpmpc/gpmpc.py:155-156);
 *   - hyp: per-GP hyper-parameter block of gpk_hyp_stride() doubles in device memory:
 *       [0..31] centre c_j, [32..63] 1/ell_j, [64] sf2 (outputscale), [65] noise (sigma^2);
 *     inputs are used as (x - c) / ell, which is what gpytorch's RBFKernel does before its distance;
 *   - batch: number of independent GPs of identical shape processed by one launch (grid.z); per-GP
 *     buffers are addressed with the given element strides, a stride of 0 shares the buffer.
 */
#ifndef GPMPC_B200_H
#define GPMPC_B200_H

#ifdef __cplusplus
extern "C" {
#endif

/* ---- library info ------------------------------------------------------------------------- */
int gpk_abi_version(void);              /* 2 */
int gpk_init(void* stream);             /* create the current device's library state up front (see above); idempotent */
int gpk_shutdown(void);                 /* free it (device-synchronising); the library stays usable */
const char* gpk_last_error(void);
long gpk_launch_count(void);            /* kernels launched by this library since load / last reset */
void gpk_reset_launch_count(void);
int gpk_pad_dim(int d);                 /* padded input dimension used by packs (d <= 32), -1 if unsupported */
int gpk_hyp_stride(void);               /* doubles per hyper-parameter block */
int gpk_train_tile(void);               /* training points per pack tile */

/* ---- (1) kernel assembly -------------------------------------------------------------------
 * mode 0: out = K_f(A, B)           replaces gp.covar_module(a, b).to_dense()      gpmpc/gpmpc.py:392-393
 * mode 1: out = K_f(A, A) + noise I replaces covar_module(X).add_diag(noise)       gpmpc/gp.py:45, and the
 *         lazy K that gpytorch densifies inside the MLL                            gpmpc/gp.py:61-62
 * kernel formula: gpmpc/gp.py:12-14.  A is dtype_a (0/1), B float64; out is rows_pad x cols_pad
 * (multiples of 64) with zeros (mode 0) or identity (mode 1) on the padding. */
int gpk_cross_kernel(int dtype_a, const void* A, long lda, long na, long strideA,
                     const double* B, long ldb, long nb, long strideB, int d, const double* hyp,
                     double* out, long ldo, long strideO, long rows_pad, long cols_pad,
                     int mode, int batch, void* stream);

/* ---- (2) blocked Cholesky ------------------------------------------------------------------
 * In-place lower Cholesky of the padded npad x npad matrix; replaces the potrf that gpytorch's
 * psd_safe_cholesky issues for gpmpc/gp.py:62.  dinv receives the inverses of the npad/128 diagonal
 * blocks of L (npad * 128 doubles).  *info_dev = first failing pivot (1-based) or 0. */
int gpk_potrf(double* A, long lda, long npad, double* dinv, int* info_dev, void* stream);

/* ---- (4) triangular inverse and K^-1 --------------------------------------------------------
 * W = L^-1 (lower) and U = W^T (upper) from the factor and dinv of gpk_potrf; scratch of
 * gpk_trtri_scratch_elems(npad) doubles.  gpk_kinv forms K^-1 = U U^T (dense, symmetric): replaces
 * torch.linalg.inv(K) at gpmpc/gp.py:46 and feeds the MLL gradient. */
long gpk_trtri_scratch_elems(long npad);
int gpk_trtri(const double* L, long ldl, long npad, const double* dinv, double* W, long ldw,
              double* U, long ldu, double* tmp, long tmp_elems, void* stream);
/* gpk_potrf followed by gpk_trtri in one call: same kernels and results, but the inverse factor is computed beside
 * the factorisation on a side stream of the library (forked after the work already enqueued on `stream`, joined
 * before this call's effects are visible to later work on `stream`), which fills the SMs the small dependent kernels
 * of the Cholesky recursion leave idle. */
int gpk_potrf_trtri(double* A, long lda, long npad, double* dinv, int* info_dev, double* W, long ldw,
                    double* U, long ldu, double* tmp, long tmp_elems, void* stream);
int gpk_kinv(const double* U, long ldu, long npad, double* Kinv, long ldk, void* stream);
/* Rows [r0, r1) and columns [0, ncols) (multiples of 128) of K^-1: what one rank of a row-split MLL gradient needs
 *  -- ncols = npad for gpk_mll_grad_reduce, ncols = r1 for gpk_mll_grad_reduce_sym; the rest of Kinv is
 * not touched. */
int gpk_kinv_rows(const double* U, long ldu, long npad, long r0, long r1, long ncols, double* Kinv, long ldk,
                  void* stream);

/* ---- (3) alpha, y^T K^-1 y, log det K --------------------------------------------------------
 * alpha = K^-1 y = U (W y); scal[0] = y^T K^-1 y, scal[1] = log det K (from diag L).  Replaces the
 * solves / inv_quad_logdet inside ExactMarginalLogLikelihood (gpmpc/gp.py:57,62) and the prediction
 * strategy's mean cache (gpmpc/gpmpc.py:444).  y, t_work, alpha have npad entries (y zero-padded). */
int gpk_solve_alpha(const double* W, long ldw, const double* U, long ldu, const double* L, long ldl,
                    long npad, const double* y, double* t_work, double* alpha, double* scal,
                    void* stream);

/* ---- (2-5 fused) one-launch MLL + gradient for small GPs -------------------------------------------
 * The whole iteration of gpmpc/gp.py:59-63 (assembly, Cholesky, alpha, y^T K^-1 y, log det K, K^-1, gradient trace)
 * for n <= gpk_mll_small_max_n() = 128 training points -- the reference's own sizes, N <= 45
 * (scripts/gp_mpc_config.yaml:24,30) -- by ONE CTA per GP out of shared memory; `batch` GPs of identical (n, d) run
 * concurrently (the three GPs gpmpc/gpmpc.py:159-162 fits one after the other).
 *   out[gp][0] = potrf info (first non-positive pivot, 1-based; 0 = ok), [1] = y^T K^-1 y, [2] = log det K,
 *   out[gp][4 + v] = the gpk_pad_dim(d) + 2 sums of gpk_mll_grad_reduce; stride gpk_mll_small_out_stride() doubles.
 *   alpha (may be NULL) receives K^-1 y, n entries per GP.  K / Kinv (both or neither) receive the dense n x n
 *   matrices K_f + noise I and its inverse -- gp.K, gp.K_inv of compute_covariances, gpmpc/gp.py:43-46. */
int gpk_mll_small_max_n(void);
int gpk_mll_small_out_stride(void);
int gpk_mll_small(const double* X, long ldx, long strideX, const double* y, long stride_y, long n, int d,
                  const double* hyp, double* out, double* alpha, long stride_alpha, double* K, double* Kinv,
                  long ldk, long strideK, int batch, void* stream);

/* ---- (5) MLL gradient trace reduction --------------------------------------------------------
 * With G = alpha alpha^T - K^-1, over rows [r0, r1) and all columns < n:
 *   out[0] = sum G_aa, out[1] = sum G_ab K_f,ab, out[2+j] = sum G_ab K_f,ab ((x_aj - x_bj)/ell_j)^2.
 * Replaces loss.backward() at gpmpc/gp.py:63 (closed form instead of autograd).  Row ranges let one
 * GP's trace be split over GPUs; the (d+2) partials are then all-reduced by the caller. */
long gpk_grad_trace_work_elems(long n, int d, long r0, long r1);
int gpk_mll_grad_reduce(const double* X, long ldx, long n, int d, const double* hyp,
                        const double* Kinv, long ldk, const double* alpha, long r0, long r1,
                        double* work, long work_elems, double* out, void* stream);
/* The same sums from the lower triangle of K^-1 only (the summand is symmetric: columns b < a counted twice, b > a
 * skipped): half the traffic and exponentials.  The partial of a row range differs from gpk_mll_grad_reduce's, the sum
 * over a partition of the rows is the same. */
int gpk_mll_grad_reduce_sym(const double* X, long ldx, long n, int d, const double* hyp,
                        const double* Kinv, long ldk, const double* alpha, long r0, long r1,
                        double* work, long work_elems, double* out, void* stream);
/* Allocation-free variant of the two entries above (sym = 0: gpk_mll_grad_reduce, sym = 1: gpk_mll_grad_reduce_sym).
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid allocation).  The per-CTA partial sums
 * live in a file-scope `static __device__` array of the library instead of a caller workspace; the call performs no
 * cudaMalloc / cudaFree and calls no allocating library routine.  Inputs and `out` (d_pad + 2 doubles, device) are the
 * caller's.  Limit: gpk_grad_trace_work_elems(n, d, r0, r1) <= gpk_mll_grad_reduce_static_max_elems() = 139264 doubles
 * (1.06 MiB, held by every GPU context that loads the library until exit) -- ceil((r1 - r0) / 64) * ceil(n / 64) * (d_pad + 2) partials: a full pass up to
 * n = 4096 at any d <= 32 (n = 11904 at d = 2), or a row slice of a larger GP; beyond it the call returns an error
 * naming the limit and launches nothing.  Every call on a device shares the one array (all slots a launch reads are
 * written by that launch first, so nothing needs clearing between calls): calls that may OVERLAP on the GPU -- two
 * streams, two host threads -- race on it and must be serialised by the caller; calls on one stream are safe. */
long gpk_mll_grad_reduce_static_max_elems(void);
int gpk_mll_grad_reduce_static(const double* X, long ldx, long n, int d, const double* hyp,
                        const double* Kinv, long ldk, const double* alpha, long r0, long r1,
                        int sym, double* out, void* stream);

/* ---- (6) fused posterior mean + input-Jacobian ------------------------------------------------
 * gpk_pack_train builds the resident training pack (scaled points, -1/2|x|^2, sf2*alpha) in the
 * prediction dtype.  gpk_predict_mean_jac evaluates mean(z) = k(z,X) alpha (gpmpc/gp.py:85) and
 * jac(z) = d mean / d z (CasADi AD of gpmpc/gp.py:12-14 inside acados, gpmpc/gpmpc.py:189-211) for q
 * queries; K_ZX is never written.  work may be NULL (then the training set is not split over CTAs). */
long gpk_pack_train_elems(long n, int d);
int gpk_pack_train(int dtype, const double* X, long ldx, long n, int d, const double* alpha,
                   const double* hyp, void* pack, long pack_elems, void* stream);
long gpk_mean_jac_work_bytes(int dtype, long q, long n, int d, int batch);
int gpk_predict_mean_jac(int dtype, const void* Z, long ldz, long q, long strideZ, int d,
                         const double* hyp, const void* pack, long stride_pack, long n,
                         void* mean, long stride_mean, void* jac, long ldj, long stride_jac,
                         void* work, long work_bytes, int batch, void* stream);

/* ---- (7) posterior variance ---------------------------------------------------------------------
 * var_f(z) = sf2 - || W k(X,z) ||^2: the diagonal that gpmpc/gpmpc.py:443-445 keeps of
 * gp.likelihood(gp(z)).covariance_matrix, without the likelihood noise (the caller adds hyp noise).
 * The Q x Q covariance is never formed.  Scratch: gpk_predict_var_work_bytes(q_chunk, npad, batch);
 * queries are processed in chunks that fit the scratch given. */
long gpk_predict_var_work_bytes(long q_chunk, long npad, int batch);
int gpk_predict_var(int dtype, const void* Z, long ldz, long q, long strideZ, int d, const double* hyp,
                    const double* X, long ldx, long n, long strideX, const double* W, long ldw,
                    long strideW, long npad, void* var, long stride_var, void* work, long work_bytes,
                    int batch, void* stream);
/* the dominant kernel of (7) on its own: colsq[tile_m][q] = sum over the 128 rows of tile_m of (W K_ZX^T)[i][q]^2 */
int gpk_var_colsq(const double* W, long ldw, long strideW, const double* kzx, long ldk, long strideK,
                  long npad, long qpad, double* colsq, long ld_colsq, long stride_colsq, int batch,
                  void* stream);
int gpk_var_finalize(int dtype, const double* colsq, long ld, long stride_colsq, int tiles_m, long q,
                     const double* hyp, void* var, long stride_var, int batch, void* stream);

/* ---- (6+7) fused prediction on the tensor pipes: mean + Jacobian + variance in one pass ---------------
 * The default prediction path.  The kernel k(z, X) (gpmpc/gp.py:12-14) is evaluated ONCE per (query, training
 * point) pair by DMMA GEMMs + exp and feeds
 *   - mean(z) = k(z,X) alpha (gpmpc/gp.py:85) and jac(z) = d mean / d z (CasADi AD inside acados,
 *     gpmpc/gpmpc.py:189-211), and
 *   - var_f(z) = sf2 - || W k(X,z) ||^2, the diagonal gpmpc/gpmpc.py:443-445 keeps of
 *     gp.likelihood(gp(z)).covariance_matrix (without the noise), whose contraction runs on the tcgen05 int8
 *     tensor cores: fp64 operands are split error-free into `digits` balanced base-256 int8 slices ("Ozaki
 *     scheme"), the S(S+1)/2 slice products accumulate exactly in int32 (TMEM) and are recombined in fp64.
 *     digits = 4 / 5 / 6 keep 30 / 38 / 46 bits (5: ~2e-9 relative variance error at N = 2048).
 * gpk_oz_slice_w builds the digit image of W = L^-1 once per fit (image layout: csrc/ozaki.cuh).
 * mean / jac may both be NULL (variance only).  Scratch: gpk_predict_i8_work_bytes(q_chunk, ...); queries are
 * processed in chunks that fit the scratch given.  mj_done_event (cudaEvent_t as void*, may be NULL) is
 * recorded on `stream` as soon as mean and jac are final, before the variance contraction of the last chunk. */
int gpk_oz_default_digits(void);
long gpk_oz_max_npad(int digits);                          /* largest npad with overflow-free int32 accumulation: 32768 at
                                                              5 digits, 26368 at 6, 43904 at 4; -1 if unsupported digits */
int gpk_oz_query_tile(void);                               /* queries per tile of the E image (128) */
long gpk_oz_w_image_bytes(long npad, int digits);          /* -1 if unsupported */
long gpk_oz_e_image_bytes(long qpad, long npad, int digits);
int gpk_oz_pieces(long qtiles_total, long npad);           /* partial-sum pieces per query tile used by the GEMM */
int gpk_oz_slice_w(const double* W, long ldw, long npad, int digits, void* img, double* wscale, void* stream);
long gpk_predict_i8_work_bytes(long q_chunk, long npad, int d, int digits, int batch);
int gpk_predict_i8(int dtype, const void* Z, long ldz, long q, long strideZ, int d, const double* hyp,
                   const double* X, long ldx, long n, long strideX, const double* alpha, long stride_alpha,
                   const void* Wimg, long strideW, const double* wscale, long stride_ws, long npad, int digits,
                   void* mean, long stride_mean, void* jac, long ldj, long stride_jac, void* var, long stride_var,
                   void* work, long work_bytes, int batch, int mul_digits, void* mj_done_event, void* stream);
/* mul_digits: 0 = multiply all `digits` digits of the image; 4 <= mul_digits < digits multiplies only the leading
 * mul_digits digits of both operands (fewer digit products, coarser variance: the caller's accuracy policy decides,
 * nothing is cut implicitly by dtype). */
/* mean + Jacobian alone on the same DMMA front kernel (no digit image, no variance): the want_var = false path in
 * float64 for d >= 9, where it is faster than the SIMT kernel of (6).  Takes X and alpha directly (no pack). */
long gpk_mean_jac_dmma_work_bytes(long q_chunk, int d, int batch);
int gpk_predict_mean_jac_dmma(int dtype, const void* Z, long ldz, long q, long strideZ, int d, const double* hyp,
                              const double* X, long ldx, long n, long strideX, const double* alpha,
                              long stride_alpha, void* mean, long stride_mean, void* jac, long ldj, long stride_jac,
                              void* work, long work_bytes, int batch, void* stream);
/* the two stages of (6+7) on their own, for tests and bench.py:
 *   gpk_oz_cross         queries -> int8 digit image of E = exp(-r^2/2), [batch][qpad/128][npad/64][digits][128x64]
 *   gpk_oz_var_partials  part[gp][slot][q], slot < 2 n_pieces: sums of (W k)_i^2 over disjoint sets of rows i;
 *                        dbg (optional) receives the raw int32 accumulators [tile][row block][digit group][128][64] */
int gpk_oz_cross(int dtype, const void* Z, long ldz, long q, long strideZ, const double* X, long ldx, long n,
                 long strideX, int d, const double* hyp, void* img, long strideE, long npad, long qpad, int digits,
                 int batch, void* stream);
int gpk_oz_var_partials(const void* Eimg, long strideE, const void* Wimg, long strideW, const double* wscale,
                        long stride_ws, const double* hyp, long npad, long qpad, int digits, double* part,
                        long ld_part, long stride_part, int n_pieces, int batch, int* dbg, void* stream);

/* ---- (f2) chance-constraint tightening, batched over rollouts ------------------------------------
 * Replaces GPMPC.propagate_constraint_limits (gpmpc/gpmpc.py:425-498) for `rollouts` independent rollouts of T
 * stages: stage disturbance covariance from the three GPs' variances and the roll / pitch angles of the stage
 * input z (gpmpc.py:447-470), the LQR-closed-loop covariance recursion (gpmpc.py:478-495) and the limits
 * -inverse_cdf |A| sqrt(diag cov) for A = [-I; I] (gpmpc.py:472-476, 487-488, 497).
 *   var   [3][rollouts*T] (GP g at var + g*stride_var): latent var_f as gpk_predict_i8 returns it
 *         (var_has_noise = 0; the likelihood noise of hyp is added twice, as gpmpc.py:444 + 485 do) or the
 *         predictive variance gp.likelihood(gp(z)) (var_has_noise = 1; noise added once more)
 *   Z     [rollouts*T][ldz] stage inputs; only columns idx_phi, idx_theta are read (6, 7 in gpmpc.py:449)
 *   hyp   3 hyper-parameter blocks;  A (nx x nx), B (nx x nu), K (nu x nx), Bd (nx x 5) row-major:
 *         discrete_dfdx, discrete_dfdu, lqr_gain, Bd of gpmpc.py:69, 84
 *   state_constraint [rollouts][2 nx][T+1], input_constraint [rollouts][2 nu][T]  (the reference's arrays per
 *         rollout).  nx <= 16, nu <= min(8, nx).  A variance rounded below zero gives 0, not NaN. */
int gpk_tighten_constraints(const double* var, long stride_var, int var_has_noise, const double* Z, long ldz,
                            int idx_phi, int idx_theta, const double* hyp, const double* A, const double* B,
                            const double* K, const double* Bd, int nx, int nu, int T, long rollouts, double dt,
                            double inverse_cdf, double* state_constraint, double* input_constraint, void* stream);

/* ---- (f1) batched discrete stage dynamics with sensitivities -----------------------------------------
 * For q independent (x, u) pairs: x_next = f_disc(x, u), A = d f_disc/dx (12 x 12), B = d f_disc/du (12 x 4) of the
 * acados model GPMPC.setup_acados_model builds (gpmpc/gpmpc.py:166-209): classical RK4 (gpmpc.py:201-206) of
 * f_cont = prior + res_dyn (gpmpc.py:193-199) with the thrust / roll / pitch GP means at z[idx_T], z[idx_R], z[idx_P]
 * (gpmpc.py:172, 187-191).  What acados evaluates per stage and SQP iteration, batched over stages and rollouts.
 *   x [q][ldx >= 12], u [q][ldu >= 4] row-major (state order x, dx, y, dy, z, dz, phi, theta, psi, dphi, dtheta, dpsi)
 *   prior[n_prior] (HOST memory): a, b, c, d, e, f, h, l (scripts/gp_mpc_config.yaml:9-17), g [, yaw c, d, e]: the attitude model
 *         of crazyflow.sim.symbolic.symbolic_attitude as the reference uses it (gpmpc.py:128-146); 9 <= n_prior <= 12
 *   hyp, pack: three GPs (thrust, roll, pitch) with inputs padded to d = 3 (zero columns for the thrust GP) and the
 *         same n; pack from gpk_pack_train in float64, GP g at pack + g*stride_pack elements.  For the FITC model
 *         (gpmpc.py:175-186) pack the inducing inputs with the weights of precompute_sparse_posterior_mean.
 *   x_next [q][12], A [q][12][12], B [q][12][4];  work: gpk_stage_rk4_work_bytes(q_chunk, n, 3) bytes, pairs are
 *         processed in chunks that fit. */
long gpk_stage_rk4_work_bytes(long q_chunk, long n, int batch);
int gpk_stage_rk4(const double* x, long ldx, const double* u, long ldu, long q, const double* prior, int n_prior,
                  double dt, const double* hyp, const void* pack, long stride_pack, long n, double* x_next,
                  double* A, double* B, void* work, long work_bytes, void* stream);

/* ---- building block exposed for tests ---------------------------------------------------------
 * C = beta C + alpha A B^T on the FP64 DMMA pipe; M, N multiples of 128, K of 16.
 * a_tri: 0 full, 1 A lower triangular, 2 A upper triangular; c_lower: only tiles on/below the diagonal. */
int gpk_dgemm_nt(const double* A, long lda, const double* B, long ldb, double* C, long ldc,
                 long M, long N, long K, double alpha, double beta, int a_tri, int c_lower, void* stream);

/* ---- fit GEMMs on the tcgen05 int8 tensor pipe ---------------------------------------------------
 * The O(N^3) steps behind gpmpc/gp.py:62 (Cholesky, solves) and gpmpc/gp.py:46 (inverse) are GEMMs of the contract
 * above.  gpk_oz_dgemm_nt runs one of them error-free on int8 digit slices (`digits` = 5/6/7 balanced base-256
 * digits per operand, power-of-two scales per row and chunk of 4096 k; exact int32 accumulation in TMEM, fp64
 * recombination): M, N multiples of 128, K of 64, a_tri != 0 needs M <= K, C must not alias A or B; when A is B or a
 * 128-aligned window of B's rows one digit image serves both operands.
 * Scratch: gpk_oz_dgemm_nt_work_bytes(M, N, K, same = (A == B), digits).
 * gpk_oz_gemm_set_workspace registers per-device scratch (gpk_oz_gemm_work_bytes(npad, digits)) that makes
 * gpk_potrf / gpk_trtri / gpk_kinv / gpk_dgemm_nt route every GEMM with M, N >= min_dim (0 = keep the default of
 * 2048) and K >= 1024 through this path; ptr == NULL or digits == 0 restores the DMMA path.  The caller owns the
 * memory and must keep it alive while registered. */
long gpk_oz_dgemm_nt_work_bytes(long M, long N, long K, int same, int digits);
int gpk_oz_dgemm_nt(const double* A, long lda, const double* B, long ldb, double* C, long ldc, long M, long N,
                    long K, double alpha, double beta, int a_tri, int c_lower, int digits, void* work,
                    long work_bytes, void* stream);
long gpk_oz_gemm_work_bytes(long npad, int digits);
int gpk_oz_gemm_set_workspace(void* ptr, long bytes, int digits, long min_dim);

#ifdef __cplusplus
}
#endif
#endif /* GPMPC_B200_H */
