/* sgp_b200.h -- C ABI of the B200-native Spatial_GP ELBO+gradient hot path.
 *
 * The reference (retinal-information-processing-lab/gaussian_processes, Spatial_GP_repo/utils.py) has no
 * FFI of its own: its boundary is the set of module-level Python callables the notebooks use (SURVEY.md
 * 8b).  This header is what a `utils`-compatible Python module binds through ctypes
 * (gaussian_processes_b200/utils.py does exactly that; INTEGRATION.md shows the stub).  Each entry point
 * cites the reference callable (utils.py line range) whose tensor work it replaces.
 *
 * Conventions
 *   - every function returns an int status (SGP_OK == 0); sgp_last_error(h) gives the message;
 *   - sizes must be positive and every leading dimension at least the row length it strides over: anything else is
 *     refused with SGP_E_ARG before a kernel is launched (the library cannot know the size of the caller's buffers);
 *   - all matrices are fp64, row-major, caller-allocated DEVICE memory unless the name ends in _host;
 *   - `ld*` are leading dimensions in elements; `stream` is a cudaStream_t passed as void*;
 *   - the library never owns user-visible memory; the handle owns scratch workspace only;
 *   - theta is 6 doubles in the reference's dict order (utils.py:824):
 *        [sigma_0, eps_0x, eps_0y, -2log2beta, -log2rho2, Amp];
 *   - there is no CPU fallback: without a CUDA device every compute call returns SGP_E_CUDA.
 */
#ifndef SGP_B200_H_
#define SGP_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define SGP_API __attribute__((visibility("default")))
#else
#define SGP_API
#endif

typedef struct sgp_handle sgp_handle;

enum {
  SGP_OK = 0,
  SGP_E_CUDA = 1,        /* CUDA runtime error (message has the cudaError string) */
  SGP_E_ARG = 2,         /* bad argument (null pointer, odd leading dimension, m too large, ...) */
  SGP_E_BOUNDS = 3,      /* theta outside [lo, hi]: reference raises ValueError (utils.py:865-867) */
  SGP_E_UNSUPPORTED = 4, /* configuration not implemented on the device path */
  SGP_E_NOMEM = 5
};

/* status bits returned in out[8] of sgp_elbo_grad and by the factorisation helpers */
enum {
  SGP_FLAG_KT_NOT_PD = 1,      /* K_tilde_b not positive definite (log_det takes the eigh fallback, utils.py:1279) */
  SGP_FLAG_KT_SMALL_PIVOT = 2, /* Cholesky pivot < 1e-10 (safe_log raises, utils.py:670) */
  SGP_FLAG_V_NOT_PD = 4,
  SGP_FLAG_V_SMALL_PIVOT = 8,
  SGP_FLAG_NAN = 16            /* NaN in the mean rate f (utils.py:1923) */
};

SGP_API int sgp_version(void);
SGP_API int sgp_create(int device, sgp_handle** out);
SGP_API int sgp_destroy(sgp_handle* h);
SGP_API const char* sgp_last_error(const sgp_handle* h);
/* rows of the image pass processed per slab (bounds the scratch memory); 0 = default (65536) */
SGP_API int sgp_set_slab_rows(sgp_handle* h, long rows);
/* options: "kl_weight" (weight of the KL adjoint in the returned gradient; ranks > 0 of an image-sharded
 * evaluation use 0 so that the all-reduced sum counts the N-independent KL term once), "profile" (1 = record
 * CUDA events at the stage boundaries of sgp_elbo_grad), "slab_rows", "path" (0 = choose, 1 = dense C products,
 * 2 = separable quadratic forms), "i8_slices" (-1 = choose, 0 = the two N x p x m products of sgp_elbo_grad on the
 * fp64 tensor pipe, 2..8 = on tcgen05.mma.kind::i8 with that many int8 slices per operand; see sgp_gemm_i8emu;
 * 8 is the error class of a DGEMM and the default, 4 gives 27 bits per operand -- the fp32-class setting),
 * "i8_slices_images" (-1 = same as "i8_slices", 2..8 = slice count of the two products that contract with the
 * images, K_lin = X (UC)^T and D = X^T R; the m-space products keep "i8_slices"),
 * "i8_slices_xtr" / "i8_slices_sig2" (-1 = default, 2..8: slice counts of the two products that feed only the gradient,
 * D = X^T R (at most the images' count) and Sigma2 = K^T diag(v) K (default 7 when "i8_slices" is on auto)),
 * "support" (pixel set the cached copies / int8 slices of the images are built on: -1 = the receptive-field mask grown
 * by two pixels, kept while later masks stay inside it -- the mask is recomputed from theta in every closure
 * (utils.py:2030) and flips pixels during a line search; 0 = exactly the mask, as the reference gathers x[:, mask];
 * 1 = the whole grid, images read in place.  Pixels of the support outside the mask carry C = 0, i.e. exact zeros),
 * "square_is_same" (1 = a call with N == m is the reference's ntilde == nt branch, K = K_tilde and a = B,
 * utils.py:2032 / 2068 -- the default, as the reference decides by the row counts alone; 0 = N == m is treated like
 * any other N: set by image-sharded callers, whose local row count says nothing about the global one),
 * "projected" (-1 = moments in the projected basis K_b = K B when q <= 0.4 m, 0 = never, 1 = always; same numbers
 * up to rounding, utils.py:2047-2068), "overlap" (1 = the q x q chains run on side streams),
 * "chol_fused" (largest matrix order whose Cholesky factor / inverse / log det run as ONE launch of a thread-block
 * cluster, csrc/chol_fused.cu; default 320, at most 1024, 0 = always one launch per block step),
 * "sep_uc" (1, the default = on the separable path U C and the inducing norms come from the Kronecker structure of C in
 * one launch; 0 = build C, slice, multiply).
 * With int8 slices the library keeps a sliced copy of device-resident images across calls (same bytes as X).     */
SGP_API int sgp_set_option(sgp_handle* h, const char* name, double value);
/* accumulated per-stage device milliseconds of sgp_elbo_grad since the last reset:
 * [0] setup + inducing-point forward, [1] per-image quadratic forms (separable kernel, or the dense X*C GEMM),
 * [2] GEMM X (U C)^T (dense path: Z U^T), [3] moments GEMM, [4] per-image moments + adjoint of the acos map,
 * [5] sum w_i x_i^T dC_k x_i (dense path: R*U GEMM), [6] GEMM X^T R (dense path: X^T Y), [7] K^T diag(v) K GEMM,
 * [8] q x q adjoint, inducing-point backward, <dC,GC> contraction, [9] arc-cosine map (streaming pass),
 * [10] with int8 slices: the X^T R product itself ([6] then holds the transposed slicing of R).                 */
SGP_API int sgp_get_profile(sgp_handle* h, double* stage_ms_out, int n, long* calls_out, int reset);
/* forget the cached masked copies and int8 slices of X / xtilde (call if their contents changed in place) */
SGP_API int sgp_invalidate_cache(sgp_handle* h);
/* same, and give the memory back (gathered copies, int8 slices, staging buffers, scratch arena) */
SGP_API int sgp_release_cache(sgp_handle* h);
/* Content identity of the image buffers (`x` / `xtilde` of varGP, utils.py:1568: constant during a fit).  The caches
 * above are keyed on (token, pointer, sizes, strides, mask): a pointer alone does not identify its content, because
 * allocators hand a freed block to the next tensor of the same size.  Callers pass a NEW token whenever the images
 * they are about to pass may differ from the last ones seen under the current token (another recording, a resample,
 * an in-place edit); the Python binding derives it from (tensor identity, torch version counter, pointer, shape).  */
SGP_API int sgp_bind_images(sgp_handle* h, unsigned long long token);
/* number of kernels this library has launched through `h` since creation / last reset */
SGP_API long sgp_launch_count(const sgp_handle* h, int reset);

/* ---- a1: localker(theta, hi, lo, n_px_side, grad)  utils.py:861-914 ---------------------------------------
 * sgp_mask_host: RF mask alpha >= 1e-3 on the n_px x n_px grid (HOST output, n_px^2 bytes) and p = mask.sum().
 * sgp_localker : C (p x p, ld = ldc >= p) and optionally dC (5 matrices, consecutive, each ldc*p.. see dc_stride)
 *                in the reference's dict order Amp, -2log2beta, -log2rho2, eps_0x, eps_0y (utils.py:910).     */
SGP_API int sgp_mask_host(const double* theta_host, int n_px, uint8_t* mask_host, int* p_out);
SGP_API int sgp_localker(sgp_handle* h, const double* theta_host, int n_px, double* C, long ldc, double* dC, long dc_stride,
                 void* stream);

/* ---- a2: acosker(theta, x1, x2, C, dC=None, diag=False)  utils.py:975-1025 ---------------------------------
 * x1 (n1 x p, ld1), x2 (n2 x p, ld2) are the MASKED images; K (n1 x n2, ldk).  symmetrize != 0 applies
 * K = (K + K^T)/2 (the reference does so whenever n1 == n2).                                                   */
SGP_API int sgp_acosker(sgp_handle* h, const double* x1, long n1, long ld1, const double* x2, long n2, long ld2, int p,
                const double* C, long ldc, double sigma_0, double* K, long ldk, int symmetrize, void* stream);
/* K as above plus Kvec_i = x1_i^T C x1_i + sigma_0^2 (utils.py:1029) in the same pass: the kernel forms it anyway as
 * X1^2 (utils.py:978), so varGP's "K, then Kvec" pair (utils.py:1677-1680, 1804-1806) costs one x1 C product.      */
SGP_API int sgp_acosker_kvec(sgp_handle* h, const double* x1, long n1, long ld1, const double* x2, long n2, long ld2,
                     int p, const double* C, long ldc, double sigma_0, double* K, long ldk, int symmetrize,
                     double* Kvec, void* stream);
/* acosker(theta, x1, x2, C) for images x1 whose Kvec (= x1_i^T C x1_i + sigma_0^2, the `diag=True` result at THIS theta:
 * the fitted model's 'Kvec') is already known: the x1 C product is not repeated, one pass over x1.  The active-learning
 * cycle's "one new column of K" (one_cell_active_training.ipynb cell 17; same values as sgp_acosker, never symmetrised:
 * n1 != n2 there).                                                                                                 */
SGP_API int sgp_acosker_known(sgp_handle* h, const double* x1, long n1, long ld1, const double* kvec1, const double* x2,
                      long n2, long ld2, int p, const double* C, long ldc, double sigma_0, double* K, long ldk,
                      void* stream);
/* ---- a3: acosker(theta, x1, x2=None, C, diag=True)  utils.py:1027-1030 : Kvec_i = x_i^T C x_i + sigma_0^2     */
SGP_API int sgp_acosker_diag(sgp_handle* h, const double* x1, long n1, long ld1, int p, const double* C, long ldc,
                     double sigma_0, double* Kvec, void* stream);

/* ---- a5: lambda_moments(x, K_tilde, KKtilde_inv, Kvec, K, C, m, V, theta)  utils.py:1089-1101 -----------------
 * lam_m = a m ; lam_var = Kvec + rowsum(a o (a V^T) - K_b o a).  a, K_b: N x q (lda, ldk).                      */
SGP_API int sgp_lambda_moments(sgp_handle* h, const double* a, long lda, const double* K_b, long ldk, const double* Kvec,
                       long N, int q, const double* m_b, const double* V_b, long ldv, double* lam_m,
                       double* lam_var, void* stream);

/* ---- a6-a8: mean_f_given_lambda_moments / compute_loglikelihood(f-param mode) / lambda0_given_logA -------------
 * One streaming pass over (lam_m, lam_var, r)  (utils.py:1138, 1243, 1253-1255, 1220-1227).
 * out_host[10] = { loglik, dloglik/dlogA, dloglik/dlambda0, lambda0_closed_form, sum r, sum f, sum r*lam_m,
 *                  nan_count, sum (lam_m + A lam_var) f, sum exp(A lam_m + A^2 lam_var/2) };
 * f (N, nullable) receives the mean rate.  An image-sharded caller all-reduces entries 4,5,6,8,9 and recombines:
 * loglik = A*[6] + lambda0*[4] - [5], dlogA = A*([6] - [8]), dlambda0 = [4] - [5], lambda0_closed = log[4] - log[9]. */
SGP_API int sgp_fparam_eval(sgp_handle* h, const double* lam_m, const double* lam_var, const double* r, long N, double logA,
                    double lambda0, double* f, double* out_host, void* stream);

/* ---- a7: compute_loglikelihood with a caller-supplied f_mean  utils.py:1240-1255 / 1269 -----------------------
 * out_host[5] = { sum r*lam_m, sum f, sum (lam_m + A lam_var) f, sum r, nan-count(f) } ; lam_var may be NULL.
 * loglik = A*out[0] + lambda0*out[3] - out[1] ; dloglik/dlogA = A*(out[0] - out[2]) ; dloglik/dlambda0 = out[3]-out[1] */
SGP_API int sgp_loglik_sums(sgp_handle* h, const double* r, const double* f, const double* lam_m, const double* lam_var,
                    long N, double logA, double* out_host, void* stream);

/* ---- a10: Estep(r, KKtilde_inv, m, f_params, f_mean, K_tilde, ...)  utils.py:1420-1422 ------------------------
 * g = A a^T (r - f) (q) ; G = A^2 a^T diag(f) a (q x q, ldg).                                                   */
SGP_API int sgp_estep_terms(sgp_handle* h, const double* a, long lda, const double* r, const double* f, long N, int q,
                    double logA, double* g, double* G, long ldg, void* stream);

/* ---- nd_utility / nd_p_r_given_xD / nd_lambda_r_mean / nd_mean_noise_entropy  utils.py:416-525 --
 * Active-learning utility of nstar candidate images from the mean `mu` and variance `sigma2` (device, length nstar)
 * of log f at each candidate, summed over the counts r[0..nr) (device).  The Lambert-W step that the reference sends
 * through scipy on the CPU (utils.py:464-466) runs on the device.  Outputs are device pointers, each may be null:
 * U[nstar] (utils.py:516), and the (nr x nstar) row-major arrays lam (utils.py:466), mask (z != inf, utils.py:451),
 * p and log p (utils.py:492-494).  Masked entries follow the reference: z, r*sigma2, r and log r! are zeroed and the
 * entry stays in the sums. */
SGP_API int sgp_nd_utility(sgp_handle* h, const double* sigma2, const double* mu, long nstar, const double* r, int nr,
                   double* U, double* lam, unsigned char* mask, double* p, double* logp, void* stream);

/* ---- small dense algebra on the q x q state (north_star item 2) -----------------------------------------------
 * In-place lower Cholesky (strict upper triangle clobbered) ; X = L^{-1} ; flags_host receives SGP_FLAG_KT_* bits. */
SGP_API int sgp_cholesky(sgp_handle* h, double* A, long ld, int n, int* flags_host, void* stream);
SGP_API int sgp_spd_inverse(sgp_handle* h, const double* A, long lda, int n, double* Ainv, long ldi, double* logdet_host,
                    int* flags_host, void* stream);
/* sgp_cholesky_noalloc -- NO counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The in-place lower Cholesky of sgp_cholesky (same kernels, same result) for n <= 1024, with its scratch (the inverses of
 * the 64 x 64 diagonal blocks of L + one status word: 512.25 KiB) in a file-scope `static __device__` array that every
 * GPU loading the library holds until the process exits.  Takes no handle; calls no cudaMalloc / cudaFree / allocating
 * library routine; does not synchronise: all work is enqueued on `stream` of the CURRENT device.
 *   A          device, n x n row-major, ld even and >= n, 16-byte aligned; lower triangle in, L out
 *   flags_dev  device int, nullable: receives the status bits (1 = pivot <= 0, 2 = pivot < 1e-10) in stream order
 *   errbuf     nullable host buffer (errbuf_len bytes) for the error message
 * n > 1024 returns SGP_E_UNSUPPORTED with a message naming the limit.  Every call on a device shares the one array and it
 * keeps the last call's contents (the status word is cleared at the start of each call): calls that OVERLAP on the GPU
 * (two streams / two threads on one device) race -- serialise them.  The first call on a device does the one-time
 * set-up (cudaGetSymbolAddress + kernel attributes): make it outside a graph capture.                              */
SGP_API int sgp_cholesky_noalloc(double* A, long ld, int n, int* flags_dev, void* stream, char* errbuf, int errbuf_len);

/* ---- generic fp64 tensor-core GEMM (exposed for tests and for the roofline measurement) -----------------------
 * C[M x N] = alpha * A * B^T-style contraction with selectable operand layouts:
 *   a_layout / b_layout = 0 : element (row, k) at ptr[row*ld + k] ; = 1 : at ptr[k*ld + row].                  */
SGP_API int sgp_gemm_f64(sgp_handle* h, int a_layout, int b_layout, const double* A, long lda, const double* B, long ldb,
                 double* C, long ldc, int M, int N, int K, double alpha, int accumulate, int small_tile,
                 void* stream);

/* ---- fp64-accurate GEMM on the int8 tensor cores (tcgen05.mma.kind::i8, exposed for tests and measurement) -----
 * layout 0: C[M x N] (ldc) = A[M x K] (lda) * B[N x K]^T (ldb), both operands row-major with K contiguous;
 * layout 1: C = A^T B with A[K x M], B[K x N] (the contraction runs over the rows, e.g. over images).  The operands are
 * split into `slices` (2..8) signed 7-bit digits per element relative to a power-of-two scale per row; the digit
 * products are accumulated exactly in int32 in tensor memory and recombined in fp64 (csrc/ozaki_i8.cu).  With 8 slices
 * the error is <= ~2^-53 * K * max|A_i.| * max|B_j.| per element, i.e. the class of a DGEMM; results are deterministic.
 * Serves the same contractions of acosker (utils.py:969-973) / lambda_moments (utils.py:1090-1101) as sgp_gemm_f64. */
SGP_API int sgp_gemm_i8emu(sgp_handle* h, int layout, const double* A, long M, long lda, const double* B, int N, long ldb,
                   long K, int slices, double* C, long ldc, void* stream);

/* ---- a11: the M-step closure body of varGP (utils.py:2030-2099) as ONE call -----------------------------------
 * theta -> loglik, KL and d(loglik - KL)/dtheta, with X, xtilde, r, B, m_b, V_b, f_params fixed.
 *   X (N x n_px^2, ldx), r (N), xtilde (m x n_px^2, ldu), B (m x q, ldb), m_b (q), V_b (q x q, ldv).
 *   lo_host / hi_host: theta limits (6 each, may be NULL = no check) -> SGP_E_BOUNDS (closure returns +inf, U:2021-2028).
 *   logdetV_host: pointer to a precomputed log det V_b, or NULL to compute it here.
 *   lam_m / lam_var / f: optional device outputs (N each).
 *   out_host[16] = { loglik, KL, g_sigma_0, g_eps_0x, g_eps_0y, g_-2log2beta, g_-log2rho2, g_Amp,
 *                    flags, p, sum r, sum f, sum r*lam_m, logdet K_tilde_b, logdet V_b,
 *                    1 if the moments went through the projected basis else 0 }
 *   with g_k = d(loglik - KL)/dtheta_k (the closure hands -g_k to LBFGS, U:2099).
 * The call synchronises `stream` before returning (the caller needs the loss as a host float).                  */
SGP_API int sgp_elbo_grad(sgp_handle* h, const double* X, long ldx, const double* r, long N, const double* xtilde, long ldu,
                  int m, int n_px, const double* theta_host, const double* lo_host, const double* hi_host,
                  const double* B, long ldb, int q, const double* m_b, const double* V_b, long ldv, double logA,
                  double lambda0, const double* logdetV_host, double* lam_m, double* lam_var, double* f,
                  double* out_host, void* stream);

/* fp32 mode (north_star: "1e-4 in fp32"; the reference itself is fp64 only, utils.py:31-33).  Same evaluation with the
 * images `X` / `xtilde` in FLOAT32 (device, row-major, ld in floats).  The library makes its fp64 working copy (and
 * the int8 slices) of them once per image set; with "i8_slices" on auto the products run on 5 int8 slices (35 bits per
 * operand -- the fp32-class engine; the slice threshold N >= 4096 still applies), the elementwise passes and the
 * quadratic forms stay fp64.  Results agree with the fp64 evaluation of the same (fp32-valued) images to 1e-4 or
 * better (tests/test_gpu_f32.py); everything else (r, B, m_b, V_b, outputs) is fp64 as in sgp_elbo_grad.            */
SGP_API int sgp_elbo_grad_f32(sgp_handle* h, const float* X, long ldx, const double* r, long N, const float* xtilde,
                              long ldu, int m, int n_px, const double* theta, const double* lo, const double* hi,
                              const double* B, long ldb, int q, const double* m_b, const double* V_b, long ldv,
                              double logA, double lambda0, const double* logdetV_host, double* lam_m, double* lam_var,
                              double* f, double* out_host, void* stream);
/* Same evaluation, ASYNCHRONOUS: the 16 result doubles are written to `out_dev` (device memory) and the call returns
 * once everything is enqueued on `stream`, without a host synchronisation.  For callers that post-process the result
 * on the device in stream order: an image-sharded evaluation all-reduces out_dev[0..8] across the GPUs and reads it back
 * once (the one exchange step per evaluation).                                                 */
SGP_API int sgp_elbo_grad_dev(sgp_handle* h, const double* X, long ldx, const double* r, long N, const double* xtilde,
                              long ldu, int m, int n_px, const double* theta, const double* B, long ldb, int q,
                              const double* m_b, const double* V_b, long ldv, double logA, double lambda0,
                              const double* logdetV_host, double* lam_m, double* lam_var, double* f, double* out_dev,
                              void* stream);
/* Population fit (BASELINE.json configs[3]; the reference fits one cell per varGP call, utils.py:1568 / `cellid`
 * utils.py:1636-1643): `ncells` independent cells that share the stimuli X (N x n_px^2, device) and the inducing images
 * xtilde.  Cell c has its own theta (theta + 6c), spike counts r[c] (N), basis B[c] (m x q[c], ld ldb[c]), variational
 * state m_b[c] / V_b[c] (ld ldv[c]), link parameters logA[c] / lambda0[c] and optionally logdetV[c] (nullable array).
 * r, B, m_b, V_b are HOST arrays of DEVICE pointers.  One pass of the fused closure per cell, back to back on `stream`
 * with no host synchronisation in between; every cell runs on the same resident images and the same int8 slices of
 * them (cells whose receptive-field mask crops the grid get C = 0 outside it).  out_host: ncells x 16 doubles, the
 * layout of sgp_elbo_grad.  No collective: cells are independent (last row).                   */
SGP_API int sgp_elbo_grad_cells(sgp_handle* h, const double* X, long ldx, long N, const double* xtilde, long ldu, int m,
                                int n_px, int ncells, const double* const* r, const double* theta,
                                const double* const* B, const long* ldb, const int* q, const double* const* m_b,
                                const double* const* V_b, const long* ldv, const double* logA, const double* lambda0,
                                const double* logdetV, double* out_host, void* stream);
/* ---- image-sharded evaluation: exchange over peer memory (north_star item 4) ---------------
 * One process per GPU; rank k holds a block of the images and calls sgp_elbo_grad / sgp_elbo_grad_dev on it.  After
 *   sgp_comm_export(h, world, handle64)            -- allocate this rank's exchange buffer, get its CUDA IPC handle
 *   <all-gather the 64-byte handles over any host channel>
 *   sgp_comm_import(h, rank, world, handles)       -- map every peer's buffer (NVLink / NVSwitch peer access)
 * the result-assembly kernel of every evaluation ALSO does the exchange step: each rank stores its 16 doubles into a
 * slot of every peer's buffer, publishes an epoch flag (release), waits for all ranks' flags (acquire, 3 s limit ->
 * status bit 32), and sums loglik, the 6 gradients and the image sums in rank order.  Every rank returns identical,
 * already reduced numbers from ONE kernel: no collective launch, no extra host round trip.  KL is replicated and
 * returned once; the KL part of the gradient must be counted once (option "kl_weight" = 0 on ranks > 0).  All ranks must
 * issue the same sequence of evaluations.  sgp_comm_close ends the mode.  Cells / host-streaming calls never exchange. */
SGP_API int sgp_comm_export(sgp_handle* h, int world, unsigned char* handle_out /* 64 bytes */);
SGP_API int sgp_comm_import(sgp_handle* h, int rank, int world, const unsigned char* handles /* world x 64 bytes */);
SGP_API int sgp_comm_close(sgp_handle* h);
/* Resident-image protocol for HOST callers.  `x`, `r` and `xtilde` are constant during a fit (varGP never modifies
 * them, utils.py:1568-2316), only theta and the projected variational state change between closure evaluations.
 * sgp_upload_images copies the three to device buffers owned by the handle ONCE (the int8 slices of the images are
 * made by the first evaluation and kept); sgp_elbo_grad_resident then sends theta (by value), B (m x q), m_b (q) and
 * V_b (q x q) from host memory per evaluation -- a few MB instead of the N x n_px^2 images -- runs the fused closure
 * and returns the 16 doubles of sgp_elbo_grad in out_host.  lo / hi: nullable theta limits (SGP_E_BOUNDS outside).   */
SGP_API int sgp_upload_images(sgp_handle* h, const double* X_host, long ldx, const double* r_host, long N,
                              const double* xtilde_host, long ldu, int m, int n_px);
SGP_API int sgp_elbo_grad_resident(sgp_handle* h, const double* theta, const double* lo, const double* hi,
                                   const double* B_host, long ldb, int q, const double* m_b_host,
                                   const double* V_b_host, long ldv, double logA, double lambda0, double* out_host);
/* Same evaluation with X / r in HOST memory (pinned or pageable): the images are streamed to the device slab by
 * slab on a copy stream that overlaps the compute stream.  Used for the end-to-end measurement.                  */
SGP_API int sgp_elbo_grad_host(sgp_handle* h, const double* X_host, long ldx, const double* r_host, long N,
                       const double* xtilde_host, long ldu, int m, int n_px, const double* theta_host,
                       const double* B_host, long ldb, int q, const double* m_b_host, const double* V_b_host,
                       long ldv, double logA, double lambda0, double* out_host);

/* ---- debugging aid for parity tests: copy a named intermediate of the last sgp_elbo_grad call ---------------
 * names: C Kt Ktb Kinv Mfull GKt GCU Sig2 bb y beta X2 ...   dst: device, capacity in elements.                */
SGP_API int sgp_debug_get(sgp_handle* h, const char* name, double* dst, long capacity, long* rows, long* cols, long* ld);

#ifdef __cplusplus
}
#endif
#endif /* SGP_B200_H_ */
