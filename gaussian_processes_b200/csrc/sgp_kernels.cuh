// Host-callable wrappers of the non-GEMM kernels of the Spatial_GP pass (definitions: sgp_kernels.cu).
// All pointers are device pointers unless stated otherwise; every launch goes to `st`.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace sgp {

// status bits reported in out[8] of the fused closure (and by the factorisation helpers)
enum StatusBits : int {
  kFlagKtNotPD = 1,        // Cholesky of K_tilde_b hit a pivot <= 0 (reference: log_det falls back to eigh, U:1279)
  kFlagKtSmallPivot = 2,   // pivot < 1e-10 (reference safe_log raises, U:670)
  kFlagVNotPD = 4,         // same for V_b
  kFlagVSmallPivot = 8,
  kFlagNaN = 16,           // NaN in f (reference raises in the f-param closure, U:1923)
  kFlagCommTimeout = 32,   // image-sharded exchange: a peer's partial result did not arrive within the time limit
};

// slots of the small device scalar buffer the closure accumulates into
enum ScalSlot : int {
  kSlotLogdetK = 0, kSlotLogdetV = 1, kSlotMb = 2, kSlotTrVK = 3,
  kSlotRLam = 4, kSlotSumF = 5, kSlotSumR = 6, kSlotNaN = 7,
  kSlotSig = 8,            // sum R + sum w + sum beta + sum R~ + sum w~   (sigma_0 gradient / (2 sigma_0))
  kSlotG = 9,              // 5 slots: <dC_k, GC> for Amp, -2log2beta, -log2rho2, eps_0x, eps_0y
  kSlotAux0 = 14, kSlotAux1 = 15, kSlotAux2 = 16, kSlotAux3 = 17,
  kSlotGW = 22,            // 5 slots: sum_i w_i x_i^T dC_k x_i (separable path), added to kSlotG at the end
  kNumSlots = 32
};

// every kernel launch of the library is counted (bench.py reports it as gpu_launches)
void note_launch(int n = 1);
long launch_counter();
// multiProcessorCount of the current device (queried once per device; 148 on B200)
int num_sms();

struct ThetaDev {  // hyper-parameters in the form the kernels need
  double s0, s02, ex, ey, eb, er, amp;
};

// ---- a1: localker --------------------------------------------------------------------------
// C (p x p, ld = ldc, padding zeroed) from masked pixel coordinates gx/gy (device, length p).
// dC (optional): 5 matrices, stride dc_stride, order Amp, -2log2beta, -log2rho2, eps_0x, eps_0y (U:910).
void build_C(double* C, long ldc, int p, const double* gx, const double* gy, const ThetaDev& th, double* dC,
             long dc_stride, cudaStream_t st);

// out[i][c] = X[i][idx[c]] (c < p), 0 for p <= c < ldo
void gather_cols(const double* X, long ldx, long nrows, const int* idx, int p, double* out, long ldo,
                 cudaStream_t st);
void gather_cols_f32(const float* X, long ldx, long nrows, const int* idx, int p, double* out, long ldo,
                     cudaStream_t st);     // same, from float32 images

// s[i] = s02 + sum_t part[t*ldpart + i] ; nrm[i] = sqrt(s[i])
void norms_finalize(const double* part, long ldpart, int nparts, long n, double s02, double* s, double* nrm,
                    cudaStream_t st);

// ---- a5-a7: per-image moments, rate, adjoint weights ---------------------------------------
// lam_var = s + 0.5*sum_{t<nparts} part1 ; lam_m = sum_{t<nparts2} part2 ; f = exp(A lam_m + A^2 lam_var/2 + lambda0)
// u = A (r - f) ; v = -A^2 f / 2 ; blockpart[blk*4 + {0,1,2,3}] = sum r*lam_m, sum f, sum r, nan-count
int moments_rows_blocks(long n);
void moments_rows(const double* s, const double* part1, const double* part2, long ldpart, int nparts, int nparts2,
                  const double* r, long n, double A, double lambda0, double* lam_m, double* lam_var, double* f,
                  double* u, double* v, double* blockpart, cudaStream_t st);

// ---- arc-cosine map, in place: K holds x1^T C x2 on entry (n x m, ld even), K = X1X2*J and cos on exit ----
void acos_forward(double* K, double* cosm, long ld, long n, int m, const double* na, const double* nb, double s02,
                  cudaStream_t st);

// ---- adjoint of the arc-cosine nonlinearity (N x m elementwise + row/column reductions) ----
// mode 0: GK_ij = u_i*bb_j + v_i*G_ij ; mode 1: GK_ij = G_ij.
// G is overwritten with R = GK*P.  rowout_i = (mode0 ? v_i : 0) + 0.5*rho_i/na_i,
// colpart_k[blk][j] = sum_{i in blk} GK_ij Q_ij na_i ; colpart_y[blk][j] = sum_i Kmat_ij u_i (mode 0)
// sumpart[blk] = sum R.   Requires m <= kBwdMaxCols.
constexpr int kBwdRows = 32;
constexpr int kBwdMaxCols = 2048;
int acos_backward_blocks(long n);
void acos_backward(int mode, double* G, const double* cosm, const double* Kmat, long ld, long n, int m,
                   const double* u, const double* v, const double* bb, const double* na, const double* nb,
                   double* rowout, double* colpart_k, double* colpart_y, long ldcol, double* sumpart,
                   cudaStream_t st, double* colmax_r = nullptr, double* colmax_w = nullptr);
// colmax_r / colmax_w (mode 0, optional; [blk][ldcol] like colpart_k): per-block column maxima of |R_ij| and of
// -v_i K_ij^2 -- the column scales of the two transposed int8 slicings that follow (oz_exps_from_colmax), so that
// neither needs its own pass over the slab.

// out[j] (+)= sum_b part[b*ld + j]   (fixed order -> deterministic)
void col_reduce(const double* part, int nblk, int m, long ld, double* out, int accumulate, cudaStream_t st);
// out[i] (+)= sum_z part[z*zstride + i], i < n
void slice_reduce(const double* part, int nz, long zstride, long n, double* out, int accumulate, cudaStream_t st);
// scal[slot] (+)= scale * sum_i x[i*stride], single CTA, fixed order
void sum_to_slot(const double* x, long n, long stride, double scale, double* scal, int slot, int accumulate,
                 cudaStream_t st);
void sum_to_slots(const double* x, long n, long stride, int nslots, double* scal, int slot0, int accumulate,
                  cudaStream_t st);   // nslots interleaved columns, slots slot0..

// ---- small dense helpers (q x q / m x m) -----------------------------------------------------
void symmetrize(const double* A, long lda, int n, double* out, long ldo, cudaStream_t st);   // out = (A + A^T)/2
void add_transposed(double* A, long lda, int rows, int cols, const double* B, long ldb, cudaStream_t st);   // A += B^T
void mirror_lower(double* A, long ld, int n, cudaStream_t st);                                // A[i][j] = A[j][i], j > i
void lincomb(double* out, long ldo, int rows, int cols, double a, const double* A, long lda, double b,
             const double* B, long ldb, double diag, cudaStream_t st);                        // out = a*A + b*B + diag*I
void rank1_add(double* out, long ldo, int rows, int cols, double alpha, const double* x, const double* y,
               cudaStream_t st);                                                             // out += alpha x y^T
void gemv(const double* A, long lda, int rows, int cols, int trans, const double* x, double* y, cudaStream_t st);
void copy_mat(const double* A, long lda, int rows, int cols, double* out, long ldo, cudaStream_t st);
// G = G + klw*(-0.5*Kinv + 0.5*KVK^T + 0.5*b b^T)     (adjoint of K_tilde_b: loglik part already in G)
void gkt_assemble(double* G, long ld, int q, const double* Kinv, const double* KVK, const double* b, double klw,
                  cudaStream_t st);
// out = 0.5*(kap_a + kap_b)/nrm + extra   (null pointers count as 0)
void col_weights(const double* kap_a, const double* kap_b, const double* extra, const double* nrm, int m,
                 double* out, cudaStream_t st);

// ---- Cholesky / triangular inverse (blocked, panel 64) --------------------------------------
// In-place lower Cholesky of A (n x n, lower triangle referenced; the strict upper triangle of the diagonal
// blocks is zeroed). flags |= notpd/small bits.  Winv: workspace of chol_winv_elems(n) doubles, receives the
// inverses of the 64 x 64 diagonal blocks of L.  Panel solve and trailing update run on the DMMA GEMM.
inline long chol_winv_elems(int n) { return (long)((n + 63) / 64) * 64 * 64; }
cudaError_t chol_init();     // once per device (kernel attribute), before the first cholesky_lower
cudaError_t cholesky_lower(double* A, long ld, int n, int* flags, int bit_notpd, int bit_small, double* Winv,
                           cudaStream_t st);
// One-launch version of cholesky_lower + tri_inverse_lower + logdet_from_chol for n <= 1024 (chol_fused.cu: a cluster of
// 8 CTAs, cluster barriers instead of kernel boundaries).  X (optional) = L^-1; slot < 0: no log det.
bool chol_inv_fused_supported(int n);
cudaError_t chol_inv_fused(double* A, long ld, int n, double* X, long ldx, double* Winv, int* flags, int bit_notpd,
                           int bit_small, double* scal, int slot, cudaStream_t st);
// X = L^{-1} (lower, strict upper zero). X must not alias L. tmp: scratch >= 64*ld doubles; Winv from cholesky_lower.
cudaError_t tri_inverse_lower(const double* L, long ld, int n, double* X, double* tmp, const double* Winv,
                              cudaStream_t st);
// scal[slot] = 2*sum log L_kk
void logdet_from_chol(const double* L, long ld, int n, double* scal, int slot, cudaStream_t st);
// scal[slot_mb] = m.b ; scal[slot_tr] = sum_ij V_ij Kinv_ij (Kinv symmetric) ; scratch: 2*q doubles
void kl_scalars(const double* m_b, const double* b, const double* V, long ldv, const double* Kinv, long ldk, int q,
                double* scal, int slot_mb, int slot_tr, double* scratch, cudaStream_t st);

// ---- ntilde == nt closure (a = B): elementwise pieces of the adjoint, m x q each (definitions: sgp_kernels.cu) ----
void same_ga(double* Ga, long ld, int rows, int cols, const double* u, const double* v, const double* m_b,
             const double* aV, const double* Kb, cudaStream_t st);                   // u m^T + v o (2 aV - Kb)
void same_combine(double* E, long ld, int rows, int cols, const double* H, const double* v, const double* a,
                  const double* BG, cudaStream_t st);                                // H - diag(v) a + BG
void scale_rows(double* out, const double* in, long ld, int rows, const double* w, cudaStream_t st);   // diag(w) in

// ---- final contraction with dC_k and result assembly -----------------------------------------
// GC = sum_z GCpart[z] + GCextra ; gpart[blk*5 + k] = sum_ab GC_ab * dC_k,ab  (k: Amp, beta, rho, ex, ey)
int contract_blocks(int p);
void contract_dC(const double* GCpart, int nz, long zstride, const double* GCextra, const double* C, long ld, int p,
                 const double* gx, const double* gy, const ThetaDev& th, double* gpart, cudaStream_t st);

// out[0] loglik, [1] KL, [2..7] d(loglik-KL)/d(sigma_0, eps_0x, eps_0y, -2log2beta, -log2rho2, Amp),
// [8] flags, [9] p, [10] sum r, [11] sum f, [12] sum r*lam_m, [13] logdet K_tilde_b, [14] logdet V_b, [15] out15 (1 = moments in the projected basis)
void finalize(const double* scal, const int* flags, double A, double lambda0, const ThetaDev& th, int p,
              double logdetV_override, int use_override, double* out, double out15, cudaStream_t st);

// ---- image-sharded evaluation: the exchange step fused into the result assembly ----------------------------------
// Every rank's finalize kernel writes its 16 result doubles straight into a slot of EVERY peer's exchange buffer
// (peer-to-peer stores over NVLink / NVSwitch; the buffers are cudaMalloc'ed and opened through CUDA IPC), publishes a
// per-rank epoch flag with release semantics, waits for the flags of all ranks (acquire loads, bounded by a time limit),
// and sums the slots in rank order -- so every rank ends with bit-identical reduced numbers, without a collective
// launch, a second kernel or a host round trip.  Slots / flags are double-buffered by epoch parity.
constexpr int kCommMaxRanks = 16;
struct CommDev {
  int rank = 0, world = 1;
  unsigned long long epoch = 0;                       // same on every rank: evaluations are issued in lock step
  double* slots[kCommMaxRanks] = {nullptr};           // base of rank i's buffer: [2][world][16] doubles
  unsigned long long* flags[kCommMaxRanks] = {nullptr};   // rank i's flags: [2][world]
};
inline size_t comm_buffer_bytes(int world) { return (size_t)2 * world * 16 * sizeof(double) + (size_t)2 * world * 8 + 256; }
void finalize_exchange(const double* scal, const int* flags, double A, double lambda0, const ThetaDev& th, int p,
                       double logdetV_override, int use_override, double* out, double out15, const CommDev& comm,
                       cudaStream_t st);


// ---- E-step / f-param helpers -------------------------------------------------------------------
// lam_m = sum_t part2 ; lam_var = Kvec + sum_t part1 - rowdot(K_b, a)        (utils.py:1090-1101)
void moments_combine(const double* Kvec, const double* part1, const double* part2, long ldpart, int nparts,
                     int nparts2, const double* a, long lda, const double* K_b, long ldk, long n, int q, double* lam_m,
                     double* lam_var, cudaStream_t st);
// one streaming pass over (lam_m, lam_var, r): out[0..6) =
//   sum r*lam_m, sum f, sum (lam_m + A lam_var) f, sum r, sum exp(A lam_m + A^2 lam_var/2), nan-count.
// blockpart: fparam_blocks(n) * 8 doubles of scratch; ticket: one zero-initialised counter (re-armed by the kernel)
int fparam_blocks(long n);
void fparam_pass(const double* lam_m, const double* lam_var, const double* r, long n, double A, double lambda0,
                 double* f, double* blockpart, unsigned* ticket, double* out, cudaStream_t st);
// same, for a caller-supplied f: out[0..5) = {sum r*lam_m, sum f, sum (lam_m + A lam_var) f, sum r, nan}
void loglik_sums_pass(const double* r, const double* f, const double* lam_m, const double* lam_var, long n, double A,
                      double* blockpart, unsigned* ticket, double* out, cudaStream_t st);
// colpart[blk][j] = sum_{i in blk} a_ij (r_i - f_i)
int colsum_blocks(long n);
void colsum_weighted(const double* a, long lda, long n, int q, const double* r, const double* f, double* colpart,
                     long ldcol, cudaStream_t st);
void scale_vec(double* x, long n, double alpha, cudaStream_t st);

// ---- separable evaluation of the per-image quadratic forms (sep_quad.cu) ------------------------------
// s_i = x_i^T C x_i + sigma_0^2, nrm_i = sqrt(s_i) and Q[i][5] = x_i^T dC_k x_i in the units of the closure's
// gradient slots (Amp, -2log2beta, -log2rho2, eps_0x, eps_0y), from the Kronecker structure of localker's C.
// X: gathered images (nrows x p, ld = ldx); off[c] = iy*sep_quad_ld(n) + ix of masked pixel c; alpha[c] = RF envelope.
constexpr int kSepMaxTiles = 6;   // n_px_side <= 48
struct SepQuadArgs {
  const double* X; long ldx; long nrows; int p;
  const int* off; const double* alpha;
  int n;
  double er, eb, ex, ey, amp, s02;
  double* s; double* nrm; double* Q;   // Q may be null (value only)
  int max_ctas = 0;                    // 0 = one CTA per SM; smaller grids leave SMs to a concurrent stream
  int full = 0;                        // 1: the mask keeps every pixel and off[] is the plain n x n layout (fast scatter)
};
bool sep_quad_supported(int n_px);
int sep_quad_ld(int n_px);
cudaError_t sep_quad(const SepQuadArgs& a, cudaStream_t st);
// ZU (m x ldz, padding zeroed) = U C, t_out[j] = u_j^T C u_j + s02, nrm_out[j] = sqrt(t_out[j]) from the separable structure
// of C (off / alpha as in SepQuadArgs); U: the gathered inducing images (m x p, ld = ldu)
cudaError_t sep_apply_C(const double* U, long ldu, int m, int p, const int* off, const double* alpha, int n, double er,
                        double amp, double s02, double* ZU, long ldz, double* t_out, double* nrm_out, cudaStream_t st);
int wq_reduce_blocks(long n);
void wq_reduce(const double* w, const double* Q, long n, double* blockpart, cudaStream_t st);

// ---- fp64-accurate GEMM on tcgen05.mma.kind::i8 (int8 slices, exact int32 accumulation in TMEM): ozaki_i8.cu ----
int oz_ldk(int K);                                   // row pitch (bytes) of one slice PAIR of an operand with K columns
size_t oz_slice_bytes(long rows, int K, int S);      // bytes of the digit array of a rows x K operand
cudaError_t oz_init();                               // opt-in shared memory of the kernel (once per process)
// q: [ceil(S/2)][rows][ldk] int8 digits (slices 2j, 2j+1 interleaved per 64-byte k-block), exps[rows]: row scales
// colvec / rowdot (optional): rowdot[r] = sum_c X[r][c] * colvec[c], taken while the row streams through
cudaError_t oz_slice_rows(const double* X, long ldx, long rows, int cols, int S, int8_t* q, int ldk, int* exps,
                          cudaStream_t st, const double* colvec = nullptr, double* rowdot = nullptr);
// C (M x N) = A * B^T from the slices of A (M rows) and B (N rows), K columns each
// slices of X^T (contraction over the rows of X): q is [ceil(S/2)][cols][oz_ldk(rows)], exps[cols]; mx: cols 8-byte words
// rowsq (optional, length rows): row r is multiplied by sqrt(|rowsq[r]|) before slicing (the sign goes into alpha)
cudaError_t oz_slice_cols(const double* X, long ldx, long rows, int cols, int S, int8_t* q, int ldk, int* exps,
                          unsigned long long* mx, cudaStream_t st, const double* rowsq = nullptr,
                          bool exps_ready = false);
// exps[j] from per-block column maxima part[b*ld + j] (squared: the maxima are of the squared entries)
void oz_exps_from_colmax(const double* part, int nblk, int cols, long ld, int squared, int* exps, cudaStream_t st);
// nz > 1: split-K, slice z of the k-blocks goes to C + z * zstride (oz_split picks nz; the effective count, which
// oz_gemm never exceeds, is ceil(nkb / ceil(nkb / nz))); accumulate: C += instead of C =
int oz_split(long M, int N, int K, int S, int lower_only = 0);
double* oz_prof_ptr();   // 8 doubles written by CTA 0 of the last oz_gemm launch when SGP_OZ_DEBUG & 64 (cycle counts)
inline int oz_eff_split(int K, int nz) { const int nkb = (K + 63) / 64, kbz = (nkb + nz - 1) / nz; return (nkb + kbz - 1) / kbz; }
// optional moments epilogue: per 16-column group g, part1[g*ldpart + row] = sum_col C*aux, part2 = sum_col aux*colvec
// row-sum epilogue: part1[g * ldpart + row] = sum over the 16-column group g of C[row][.] * aux[row][.]
struct OzMoments { const double* aux; long ldaux; double* part1; long ldpart; };
cudaError_t oz_gemm(const int8_t* qa, const int* ea, long M, const int8_t* qb, const int* eb, int N, int K, int ldk,
                    int S, double* C, long ldc, int nz, long zstride, int accumulate, cudaStream_t st,
                    const OzMoments* mom = nullptr, int max_ctas = 0, double alpha = 1.0, int lower_only = 0);

// ---- active-learning utility (utils.py:416-525): U[nstar] and optional (r, nstar) arrays lam / mask / p / log p ----
void nd_utility(const double* sigma2, const double* mu, const double* r, int nr, long nstar, double* U, double* lam,
                unsigned char* mask, double* p, double* logp, cudaStream_t st);

}  // namespace sgp
