// C ABI of the Spatial_GP hot path (include/sgp_b200.h): handle, workspace, and the orchestration of the
// fused M-step closure.  Host C++ only launches kernels; there is no CPU compute path.
#include "../../include/sgp_b200.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <initializer_list>
#include <vector>

#include "gemm_f64.cuh"
#include "sgp_kernels.cuh"

using namespace sgp;

namespace {

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  cudaError_t reserve(size_t need) {
    if (need <= bytes) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; bytes = 0;
    cudaError_t e = cudaMalloc(&p, need);
    if (e == cudaSuccess) bytes = need;
    if (e == cudaSuccess && poison_new()) e = cudaMemset(p, 0xFF, need);   // SGP_POISON: fresh memory is not zero either
    return e;
  }
  static bool poison_new() { static const bool on = getenv("SGP_POISON") && atoi(getenv("SGP_POISON")) != 0; return on; }
  void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
};

// SGP_POISON=1 (testing only): every entry point fills its work area with 0xFF bytes (NaN as a double, -1 as a digit or
// an exponent) before using it, so a kernel that reads scratch it did not write in THIS call -- stale values from an
// earlier call, unwritten padding multiplied by zero -- shows up as NaN / wrong results in the parity tests.
const int g_poison = getenv("SGP_POISON") ? atoi(getenv("SGP_POISON")) : 0;
#define RESERVE_ARENA(need)                                                                   \
  do {                                                                                        \
    CU(h->arena.reserve(need));                                                               \
    if (g_poison) {   /* synchronised: uploads on the copy stream must not race with the fill */ \
      CU(cudaMemsetAsync(h->arena.p, 0xFF, h->arena.bytes, st));                              \
      CU(cudaStreamSynchronize(st));                                                          \
    }                                                                                         \
  } while (0)

struct DebugEntry { const double* ptr; long rows, cols, ld; };

inline long rup(long x, long a) { return (x + a - 1) / a * a; }

}  // namespace

struct sgp_handle {
  int device = 0;
  std::string err;
  long slab_rows = 65536;
  long stream_slab_rows = 8192;    // host-streaming variant: smaller slabs so H2D copies overlap the compute earlier
  long launches_base = 0;
  bool gemm_ready = false;
  // scratch arena (re-carved every call) + persistent caches
  DevBuf arena, xg, ug, gcpart, s2part, dpart;
  int path = 0;             // 0 = choose per (n, p); 1 = dense C products; 2 = separable quadratic forms
  // the two N x p x m products on tcgen05.mma.kind::i8 (ozaki_i8.cu): -1 = when the slab is large, 0 = never
  // (fp64 DMMA), 2..8 = always, with that many int8 slices
  int i8_slices = -1;
  // slice count of the two products with the images (K_lin = X (UC)^T, D = X^T R): -1 = same as i8_slices
  int i8_slices_images = -1;
  // slice counts of the two products that feed ONLY the gradient: D = X^T R (<= the images' count: it reads the leading
  // slices of the persistent X^T digits) and Sigma2 = K^T diag(v) K (<= i8_slices); -1 = same as the parent option
  int i8_slices_xtr = -1, i8_slices_sig2 = -1;
  DevBuf xq, xe, xtq, xte;  // int8 slices of the images (row-wise and transposed per slab) + their exponents
  bool oz_valid = false; int oz_S = 0; long oz_slab = 0;
  // cache key of the gathered copies
  const double* cache_X = nullptr; long cache_N = 0, cache_ldx = 0;
  const double* cache_U = nullptr; long cache_m = 0, cache_ldu = 0;
  std::vector<uint8_t> cache_mask;
  bool cache_valid = false;
  // content identity of the image buffers (sgp_bind_images): part of the key of the gathered copies and of the
  // persistent int8 slices, because an address alone does not identify its content (allocators recycle blocks)
  unsigned long long image_token = 0, cache_token = 0;
  int* d_flags = nullptr;
  double* d_scal = nullptr;
  unsigned* d_ticket = nullptr;        // last-block-done counter of the streaming reductions (zero between launches)
  double* d_out = nullptr;
  double* h_out = nullptr;  // pinned
  // pinned staging ring for the small per-evaluation uploads (pixel lists of the support): uploads from PAGEABLE memory
  // synchronise the stream first, which would serialise back-to-back evaluations (cells, asynchronous calls)
  static constexpr int kPinSlots = 4;
  char* pin_base = nullptr; size_t pin_slot_bytes = 0; int pin_next = 0;
  cudaEvent_t pin_ev[kPinSlots] = {nullptr, nullptr, nullptr, nullptr}; bool pin_used[kPinSlots] = {false, false, false, false};
  DevBuf cells_out;
  // image-sharded exchange over peer memory (sgp_comm_*): my buffer, the peers' mappings, the lock-step epoch
  DevBuf comm_buf; CommDev comm; bool comm_on = false; void* comm_peer[kCommMaxRanks] = {nullptr};
  // resident-image protocol (sgp_upload_images / sgp_elbo_grad_resident): the fit's constant operands live here
  DevBuf res_x, res_r, res_u; long res_N = 0, res_m = 0; int res_npx = 0;
  std::map<std::string, DebugEntry> dbg;
  // second stream + events for the host-streaming variant
  cudaStream_t copy_stream = nullptr, compute_stream = nullptr, side_stream = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  cudaStream_t chain_stream = nullptr; cudaEvent_t ev_chain_fork = nullptr, ev_chain_join = nullptr;
  int overlap = 1;          // 1 = run the q x q inducing-point chain next to the first slab's quadratic forms
  int support = -1;         // pixel set of the cached image copies: -1 = mask grown by 2 px, reused while the mask stays inside; 0 = the mask itself; 1 = whole grid
  // N == m: 1 = the reference's ntilde == nt branch (K = K_tilde, a = B); 0 = treat it as any other N (an image SHARD
  // that happens to hold m rows is not the active-learning configuration)
  int square_is_same = 1;
  int chol_fused = getenv("SGP_CHOL_FUSED") ? atoi(getenv("SGP_CHOL_FUSED")) : 320;   // one-launch Cholesky / inverse up to this order
  int sep_uc = getenv("SGP_SEP_UC") ? atoi(getenv("SGP_SEP_UC")) : 1;   // U C from the separable structure of C (one launch)
  int projected = -1;       // moments in the projected basis (K_b = K B): -1 = when q is small enough to pay, 0 = never, 1 = always
  DevBuf hx[2], hr, small;
  cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_used[2] = {nullptr, nullptr}, ev_small = nullptr;
  // options
  double kl_weight = 1.0;   // weight of the KL adjoint in the gradient (image-sharded ranks > 0 use 0)
  bool profile = false;     // record CUDA events at stage boundaries of the closure
  std::vector<cudaEvent_t> prof_ev; std::vector<int> prof_stage; size_t prof_used = 0;
  double stage_ms[16] = {0}; long stage_calls = 0;
};

namespace {

int fail(sgp_handle* h, int code, const std::string& msg) {
  if (h) h->err = msg;
  return code;
}
#define CU(call)                                                                          \
  do {                                                                                    \
    cudaError_t e__ = (call);                                                             \
    if (e__ != cudaSuccess)                                                               \
      return fail(h, SGP_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__));    \
  } while (0)

// torch.linspace(-1, 1, n) in fp64 (ATen: start + step*i below the midpoint, end - step*(n-1-i) above)
inline double lin_coord(int i, int n) {
  if (n == 1) return -1.0;
  const double step = 2.0 / (double)(n - 1);
  return (i < n / 2) ? (-1.0 + step * i) : (1.0 - step * (n - 1 - i));
}

ThetaDev make_theta(const double* t) {
  ThetaDev th;
  th.s0 = t[0]; th.s02 = t[0] * t[0];
  th.ex = t[1]; th.ey = t[2];
  th.eb = std::exp(t[3]); th.er = std::exp(t[4]);
  th.amp = t[5];
  return th;
}

// RF mask of localker (U:880-883), pixel k = iy*n + ix has x = lin[ix], y = lin[iy] (U:876-879)
int compute_mask(const double* theta, int n, std::vector<uint8_t>& mask, std::vector<int>& idx,
                 std::vector<double>& gx, std::vector<double>& gy) {
  const ThetaDev th = make_theta(theta);
  mask.assign((size_t)n * n, 0);
  idx.clear(); gx.clear(); gy.clear();
  for (int iy = 0; iy < n; ++iy)
    for (int ix = 0; ix < n; ++ix) {
      const double x = lin_coord(ix, n), y = lin_coord(iy, n);
      const double la = -th.eb * ((x - th.ex) * (x - th.ex) + (y - th.ey) * (y - th.ey));
      if (std::exp(la) >= kMaskThr) {
        mask[(size_t)iy * n + ix] = 1;
        idx.push_back(iy * n + ix);
        gx.push_back(x); gy.push_back(y);
      }
    }
  return (int)idx.size();
}

struct Arena {
  char* base; size_t off = 0, cap;
  Arena(void* b, size_t c) : base((char*)b), cap(c) {}
  template <class T> T* take(size_t n) {
    off = (off + 255) & ~size_t(255);
    T* p = (T*)(base + off);
    off += n * sizeof(T);
    return p;
  }
};

cudaError_t gemm(cudaStream_t st, int al, int bl, int mode, bool small, int nsplit, GemmArgs g) {
  if (g.k_chunk == 0) g.k_chunk = (int)rup(g.K, 16);
  if (g.alpha == 0.0) g.alpha = 1.0;
  return gemm_f64_launch(g, GemmLaunch{al, bl, mode, small, nsplit}, st);
}
// plain C = alpha * op(A) op(B) helper
cudaError_t mm(cudaStream_t st, int al, int bl, const double* A, long lda, const double* B, long ldb, double* C,
               long ldc, int M, int N, int K, double alpha = 1.0, bool small = true) {
  GemmArgs g{};
  g.A = A; g.lda = lda; g.B = B; g.ldb = ldb; g.C = C; g.ldc = ldc; g.M = M; g.N = N; g.K = K; g.alpha = alpha;
  return gemm(st, al, bl, kStore, small, 1, g);
}

// Cholesky factor (in place, lower), optionally L^-1 (X) and log det (scal[slot], slot >= 0): ONE launch of the cluster
// kernel (chol_fused.cu) up to option "chol_fused" rows (default 320: beyond that the trailing updates want more than the
// cluster's 8 SMs and the launch-per-step path is faster), else cholesky_lower / logdet_from_chol / tri_inverse_lower.
cudaError_t factor_any(sgp_handle* h, double* L, long ld, int n, double* X, double* tri_tmp, double* winv, int bit_notpd,
                       int bit_small, int slot, cudaStream_t st) {
  if (n <= h->chol_fused && chol_inv_fused_supported(n))
    return chol_inv_fused(L, ld, n, X, ld, winv, h->d_flags, bit_notpd, bit_small, h->d_scal, slot, st);
  cudaError_t e = cholesky_lower(L, ld, n, h->d_flags, bit_notpd, bit_small, winv, st);
  if (e != cudaSuccess) return e;
  if (slot >= 0) logdet_from_chol(L, ld, n, h->d_scal, slot, st);
  return X ? tri_inverse_lower(L, ld, n, X, tri_tmp, winv, st) : cudaSuccess;
}

int ensure_ready(sgp_handle* h) {
  CU(cudaSetDevice(h->device));
  if (!h->gemm_ready) {
    CU(gemm_f64_init());
    CU(chol_init());
    CU(oz_init());
    CU(cudaMalloc(&h->d_flags, sizeof(int) * 4));
    CU(cudaMalloc(&h->d_scal, sizeof(double) * kNumSlots));
    CU(cudaMalloc(&h->d_ticket, 64));
    CU(cudaMemset(h->d_ticket, 0, 64));
    CU(cudaMalloc(&h->d_out, sizeof(double) * 16));
    CU(cudaMallocHost(&h->h_out, sizeof(double) * 16));
    h->gemm_ready = true;
  }
  return SGP_OK;
}

// CTA tiles of a large (64 x 128) product; `lower` counts only the tiles a symmetric product computes
int gemm_tiles(int M, int N, bool lower = false) {
  const int tm = (M + 63) / 64, tn = (N + 127) / 128;
  if (!lower) return tm * tn;
  int c = 0;
  for (int a = 0; a < tm; ++a)
    for (int b = 0; b < tn; ++b)
      if (b * 128 <= a * 64 + 63) ++c;
  return c;
}

// pick a split-K factor so tiles*nz fills the 148 SMs x 2 resident CTAs in (nearly) whole waves
int pick_split(int tiles, long K) {
  int best = 1; double best_score = -1.0;
  for (int nz = 1; nz <= 64; ++nz) {
    if (nz > 1 && K / nz < 256) break;
    const double ctas = (double)tiles * nz;
    const double waves = ctas / (2.0 * num_sms());
    const double eff = waves / std::ceil(waves);
    const double score = eff - 0.002 * nz;
    if (score > best_score) { best_score = score; best = nz; }
  }
  return best;
}

}  // namespace

extern "C" {

int sgp_version(void) { return 100; }

int sgp_create(int device, sgp_handle** out) {
  if (!out) return SGP_E_ARG;
  sgp_handle* h = new sgp_handle();
  h->device = device;
  h->launches_base = launch_counter();
  *out = h;
  return SGP_OK;
}

int sgp_destroy(sgp_handle* h) {
  if (!h) return SGP_OK;
  if (h->gemm_ready) {
    cudaSetDevice(h->device);
    h->arena.release(); h->xg.release(); h->ug.release(); h->gcpart.release(); h->s2part.release(); h->dpart.release();
    h->hx[0].release(); h->hx[1].release(); h->hr.release(); h->small.release();
    h->xq.release(); h->xe.release(); h->xtq.release(); h->xte.release();
    if (h->d_flags) cudaFree(h->d_flags);
    if (h->d_scal) cudaFree(h->d_scal);
    if (h->d_ticket) cudaFree(h->d_ticket);
    if (h->d_out) cudaFree(h->d_out);
    if (h->h_out) cudaFreeHost(h->h_out);
    if (h->pin_base) cudaFreeHost(h->pin_base);
    for (int i = 0; i < sgp_handle::kPinSlots; ++i) if (h->pin_ev[i]) cudaEventDestroy(h->pin_ev[i]);
    for (int i = 0; i < kCommMaxRanks; ++i) if (h->comm_peer[i]) cudaIpcCloseMemHandle(h->comm_peer[i]);
    h->comm_buf.release();
    h->cells_out.release(); h->res_x.release(); h->res_r.release(); h->res_u.release();
    for (int i = 0; i < 2; ++i) {
      if (h->ev_copied[i]) cudaEventDestroy(h->ev_copied[i]);
      if (h->ev_used[i]) cudaEventDestroy(h->ev_used[i]);
    }
    if (h->ev_small) cudaEventDestroy(h->ev_small);
    for (cudaEvent_t e : {h->ev_fork, h->ev_join, h->ev_chain_fork, h->ev_chain_join}) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : h->prof_ev) if (e) cudaEventDestroy(e);
    for (cudaStream_t s : {h->copy_stream, h->compute_stream, h->side_stream, h->chain_stream}) if (s) cudaStreamDestroy(s);
  }
  delete h;
  return SGP_OK;
}

const char* sgp_last_error(const sgp_handle* h) { return h ? h->err.c_str() : "null handle"; }

int sgp_set_slab_rows(sgp_handle* h, long rows) {
  if (!h) return SGP_E_ARG;
  h->slab_rows = rows > 0 ? rup(rows, 128) : 65536;
  h->stream_slab_rows = rows > 0 ? std::min<long>(rup(rows, 128), 16384) : 8192;
  return SGP_OK;
}

int sgp_set_option(sgp_handle* h, const char* name, double value) {
  if (!h || !name) return SGP_E_ARG;
  const std::string n(name);
  if (n == "kl_weight") h->kl_weight = value;
  else if (n == "profile") h->profile = value != 0.0;
  else if (n == "path") h->path = (int)value;
  else if (n == "overlap") h->overlap = value != 0.0;
  else if (n == "projected") h->projected = (int)value;
  else if (n == "support") h->support = (int)value;
  else if (n == "square_is_same") h->square_is_same = value != 0.0;
  else if (n == "chol_fused") h->chol_fused = (int)value;
  else if (n == "sep_uc") h->sep_uc = value != 0.0;
  else if (n == "i8_slices") {
    if (value != -1 && value != 0 && (value < 2 || value > 8)) return fail(h, SGP_E_ARG, "sgp_set_option: i8_slices must be -1, 0 or 2..8");
    h->i8_slices = (int)value;
  }
  else if (n == "i8_slices_xtr" || n == "i8_slices_sig2") {
    if (value != -1 && (value < 2 || value > 8)) return fail(h, SGP_E_ARG, "sgp_set_option: " + n + " must be -1 or 2..8");
    (n == "i8_slices_xtr" ? h->i8_slices_xtr : h->i8_slices_sig2) = (int)value;
  }
  else if (n == "i8_slices_images") {
    if (value != -1 && (value < 2 || value > 8)) return fail(h, SGP_E_ARG, "sgp_set_option: i8_slices_images must be -1 or 2..8");
    h->i8_slices_images = (int)value;
  }
  else if (n == "slab_rows") return sgp_set_slab_rows(h, (long)value);
  else return fail(h, SGP_E_ARG, "sgp_set_option: unknown option " + n);
  return SGP_OK;
}

int sgp_get_profile(sgp_handle* h, double* stage_ms_out, int n, long* calls_out, int reset) {
  if (!h) return SGP_E_ARG;
  for (int i = 0; i < n && i < 16; ++i) stage_ms_out[i] = h->stage_ms[i];
  if (calls_out) *calls_out = h->stage_calls;
  if (reset) { for (double& v : h->stage_ms) v = 0.0; h->stage_calls = 0; }
  return SGP_OK;
}

int sgp_invalidate_cache(sgp_handle* h) {
  if (!h) return SGP_E_ARG;
  h->cache_valid = false;
  h->oz_valid = false;
  return SGP_OK;
}

int sgp_release_cache(sgp_handle* h) {
  if (!h) return SGP_E_ARG;
  h->cache_valid = false;
  h->oz_valid = false;
  if (h->gemm_ready) {
    if (cudaSetDevice(h->device) != cudaSuccess) return fail(h, SGP_E_CUDA, "sgp_release_cache: cudaSetDevice failed");
    cudaDeviceSynchronize();   // nothing of this handle may still be reading the buffers
    h->xg.release(); h->ug.release(); h->xq.release(); h->xe.release(); h->xtq.release(); h->xte.release();
    h->hx[0].release(); h->hx[1].release(); h->hr.release();
    h->arena.release(); h->gcpart.release(); h->s2part.release(); h->dpart.release();
  }
  return SGP_OK;
}

int sgp_bind_images(sgp_handle* h, unsigned long long token) {
  if (!h) return SGP_E_ARG;
  h->image_token = token;
  return SGP_OK;
}

long sgp_launch_count(const sgp_handle* h, int reset) {
  if (!h) return 0;
  const long now = launch_counter();
  const long n = now - h->launches_base;
  if (reset) const_cast<sgp_handle*>(h)->launches_base = now;
  return n;
}

int sgp_mask_host(const double* theta, int n_px, uint8_t* mask_host, int* p_out) {
  if (!theta || n_px <= 0) return SGP_E_ARG;
  std::vector<uint8_t> mask; std::vector<int> idx; std::vector<double> gx, gy;
  const int p = compute_mask(theta, n_px, mask, idx, gx, gy);
  if (mask_host) std::memcpy(mask_host, mask.data(), mask.size());
  if (p_out) *p_out = p;
  return SGP_OK;
}

int sgp_localker(sgp_handle* h, const double* theta, int n_px, double* C, long ldc, double* dC, long dc_stride,
                 void* stream) {
  if (!h || !theta || !C) return fail(h, SGP_E_ARG, "sgp_localker: null argument");
  if (n_px <= 0) return fail(h, SGP_E_ARG, "sgp_localker: n_px must be positive");
  int rc = ensure_ready(h); if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  std::vector<uint8_t> mask; std::vector<int> idx; std::vector<double> gx, gy;
  const int p = compute_mask(theta, n_px, mask, idx, gx, gy);
  if (ldc < p) return fail(h, SGP_E_ARG, "sgp_localker: ldc < p");
  RESERVE_ARENA((size_t)(2 * p + 64) * sizeof(double));
  double* dgx = (double*)h->arena.p; double* dgy = dgx + p;
  CU(cudaMemcpyAsync(dgx, gx.data(), sizeof(double) * p, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(dgy, gy.data(), sizeof(double) * p, cudaMemcpyHostToDevice, st));
  // build_C writes an ldc x ldc square; the caller's buffer is p rows x ldc -> clamp rows via p-sized launch
  const ThetaDev th = make_theta(theta);
  if (ldc == p) {
    build_C(C, ldc, p, dgx, dgy, th, dC, dc_stride, st);
  } else {
    return fail(h, SGP_E_ARG, "sgp_localker: ldc must equal p");
  }
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(st));  // gx/gy host vectors go out of scope
  return SGP_OK;
}

// Slice count of the int8 tensor-core path for a product with `rows` long-side rows and inner sizes n, k:
// 0 = use the DMMA GEMM.  Same policy as the fused closure (option "i8_slices": -1 auto, 0 off, 2..8 forced).
static int oz_pick(const sgp_handle* h, long rows, long n, long k) {
  if (h->i8_slices == 0 || n < 1 || k < 1) return 0;
  if (h->i8_slices > 0) return h->i8_slices;
  return (rows >= 4096 && n >= 64 && k >= 64) ? 8 : 0;
}

// C = alpha * op(A) op(B) (+ C) on the int8 slices; al/bl as in sgp_gemm_f64.  One pass over K (no split), so
// K * S * 4096 must stay below 2^31.
static int oz_mm(sgp_handle* h, cudaStream_t st, int al, int bl, const double* A, long lda, long M, const double* B,
                 long ldb, int N, int K, int S, double* C, long ldc, double alpha, int accumulate) {
  const int ldk = oz_ldk(K);
  size_t need = 0;
  auto acc = [&](size_t bytes) { need = ((need + 1023) & ~size_t(1023)) + bytes; };
  acc(oz_slice_bytes(M, K, S)); acc(oz_slice_bytes(N, K, S));
  acc((size_t)M * sizeof(int)); acc((size_t)N * sizeof(int)); acc((size_t)std::max<long>(M, N) * 8);
  need += 4096;
  RESERVE_ARENA(need);
  Arena ar(h->arena.p, h->arena.bytes);
  int8_t* qa = ar.take<int8_t>(oz_slice_bytes(M, K, S));
  int8_t* qb = ar.take<int8_t>(oz_slice_bytes(N, K, S));
  int* ea = ar.take<int>(M);
  int* eb = ar.take<int>(N);
  unsigned long long* mx = ar.take<unsigned long long>(std::max<long>(M, N));
  if (al == 0) CU(oz_slice_rows(A, lda, M, K, S, qa, ldk, ea, st));
  else CU(oz_slice_cols(A, lda, K, (int)M, S, qa, ldk, ea, mx, st));
  if (bl == 0) CU(oz_slice_rows(B, ldb, N, K, S, qb, ldk, eb, st));
  else CU(oz_slice_cols(B, ldb, K, N, S, qb, ldk, eb, mx, st));
  CU(oz_gemm(qa, ea, M, qb, eb, N, K, ldk, S, C, ldc, 1, 0, accumulate, st, nullptr, 0, alpha));
  CU(cudaGetLastError());
  return SGP_OK;
}

int sgp_gemm_f64(sgp_handle* h, int al, int bl, const double* A, long lda, const double* B, long ldb, double* C,
                 long ldc, int M, int N, int K, double alpha, int accumulate, int small_tile, void* stream) {
  if (!h || !A || !B || !C) return fail(h, SGP_E_ARG, "sgp_gemm_f64: null argument");
  if (M <= 0 || N <= 0 || K <= 0) return fail(h, SGP_E_ARG, "sgp_gemm_f64: sizes must be positive");
  // layouts: A is M x K (al == 0) or K x M (al == 1); B is N x K (bl == 0) or K x N (bl == 1)
  if (lda < (al ? M : K) || ldb < (bl ? N : K) || ldc < N)
    return fail(h, SGP_E_ARG, "sgp_gemm_f64: leading dimension smaller than the row length");
  int rc = ensure_ready(h); if (rc) return rc;
  // image-sized products (K_b = K B, a = K_b K_tilde_b^-1: utils.py:1693-1694) go to the int8 tensor-core path
  if (al == 0 && !small_tile && h->i8_slices != 0 && M >= 4096 && N >= 64 && K >= 64 &&
      rup(K, 64) * 8L * 4096 < (1L << 31))
    return oz_mm(h, (cudaStream_t)stream, al, bl, A, lda, M, B, ldb, N, K, h->i8_slices > 0 ? h->i8_slices : 8, C, ldc,
                 alpha, accumulate);
  GemmArgs g{};
  g.A = A; g.lda = lda; g.B = B; g.ldb = ldb; g.C = C; g.ldc = ldc; g.M = M; g.N = N; g.K = K;
  g.alpha = alpha; g.accumulate = accumulate; g.k_chunk = (int)rup(K, 16);
  CU(gemm_f64_launch(g, GemmLaunch{al, bl, kStore, small_tile != 0, 1}, (cudaStream_t)stream));
  return SGP_OK;
}

int sgp_gemm_i8emu(sgp_handle* h, int layout, const double* A, long M, long lda, const double* B, int N, long ldb,
                   long K, int slices, double* C, long ldc, void* stream) {
  if (!h || !A || !B || !C) return fail(h, SGP_E_ARG, "sgp_gemm_i8emu: null argument");
  if (slices < 2 || slices > 8) return fail(h, SGP_E_ARG, "sgp_gemm_i8emu: slices must be in 2..8");
  if (M <= 0 || N <= 0 || K <= 0 || K > (1L << 30)) return fail(h, SGP_E_ARG, "sgp_gemm_i8emu: bad sizes");
  if (ldc < N || lda < (layout ? M : K) || ldb < (layout ? N : K))
    return fail(h, SGP_E_ARG, "sgp_gemm_i8emu: leading dimension smaller than the row length");
  int rc = ensure_ready(h); if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int ldk = oz_ldk((int)K);
  const int nz = oz_split(M, N, (int)K, slices);
  const long zs = M * ldc;
  size_t need = 0;
  auto acc = [&](size_t bytes) { need = ((need + 1023) & ~size_t(1023)) + bytes; };
  acc(oz_slice_bytes(M, (int)K, slices)); acc(oz_slice_bytes(N, (int)K, slices));
  acc((size_t)M * sizeof(int)); acc((size_t)N * sizeof(int)); acc((size_t)std::max<long>(M, N) * 8);
  acc(nz > 1 ? (size_t)nz * zs * sizeof(double) : 8);
  need += 4096;
  RESERVE_ARENA(need);
  Arena ar(h->arena.p, h->arena.bytes);
  int8_t* qa = ar.take<int8_t>(oz_slice_bytes(M, (int)K, slices));
  int8_t* qb = ar.take<int8_t>(oz_slice_bytes(N, (int)K, slices));
  int* ea = ar.take<int>(M);
  int* eb = ar.take<int>(N);
  unsigned long long* mx = ar.take<unsigned long long>(std::max<long>(M, N));
  double* part = ar.take<double>(nz > 1 ? (size_t)nz * zs : 1);
  if (layout == 0) {          // A: M x K, B: N x K (K contiguous)
    CU(oz_slice_rows(A, lda, M, (int)K, slices, qa, ldk, ea, st));
    CU(oz_slice_rows(B, ldb, N, (int)K, slices, qb, ldk, eb, st));
  } else {                    // A: K x M, B: K x N (contraction over the rows)
    CU(oz_slice_cols(A, lda, K, (int)M, slices, qa, ldk, ea, mx, st));
    CU(oz_slice_cols(B, ldb, K, N, slices, qb, ldk, eb, mx, st));
  }
  if (nz > 1) {
    CU(oz_gemm(qa, ea, M, qb, eb, N, (int)K, ldk, slices, part, ldc, nz, zs, 0, st));
    const int kbz = (((int)K + 63) / 64 + nz - 1) / nz;
    slice_reduce(part, (((int)K + 63) / 64 + kbz - 1) / kbz, zs, zs, C, 0, st);
  } else {
    CU(oz_gemm(qa, ea, M, qb, eb, N, (int)K, ldk, slices, C, ldc, 1, 0, 0, st));
  }
  CU(cudaGetLastError());
  return SGP_OK;
}

int sgp_cholesky(sgp_handle* h, double* A, long ld, int n, int* flags_host, void* stream) {
  if (!h || !A) return fail(h, SGP_E_ARG, "sgp_cholesky: null argument");
  if (n <= 0 || ld < n) return fail(h, SGP_E_ARG, "sgp_cholesky: need n > 0 and ld >= n");
  int rc = ensure_ready(h); if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  CU(cudaMemsetAsync(h->d_flags, 0, sizeof(int) * 4, st));
  RESERVE_ARENA((size_t)chol_winv_elems(n) * sizeof(double) + 4096);
  CU(factor_any(h, A, ld, n, nullptr, nullptr, (double*)h->arena.p, kFlagKtNotPD, kFlagKtSmallPivot, -1, st));
  if (flags_host) {
    CU(cudaMemcpyAsync(flags_host, h->d_flags, sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
  }
  return SGP_OK;
}

int sgp_spd_inverse(sgp_handle* h, const double* A, long lda, int n, double* Ainv, long ldi, double* logdet_host,
                    int* flags_host, void* stream) {
  if (!h || !A || !Ainv) return fail(h, SGP_E_ARG, "sgp_spd_inverse: null argument");
  if (n <= 0 || lda < n || ldi < n) return fail(h, SGP_E_ARG, "sgp_spd_inverse: need n > 0 and leading dimensions >= n");
  int rc = ensure_ready(h); if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const long ld = rup(n, 16);
  RESERVE_ARENA((size_t)(3 * ld * ld + 64 * ld + chol_winv_elems(n) + 1024) * sizeof(double));
  Arena ar(h->arena.p, h->arena.bytes);
  double* L = ar.take<double>(ld * ld);
  double* X = ar.take<double>(ld * ld);
  double* R = ar.take<double>(ld * ld);
  double* tmp = ar.take<double>(64 * ld);
  double* winv = ar.take<double>(chol_winv_elems(n));
  CU(cudaMemsetAsync(h->d_flags, 0, sizeof(int) * 4, st));
  copy_mat(A, lda, n, n, L, ld, st);
  CU(factor_any(h, L, ld, n, X, tmp, winv, kFlagKtNotPD, kFlagKtSmallPivot, kSlotLogdetK, st));
  CU(mm(st, 1, 1, X, ld, X, ld, R, ld, n, n, n));
  symmetrize(R, ld, n, Ainv, ldi, st);
  CU(cudaGetLastError());
  if (logdet_host) CU(cudaMemcpyAsync(logdet_host, h->d_scal + kSlotLogdetK, sizeof(double), cudaMemcpyDeviceToHost, st));
  if (flags_host) CU(cudaMemcpyAsync(flags_host, h->d_flags, sizeof(int), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return SGP_OK;
}

int sgp_debug_get(sgp_handle* h, const char* name, double* dst, long capacity, long* rows, long* cols, long* ld) {
  if (!h || !name) return SGP_E_ARG;
  if (std::string(name) == "oz_prof") {   // cycle counters of the last int8 GEMM launch (SGP_OZ_DEBUG & 64)
    if (rows) *rows = 1;
    if (cols) *cols = 8;
    if (ld) *ld = 8;
    if (dst) {
      if (capacity < 8 || !oz_prof_ptr()) return fail(h, SGP_E_ARG, "sgp_debug_get: oz_prof unavailable");
      CU(cudaMemcpy(dst, oz_prof_ptr(), sizeof(double) * 8, cudaMemcpyDeviceToDevice));
    }
    return SGP_OK;
  }
  auto it = h->dbg.find(name);
  if (it == h->dbg.end()) return fail(h, SGP_E_ARG, std::string("sgp_debug_get: unknown name ") + name);
  const DebugEntry& e = it->second;
  if (rows) *rows = e.rows;
  if (cols) *cols = e.cols;
  if (ld) *ld = e.ld;
  if (dst) {
    const long need = e.rows * e.ld;
    if (capacity < need) return fail(h, SGP_E_ARG, "sgp_debug_get: capacity too small");
    CU(cudaMemcpy(dst, e.ptr, sizeof(double) * need, cudaMemcpyDeviceToDevice));
  }
  return SGP_OK;
}

}  // extern "C"

// =====================================================================================================
// the fused closure
// =====================================================================================================
namespace {

cudaError_t ensure_copy_stream(sgp_handle* h) {
  if (h->copy_stream) return cudaSuccess;
  cudaError_t e = cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) return e;
  for (int i = 0; i < 2; ++i) {
    if ((e = cudaEventCreateWithFlags(&h->ev_copied[i], cudaEventDisableTiming)) != cudaSuccess) return e;
    if ((e = cudaEventCreateWithFlags(&h->ev_used[i], cudaEventDisableTiming)) != cudaSuccess) return e;
  }
  return cudaEventCreateWithFlags(&h->ev_small, cudaEventDisableTiming);
}

struct ClosureIn {
  const double* X; long ldx; const double* r; long N;        // device (or per-slab device staging when streaming)
  const double* U; long ldu; int m; int n_px;
  const double* theta; const double* B; long ldb; int q; const double* m_b; const double* V; long ldv;
  double logA, lambda0; const double* logdetV_host;
  double* lam_m; double* lam_var; double* f;
  // host streaming (X_host != nullptr): X / r above are ignored
  const double* X_host = nullptr; const double* r_host = nullptr;
  bool first_copy_issued = false;   // the caller already started the H2D copy of slab 0 into hx[0]
  // asynchronous mode: the 16 result doubles go to out_dev (device memory) and the call returns without waiting for the
  // stream -- the caller reads / reduces them in stream order (cells loop, device-side all-reduce of the image shards)
  double* out_dev = nullptr;
  int force_support = -2;           // -2: the handle's option; otherwise overrides it for this call
  // float32 images (sgp_elbo_grad_f32): X / U above then hold the same addresses (cache key only) and every read of the
  // images goes through the fp64 working copy that the gather makes once per image set
  const float* X32 = nullptr; const float* U32 = nullptr;
  bool exchange = false;            // image shard of a larger evaluation: reduce the result with the peers (sgp_comm_*)
};

// one slot of the pinned upload ring, at least `bytes` large; waits until the copies last issued from it are done
int pin_acquire(sgp_handle* h, size_t bytes, char** out, int* slot_out) {
  if (bytes > h->pin_slot_bytes) {
    CU(cudaDeviceSynchronize());
    if (h->pin_base) cudaFreeHost(h->pin_base);
    h->pin_base = nullptr; h->pin_slot_bytes = 0;
    const size_t sb = (bytes + 65535) & ~size_t(65535);
    CU(cudaMallocHost((void**)&h->pin_base, sb * sgp_handle::kPinSlots));
    h->pin_slot_bytes = sb;
    for (int i = 0; i < sgp_handle::kPinSlots; ++i) h->pin_used[i] = false;
  }
  const int sl = h->pin_next;
  h->pin_next = (h->pin_next + 1) % sgp_handle::kPinSlots;
  if (!h->pin_ev[sl]) CU(cudaEventCreateWithFlags(&h->pin_ev[sl], cudaEventDisableTiming));
  if (h->pin_used[sl]) CU(cudaEventSynchronize(h->pin_ev[sl]));
  *out = h->pin_base + (size_t)sl * h->pin_slot_bytes;
  *slot_out = sl;
  return SGP_OK;
}

int run_closure_same(sgp_handle* h, const ClosureIn& in, double* out_host, cudaStream_t st);

int run_closure(sgp_handle* h, const ClosureIn& in, double* out_host, cudaStream_t st) {
  const double t_entry = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
  int rc = ensure_ready(h); if (rc) return rc;
  h->prof_used = 0;
  auto mark = [&](int stage) {   // "stage begins now" (only when profiling)
    if (!h->profile) return;
    if (h->prof_used == h->prof_ev.size()) {
      cudaEvent_t e; cudaEventCreate(&e); h->prof_ev.push_back(e); h->prof_stage.push_back(0);
    }
    h->prof_stage[h->prof_used] = stage;
    cudaEventRecord(h->prof_ev[h->prof_used++], st);
  };
  mark(0);
  const bool streaming = in.X_host != nullptr;
  const long N = in.N; const int m = in.m, q = in.q, n = in.n_px; const long P = (long)n * n;
  if (N <= 0 || m <= 0 || q <= 0 || q > m) return fail(h, SGP_E_ARG, "sgp_elbo_grad: bad sizes");
  if (m > kBwdMaxCols) return fail(h, SGP_E_UNSUPPORTED, "sgp_elbo_grad: m > 2048 not supported");
  // a row of images must hold n_px^2 pixels: a smaller leading dimension means the caller's n_px does not match its images
  // (the reference fails on the shape mismatch in x[:, mask], U:2031); reading on would run past the buffers
  if (n <= 0 || in.ldx < P || in.ldu < P || in.ldb < q || in.ldv < q)
    return fail(h, SGP_E_ARG, "sgp_elbo_grad: leading dimension smaller than the row length (n_px^2 pixels per image)");
  // ntilde == nt: the reference's a = B branch (U:2032, U:2068) -- the configuration its active-learning notebook
  // always runs (m == N: every image in use is an inducing image).  Small problem, own routine.
  if (N == m && h->square_is_same) {
    if (streaming) return fail(h, SGP_E_UNSUPPORTED, "sgp_elbo_grad_host: ntilde == nt needs device-resident images");
    return run_closure_same(h, in, out_host, st);
  }

  // ---- host: mask / coordinates (U:880-887) ------------------------------------------------------
  std::vector<uint8_t> mask; std::vector<int> idx; std::vector<double> gx, gy;
  const int p_true = compute_mask(in.theta, n, mask, idx, gx, gy);
  if (p_true == 0) return fail(h, SGP_E_ARG, "sgp_elbo_grad: empty receptive-field mask");
  // ---- support: the pixel set the evaluation runs on, a SUPERSET of the mask ------------------------------
  // The mask is recomputed from theta in every closure (U:2030) and flips pixels inside an LBFGS line search (parity
  // trap 5).  Gathering X[:, mask] and slicing it again at every flip costs more than a few extra columns, so the
  // gathered copy / int8 slices are built on a support S >= mask (the mask grown by two pixels, or the whole grid) and
  // kept while the mask stays inside S.  Pixels of S outside the mask get C = 0 rows / columns (coordinates at 1e30:
  // alpha = exp(-e^beta * 1e60) = 0 exactly, so C, every dC_k and the separable weights vanish there): they add exact
  // zeros to every product, and the result is the one the reference computes on x[:, mask].
  std::vector<uint8_t> support = mask;
  const int support_opt = in.force_support != -2 ? in.force_support : h->support;
  if (p_true < (int)P && !streaming && support_opt != 0) {
    const bool same_images = h->cache_valid && h->cache_token == h->image_token && h->cache_X == in.X && h->cache_N == N &&
                             h->cache_ldx == in.ldx && h->cache_U == in.U && h->cache_m == m && h->cache_ldu == in.ldu &&
                             h->cache_mask.size() == (size_t)P;
    bool reuse = false;
    if (same_images && support_opt < 0) {
      long cs = 0; bool inside = true;
      for (long k = 0; k < P; ++k) { cs += h->cache_mask[k]; if (mask[k] && !h->cache_mask[k]) inside = false; }
      reuse = inside && 10L * p_true >= 6L * cs;     // and the support has not become wastefully large for this mask
    }
    if (reuse) support = h->cache_mask;
    else if (support_opt > 0) support.assign((size_t)P, 1);
    else {
      for (int pass = 0; pass < 2; ++pass) {         // grow by one pixel ring per pass (8-neighbourhood)
        std::vector<uint8_t> g = support;
        for (int iy = 0; iy < n; ++iy)
          for (int ix = 0; ix < n; ++ix) {
            if (support[(size_t)iy * n + ix]) continue;
            bool near = false;
            for (int dy = -1; dy <= 1 && !near; ++dy)
              for (int dx = -1; dx <= 1; ++dx) {
                const int y = iy + dy, x = ix + dx;
                if (y >= 0 && y < n && x >= 0 && x < n && support[(size_t)y * n + x]) { near = true; break; }
              }
            if (near) g[(size_t)iy * n + ix] = 1;
          }
        support.swap(g);
      }
      long cs = 0; for (long k = 0; k < P; ++k) cs += support[k];
      if (10 * cs >= 9 * P) support.assign((size_t)P, 1);   // nearly everything: read the images in place instead
    }
  }
  if (support != mask) {      // index / coordinate lists over the support; 1e30 marks the pixels outside the mask
    idx.clear(); gx.clear(); gy.clear();
    for (int iy = 0; iy < n; ++iy)
      for (int ix = 0; ix < n; ++ix) {
        const size_t k = (size_t)iy * n + ix;
        if (!support[k]) continue;
        idx.push_back((int)k);
        gx.push_back(mask[k] ? lin_coord(ix, n) : 1.0e30);
        gy.push_back(mask[k] ? lin_coord(iy, n) : 1.0e30);
      }
  }
  const int p = (int)idx.size();     // from here on "p" is the size of the support; p_true is what the reference calls p
  const ThetaDev th = make_theta(in.theta);
  const double A = std::exp(in.logA);
  const long ldp = rup(p, 16), ldm = rup(m, 16), ldq = rup(q, 16);
  const long slab = std::min<long>(streaming ? h->stream_slab_rows : h->slab_rows, rup(N, 128));
  const long ldpartN = rup(slab, 16), ldpartU = ldm;
  const int ntp = (int)((p + 127) / 128), ntm = (int)((m + 127) / 128);
  const int ntp_s = (int)((p + 63) / 64);
  const bool full_mask = (p == P);            // of the SUPPORT: the images are read in place, no gathered copy
  // separable path (sep_quad.cu): 8 n^3 flops per image for the quadratic forms instead of 4 p^2 dense ones
  bool use_sep = sep_quad_supported(n) && (12.0 * n * n * n < 4.0 * (double)p * p);
  if (h->path == 1) use_sep = false;
  if (h->path == 2) use_sep = sep_quad_supported(n);
  // int8-slice tensor-core products for K_lin = X (UC)^T and D = X^T R (separable path only)
  // ozM: the two m-space products (T = K Mfull^T, Sig2 = K^T diag(v) K) -- on either path;  ozS: the two products with the
  // images (separable path only: the dense path's X C / X^T Y keep the DMMA GEMM with its fused epilogues)
  int ozM = 0;
  if (h->i8_slices != 0) ozM = h->i8_slices > 0 ? h->i8_slices : (slab >= 4096 && m >= 64 ? 8 : 0);
  const int ozS = use_sep ? ozM : 0;
  // ozI: slice count of the two products that contract with the images (K_lin = X (UC)^T, D = X^T R): they are 4Npm of
  // the executed flops and sit upstream of the cancellation in lam_var (which lives in T = K Mfull^T, kept at ozM)
  // Default (both options on auto): 7 slices = 28 slice pairs instead of 36.  Measured against the real reference at
  // n = 40, m = 500 / 1000 (tests/test_gpu_bigsize.py, profiles/r2_parity_report_v1.jsonl): gradients 3e-11 of their
  // cancellation scale, lam_m 4e-12 -- the same class as 8 slices, because the moments' cancellation is downstream.
  const int ozI = ozS ? (h->i8_slices_images > 0 ? h->i8_slices_images : (h->i8_slices < 0 ? 7 : ozS)) : 0;
  // Moments in the projected basis when the kept subspace is small (q < m): K_b = K B, T_b = K_b Mg^T, T = T_b B^T and
  // K_b^T diag(v) K_b cost 4Nmq + 3Nq^2 flops against the 3Nm^2 of K Mfull^T and K^T diag(v) K (break-even q = 0.54 m;
  // the extra slicing passes move it down).  Same numbers up to rounding: Mfull = B Mg B^T.
  const bool proj = ozM > 0 && h->projected != 0 && q >= 16 && (h->projected > 0 || 5 * (long)q <= 2 * (long)m);
  const int ldkP = oz_ldk(p), ldkN = oz_ldk((int)slab);
  const bool chain_overlap = use_sep && h->overlap != 0;
  const bool tail_overlap = chain_overlap && ozS > 0;
  const int ozX = ozI ? (h->i8_slices_xtr > 0 ? std::min(h->i8_slices_xtr, ozI) : ozI) : 0;      // D = X^T R
  // Sigma2 enters only the gradient (through S2b), not lam_var: with everything on auto it takes 7 slices as well.
  // Measured against the real reference (scripts/probe_gradslices.py, round 2): gradients <= 5e-11 of their
  // cancellation scale on every fixture (8 slices: 3e-11).  X^T R at 6 slices would save another 0.33 ms but leaves
  // only a 3x margin (3e-10 at m = 1000), so it keeps the images' count.
  const int ozW = ozM ? (h->i8_slices_sig2 > 0 ? std::min(h->i8_slices_sig2, ozM) : (h->i8_slices < 0 ? std::min(7, ozM) : ozM)) : 0;
  const int oz_pairs = (std::max(ozM, ozI) + 1) / 2, oz_pairsI = (ozI + 1) / 2;
  const int nzO = ozS ? oz_split(p, m, (int)std::min(slab, N), ozI) : 1;
  const int nzK = ozM ? (proj ? oz_split(q, q, (int)std::min(slab, N), ozM, 1) : oz_split(m, m, (int)std::min(slab, N), ozM, 1)) : 1;
  std::vector<int> off; std::vector<double> alpha;
  if (use_sep) {
    const int ldq_sep = sep_quad_ld(n);
    off.resize(p); alpha.resize(p);
    for (int c = 0; c < p; ++c) {
      const int iy = idx[c] / n, ix = idx[c] % n;
      off[c] = iy * ldq_sep + ix;
      alpha[c] = mask[idx[c]] ? std::exp(-th.eb * ((gx[c] - th.ex) * (gx[c] - th.ex) + (gy[c] - th.ey) * (gy[c] - th.ey)))
                              : 0.0;
    }
  }

  // ---- gathered / aligned copies of the images (cached across calls while the mask is unchanged) --
  const bool direct_X = !streaming && !in.X32 && full_mask && (in.ldx % 2 == 0) && ((uintptr_t)in.X % 16 == 0);
  const bool direct_U = !in.U32 && full_mask && (in.ldu % 2 == 0) && ((uintptr_t)in.U % 16 == 0);
  bool regather = !h->cache_valid || h->cache_token != h->image_token || h->cache_mask != support || h->cache_X != in.X ||
                  h->cache_N != N || h->cache_ldx != in.ldx || h->cache_U != in.U || h->cache_m != m ||
                  h->cache_ldu != in.ldu;
  if (streaming) regather = true;
  // the caches count as valid only once this call has completed (a failure half way must not leave a 'valid' key)
  const bool cache_was_valid = h->cache_valid && !regather;
  const bool oz_was_valid = h->oz_valid && !regather;
  h->cache_valid = false; h->oz_valid = false;
  const double* Xg = in.X; long ldxg = in.ldx;
  const double* Ug = in.U; long ldug = in.ldu;
  if (!direct_X && !streaming) { CU(h->xg.reserve((size_t)N * ldp * sizeof(double))); Xg = (double*)h->xg.p; ldxg = ldp; }
  if (!direct_U) { CU(h->ug.reserve((size_t)m * ldp * sizeof(double))); Ug = (double*)h->ug.p; ldug = ldp; }

  // ---- arena ---------------------------------------------------------------------------------------
  const int nblk_bwd = acos_backward_blocks(slab), nblk_bwdU = acos_backward_blocks(m);
  const int nblk_mom = moments_rows_blocks(slab);
  size_t need = 0;
  auto acc = [&](size_t nelem) { need = ((need + 255) & ~size_t(255)) + nelem * sizeof(double); };
  acc(p); acc(p); acc((size_t)(p + 1) / 2 + 8);                       // gx, gy, idx
  acc((size_t)p * ldp);                                               // C
  acc((size_t)m * ldq); acc((size_t)q * ldq);                         // Bw, Vw
  acc((size_t)m * ldp); acc((size_t)std::max(ntp_s, (p + 15) / 16) * ldpartU); acc(m); acc(m); // ZU/YU, partU, tU, X2
  for (int i = 0; i < 6; ++i) acc((size_t)m * ldm);                   // Kraw, Kt, cosU, Mfull, Sig2, GKt
  for (int i = 0; i < 3; ++i) acc((size_t)m * ldq);                   // tmp_mq x3
  for (int i = 0; i < 14; ++i) acc((size_t)q * ldq);                  // q x q scratch
  acc(64 * ldq); acc(chol_winv_elems(q)); acc(chol_winv_elems(q));    // tri-inverse tmp, inverses of L's diagonal blocks (x2)
  for (int i = 0; i < 12; ++i) acc(ldm);                              // small vectors
  acc(use_sep ? 1 : (size_t)slab * ldp);                              // Z / Y (dense path only)
  if (use_sep) {
    acc((size_t)(p + 1) / 2 + 8); acc(p); acc((size_t)slab * 5); acc((size_t)wq_reduce_blocks(slab) * 5);
    acc((size_t)p * ldm);                                             // D = X^T R
  }
  for (int i = 0; i < 3; ++i) acc((size_t)slab * ldm);                // K, cos, T/R
  for (int i = 0; i < 6; ++i) acc(slab);                              // s, X1, u, v, w (+1)
  for (int i = 0; i < 3; ++i) acc(N);                                 // lam_m, lam_var, f (when not supplied)
  const int ng16 = (m + 15) / 16;                                     // 16-column groups of the int8 moments epilogue
  acc((size_t)std::max(std::max(ntp, ntm), ozM ? ng16 : 0) * ldpartN * 3);   // partN, pm1, pm2
  acc((size_t)nblk_mom * 4); acc((size_t)nblk_bwd * ldm * 4); acc(nblk_bwd);
  acc((size_t)nblk_bwdU * ldm); acc(nblk_bwdU);
  acc((size_t)p * ldp);                                               // GCU
  acc((size_t)p * 5);                                                 // gpart
  if (streaming) { acc((size_t)slab * ldp); }                         // gathered slab when streaming from host
  if (ozM) {                                                          // slices of UC and of R^T, exponents, col-max scratch
    acc(oz_slice_bytes(m, p, std::max(ozM, ozI)) / 8 + 16); acc((size_t)oz_pairs * m * ldkN / 8 + 16); acc(m); acc(m);
    acc((size_t)std::max(p, m) + 8);
    acc(oz_slice_bytes(slab, m, ozM) / 8 + 16); acc(oz_slice_bytes(m, m, ozM) / 8 + 16); acc(slab / 2 + 8); acc(m);   // K, Mfull slices
  }
  if (ozS) {                                                          // slices of Ug, C, Ug^T, D for ZU = Ug C and GC = D Ug
    acc(oz_slice_bytes(m, p, ozS) / 8 + 16); acc(oz_slice_bytes(p, p, ozS) / 8 + 16);
    acc(oz_slice_bytes(p, m, ozS) / 8 + 16); acc(oz_slice_bytes(p, m, ozS) / 8 + 16);
    acc(m); acc(p); acc(p); acc(p);
  }
  if (proj) {                                                         // K_b, T_b and the slices of K_b, T_b, B, B^T, Mg
    acc((size_t)slab * ldq); acc((size_t)slab * ldq);
    acc(oz_slice_bytes(slab, q, ozM) / 8 + 16); acc(oz_slice_bytes(slab, q, ozM) / 8 + 16);
    acc(oz_slice_bytes(m, q, ozM) / 8 + 16); acc(oz_slice_bytes(q, m, ozM) / 8 + 16); acc(oz_slice_bytes(q, q, ozM) / 8 + 16);
    acc((size_t)(ozM + 1) / 2 * q * ldkN / 8 + 16);
    acc(slab / 2 + 8); acc(slab / 2 + 8); acc(m); acc(q); acc(q); acc(q);
  }
  need += 4096;
  RESERVE_ARENA(need);
  Arena ar(h->arena.p, h->arena.bytes);

  double* d_gx = ar.take<double>(p); double* d_gy = ar.take<double>(p);
  int* d_idx = (int*)ar.take<double>((size_t)(p + 1) / 2 + 8);
  double* C = ar.take<double>((size_t)p * ldp);
  double* Bw = ar.take<double>((size_t)m * ldq); double* Vw = ar.take<double>((size_t)q * ldq);
  double* ZU = ar.take<double>((size_t)m * ldp);
  double* partU = ar.take<double>((size_t)std::max(ntp_s, (p + 15) / 16) * ldpartU);
  double* tU = ar.take<double>(m); double* X2 = ar.take<double>(m);
  double* Kraw = ar.take<double>((size_t)m * ldm); double* Kt = ar.take<double>((size_t)m * ldm);
  double* cosU = ar.take<double>((size_t)m * ldm); double* Mfull = ar.take<double>((size_t)m * ldm);
  double* Sig2 = ar.take<double>((size_t)m * ldm); double* GKt = ar.take<double>((size_t)m * ldm);
  double* tmq1 = ar.take<double>((size_t)m * ldq); double* tmq2 = ar.take<double>((size_t)m * ldq);
  double* tmq3 = ar.take<double>((size_t)m * ldq);
  double* qq[14]; for (int i = 0; i < 14; ++i) qq[i] = ar.take<double>((size_t)q * ldq);
  double* tri_tmp = ar.take<double>(64 * ldq);
  double* winv = ar.take<double>(chol_winv_elems(q));
  double* winv2 = ar.take<double>(chol_winv_elems(q));
  double* vecs[12]; for (int i = 0; i < 12; ++i) vecs[i] = ar.take<double>(ldm);
  double* Zs = ar.take<double>(use_sep ? 1 : (size_t)slab * ldp);
  int* d_off = nullptr; double* d_alpha = nullptr; double* Qs = nullptr; double* wqpart = nullptr; double* Dm = nullptr;
  if (use_sep) {
    d_off = (int*)ar.take<double>((size_t)(p + 1) / 2 + 8);
    d_alpha = ar.take<double>(p);
    Qs = ar.take<double>((size_t)slab * 5);
    wqpart = ar.take<double>((size_t)wq_reduce_blocks(slab) * 5);
    Dm = ar.take<double>((size_t)p * ldm);
  }
  double* Ks = ar.take<double>((size_t)slab * ldm); double* cos_s = ar.take<double>((size_t)slab * ldm);
  double* Ts = ar.take<double>((size_t)slab * ldm);
  double* s_s = ar.take<double>(slab); double* X1 = ar.take<double>(slab); double* u_s = ar.take<double>(slab);
  double* v_s = ar.take<double>(slab); double* w_s = ar.take<double>(slab); (void)ar.take<double>(slab);
  double* lam_m = in.lam_m ? in.lam_m : ar.take<double>(N);
  double* lam_var = in.lam_var ? in.lam_var : ar.take<double>(N);
  double* fvec = in.f ? in.f : ar.take<double>(N);
  const int ntmax = std::max(std::max(ntp, ntm), ozM ? ng16 : 0);
  double* partN = ar.take<double>((size_t)ntmax * ldpartN); double* pm1 = ar.take<double>((size_t)ntmax * ldpartN);
  double* pm2 = ar.take<double>((size_t)ntmax * ldpartN);
  double* blockpart = ar.take<double>((size_t)nblk_mom * 4);
  double* colk = ar.take<double>((size_t)nblk_bwd * ldm); double* coly = ar.take<double>((size_t)nblk_bwd * ldm);
  double* cmaxR = ar.take<double>((size_t)nblk_bwd * ldm); double* cmaxW = ar.take<double>((size_t)nblk_bwd * ldm);
  double* sumpartR = ar.take<double>(nblk_bwd);
  double* colkU = ar.take<double>((size_t)nblk_bwdU * ldm); double* sumpartU = ar.take<double>(nblk_bwdU);
  double* GCU = ar.take<double>((size_t)p * ldp);
  double* gpart = ar.take<double>((size_t)p * 5);
  double* Xslab_stream = streaming ? ar.take<double>((size_t)slab * ldp) : nullptr;
  int8_t* kq = nullptr; int8_t* mq = nullptr; int* ke = nullptr; int* me = nullptr;
  int8_t* zuq = nullptr; int8_t* rtq = nullptr; int* zue = nullptr; int* rte = nullptr; unsigned long long* ozmx = nullptr;
  if (ozM) {
    zuq = (int8_t*)ar.take<double>(oz_slice_bytes(m, p, std::max(ozM, ozI)) / 8 + 16);
    rtq = (int8_t*)ar.take<double>((size_t)oz_pairs * m * ldkN / 8 + 16);
    zue = (int*)ar.take<double>(m); rte = (int*)ar.take<double>(m);
    ozmx = (unsigned long long*)ar.take<double>((size_t)std::max(p, m) + 8);
    kq = (int8_t*)ar.take<double>(oz_slice_bytes(slab, m, ozM) / 8 + 16);
    mq = (int8_t*)ar.take<double>(oz_slice_bytes(m, m, ozM) / 8 + 16);
    ke = (int*)ar.take<double>(slab / 2 + 8); me = (int*)ar.take<double>(m);
  }

  int8_t* ugq = nullptr; int8_t* cq = nullptr; int8_t* utq = nullptr; int8_t* dq = nullptr;
  int* uge = nullptr; int* ce = nullptr; int* ute = nullptr; int* de = nullptr;
  if (ozS) {
    ugq = (int8_t*)ar.take<double>(oz_slice_bytes(m, p, ozS) / 8 + 16);
    cq = (int8_t*)ar.take<double>(oz_slice_bytes(p, p, ozS) / 8 + 16);
    utq = (int8_t*)ar.take<double>(oz_slice_bytes(p, m, ozS) / 8 + 16);
    dq = (int8_t*)ar.take<double>(oz_slice_bytes(p, m, ozS) / 8 + 16);
    uge = (int*)ar.take<double>(m); ce = (int*)ar.take<double>(p); ute = (int*)ar.take<double>(p); de = (int*)ar.take<double>(p);
  }
  double* Kb = nullptr; double* Tb = nullptr;
  int8_t* kbq = nullptr; int8_t* tbq = nullptr; int8_t* brq = nullptr; int8_t* btq = nullptr; int8_t* mgq = nullptr;
  int8_t* kbtq = nullptr; int* kbe = nullptr; int* tbe = nullptr; int* bre = nullptr; int* bte = nullptr; int* mge = nullptr;
  int* kbte = nullptr;
  if (proj) {
    Kb = ar.take<double>((size_t)slab * ldq); Tb = ar.take<double>((size_t)slab * ldq);
    kbq = (int8_t*)ar.take<double>(oz_slice_bytes(slab, q, ozM) / 8 + 16);
    tbq = (int8_t*)ar.take<double>(oz_slice_bytes(slab, q, ozM) / 8 + 16);
    brq = (int8_t*)ar.take<double>(oz_slice_bytes(m, q, ozM) / 8 + 16);
    btq = (int8_t*)ar.take<double>(oz_slice_bytes(q, m, ozM) / 8 + 16);
    mgq = (int8_t*)ar.take<double>(oz_slice_bytes(q, q, ozM) / 8 + 16);
    kbtq = (int8_t*)ar.take<double>((size_t)(ozM + 1) / 2 * q * ldkN / 8 + 16);
    kbe = (int*)ar.take<double>(slab / 2 + 8); tbe = (int*)ar.take<double>(slab / 2 + 8);
    bre = (int*)ar.take<double>(m); bte = (int*)ar.take<double>(q); mge = (int*)ar.take<double>(q);
    kbte = (int*)ar.take<double>(q);
  }

  double* Ktb_raw = qq[0]; double* Ktb = qq[1]; double* Lk = qq[2]; double* Linv = qq[3]; double* Kinv_raw = qq[4];
  double* Kinv = qq[5]; double* KV = qq[6]; double* KVK = qq[7]; double* Mg = qq[8];   // qq[9]: Cholesky factor of V_b (side stream)
  double* S2b = qq[10]; double* W2 = qq[11]; double* inner = qq[12]; double* Sq = qq[13];
  double* GKtb = Ktb_raw;  // reused after the forward pass
  double* bq = vecs[0]; double* bb = vecs[1]; double* kappa = vecs[2]; double* yv = vecs[3]; double* beta = vecs[4];
  double* y_b = vecs[5]; double* wr = vecs[6]; double* kappaU = vecs[7]; double* wc = vecs[8]; double* wU = vecs[9];

  // ---- uploads -------------------------------------------------------------------------------------
  // Host-streaming calls send EVERY host->device copy through the copy stream, in the order they are needed: the
  // DMA engine serves one copy at a time, and a small upload queued behind a 200 MB slab would stall the compute
  // stream for the slab's whole duration.
  if (streaming) CU(ensure_copy_stream(h));
  cudaStream_t up = streaming ? h->copy_stream : st;
  {
    char* pin = nullptr; int pslot = 0;
    const size_t pb = (size_t)p * 8;
    rc = pin_acquire(h, 5 * pb + 64, &pin, &pslot); if (rc) return rc;
    std::memcpy(pin, gx.data(), pb); std::memcpy(pin + pb, gy.data(), pb); std::memcpy(pin + 2 * pb, idx.data(), (size_t)p * 4);
    CU(cudaMemcpyAsync(d_gx, pin, pb, cudaMemcpyHostToDevice, up));
    CU(cudaMemcpyAsync(d_gy, pin + pb, pb, cudaMemcpyHostToDevice, up));
    CU(cudaMemcpyAsync(d_idx, pin + 2 * pb, (size_t)p * 4, cudaMemcpyHostToDevice, up));
    if (use_sep) {
      std::memcpy(pin + 3 * pb, off.data(), (size_t)p * 4); std::memcpy(pin + 4 * pb, alpha.data(), pb);
      CU(cudaMemcpyAsync(d_off, pin + 3 * pb, (size_t)p * 4, cudaMemcpyHostToDevice, up));
      CU(cudaMemcpyAsync(d_alpha, pin + 4 * pb, pb, cudaMemcpyHostToDevice, up));
    }
    CU(cudaEventRecord(h->pin_ev[pslot], up));
    h->pin_used[pslot] = true;
  }
  CU(cudaMemsetAsync(h->d_flags, 0, sizeof(int) * 4, st));
  CU(cudaMemsetAsync(h->d_scal, 0, sizeof(double) * kNumSlots, st));
  // ---- host streaming resources ------------------------------------------------------------------------
  if (streaming) {
    for (int i = 0; i < 2; ++i) CU(h->hx[i].reserve((size_t)slab * P * sizeof(double)));
    CU(h->hr.reserve((size_t)N * sizeof(double)));
    CU(cudaMemcpyAsync(h->hr.p, in.r_host, sizeof(double) * N, cudaMemcpyHostToDevice, up));
    CU(cudaEventRecord(h->ev_small, up));            // the compute stream starts once the small operands are in
    CU(cudaStreamWaitEvent(st, h->ev_small, 0));
  }
  auto issue_copy = [&](int si) -> cudaError_t {  // H2D of slab si into staging buffer si%2
    const long i0 = (long)si * slab;
    const long ns = std::min(slab, N - i0);
    const int bsel = si & 1;
    cudaError_t e;
    if (si >= 2 && (e = cudaStreamWaitEvent(h->copy_stream, h->ev_used[bsel], 0)) != cudaSuccess) return e;
    if (in.ldx == P)
      e = cudaMemcpyAsync(h->hx[bsel].p, in.X_host + i0 * in.ldx, (size_t)ns * P * sizeof(double),
                          cudaMemcpyHostToDevice, h->copy_stream);
    else
      e = cudaMemcpy2DAsync(h->hx[bsel].p, P * sizeof(double), in.X_host + i0 * in.ldx, in.ldx * sizeof(double),
                            P * sizeof(double), ns, cudaMemcpyHostToDevice, h->copy_stream);
    if (e != cudaSuccess) return e;
    return cudaEventRecord(h->ev_copied[bsel], h->copy_stream);
  };
  const int nslabs = (int)((N + slab - 1) / slab);
  static const bool dbg_t = getenv("SGP_TIMING") != nullptr;
  auto now_ms = []() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  const double t_issue = now_ms();
  if (streaming && !in.first_copy_issued) CU(issue_copy(0));

  if (regather) {
    if (!direct_X && !streaming) {
      if (in.X32) gather_cols_f32(in.X32, in.ldx, N, d_idx, p, (double*)Xg, ldp, st);
      else gather_cols(in.X, in.ldx, N, d_idx, p, (double*)Xg, ldp, st);
    }
    if (!direct_U) {
      if (in.U32) gather_cols_f32(in.U32, in.ldu, m, d_idx, p, (double*)Ug, ldp, st);
      else gather_cols(in.U, in.ldu, m, d_idx, p, (double*)Ug, ldp, st);
    }
    h->cache_mask = support; h->cache_X = in.X; h->cache_N = N; h->cache_ldx = in.ldx;
    h->cache_U = in.U; h->cache_m = m; h->cache_ldu = in.ldu; h->cache_token = h->image_token;
  }
  (void)cache_was_valid;
  copy_mat(in.B, in.ldb, m, q, Bw, ldq, st);
  copy_mat(in.V, in.ldv, q, q, Vw, ldq, st);
  // log det V_b does not depend on theta: its Cholesky runs on a side stream next to the K_tilde_b chain.  Only the
  // fork is recorded here; the ~25 launches are ENQUEUED after the main stream's first kernels (enqueue_side below):
  // the host needs ~3 us per launch, and everything queued ahead of build_C delays the whole evaluation.
  bool side_pending = false;
  if (!in.logdetV_host) {
    if (!h->side_stream) {
      CU(cudaStreamCreateWithFlags(&h->side_stream, cudaStreamNonBlocking));
      CU(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
    }
    CU(cudaEventRecord(h->ev_fork, st));
    side_pending = true;
  }
  auto enqueue_side = [&]() -> int {
    if (!side_pending) return SGP_OK;
    CU(cudaStreamWaitEvent(h->side_stream, h->ev_fork, 0));
    double* Lv_side = qq[9];
    copy_mat(Vw, ldq, q, q, Lv_side, ldq, h->side_stream);
    CU(factor_any(h, Lv_side, ldq, q, nullptr, nullptr, winv2, kFlagVNotPD, kFlagVSmallPivot, kSlotLogdetV, h->side_stream));
    CU(cudaEventRecord(h->ev_join, h->side_stream));
    return SGP_OK;
  };

  // ---- a1: C(theta) ---------------------------------------------------------------------------------
  // Separable path: nothing in front of the tail needs C itself (the image pass takes its quadratic forms and U C from the
  // Kronecker structure), so C is built at the head of the inducing-point chain, off the critical path.
  const bool sep_uc = use_sep && h->sep_uc != 0;
  if (!sep_uc) build_C(C, ldp, p, d_gx, d_gy, th, nullptr, 0, st);

  const bool smallU = true;  // inducing-point GEMMs are small: 64x64 tiles
  // ---- inducing-point path: K_tilde, K_tilde_b, its inverse, and the moment matrices ------------------
  {
    // ZU = Ug C ; partial row dots for t_j = u_j^T C u_j.  On the critical path of every evaluation (nothing else can
    // start before it): int8-slice engine when it is on (C is symmetric, so its rows are the B operand's rows)
    if (sep_uc) {
      // ZU = U C, t_j = u_j^T C u_j + sigma_0^2, X2 = sqrt(t): one launch (sep_quad.cu) instead of five dependent ones
      CU(sep_apply_C(Ug, ldug, m, p, d_off, d_alpha, n, th.er, th.amp, th.s02, ZU, ldp, tU, X2, st));
    } else if (ozS) {
      CU(oz_slice_rows(Ug, ldug, m, p, ozS, ugq, ldkP, uge, st));
      CU(oz_slice_rows(C, ldp, p, p, ozS, cq, ldkP, ce, st));
      const OzMoments mom{Ug, ldug, partU, ldpartU};
      CU(oz_gemm(ugq, uge, m, cq, ce, p, p, ldkP, ozS, ZU, ldp, 1, 0, 0, st, &mom));
      norms_finalize(partU, ldpartU, (p + 15) / 16, m, th.s02, tU, X2, st);
    } else {
      GemmArgs g{};
      g.A = Ug; g.lda = ldug; g.B = C; g.ldb = ldp; g.C = ZU; g.ldc = ldp; g.M = m; g.N = p; g.K = p;
      g.aux = Ug; g.ldaux = ldug; g.part1 = partU; g.ldpart = ldpartU;
      CU(gemm(st, 0, 1, kRowdot, smallU, 1, g));
      norms_finalize(partU, ldpartU, ntp_s, m, th.s02, tU, X2, st);
    }
    // Everything below only feeds the moments (Mfull, bb) and the KL scalars: a chain of ~100 small launches that
    // is latency-bound on a handful of SMs.  It runs on its own stream while the main stream starts the image pass
    // (quadratic forms of the first slab on a slightly smaller grid); they join before the first moments product.
    if (chain_overlap) {
      if (!h->chain_stream) {
        CU(cudaStreamCreateWithFlags(&h->chain_stream, cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&h->ev_chain_fork, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&h->ev_chain_join, cudaEventDisableTiming));
      }
      CU(cudaEventRecord(h->ev_chain_fork, st));
    }
  }
  // The chain itself (~100 launches, ~0.3 ms of host enqueue time).  With chain_overlap it is enqueued from inside the
  // image loop, AFTER the first slab's quadratic forms / K_lin product / arc-cosine map have been queued on the main
  // stream (none of them needs it): queued first, it kept the main stream waiting for the host.
  bool chain_enqueued = false;
  auto enqueue_chain = [&]() -> int {
    if (chain_enqueued) return SGP_OK;
    chain_enqueued = true;
    cudaStream_t st_main = st;
    cudaStream_t st = chain_overlap ? h->chain_stream : st_main;   // shadows the outer stream inside this lambda
    if (chain_overlap) CU(cudaStreamWaitEvent(h->chain_stream, h->ev_chain_fork, 0));
    if (sep_uc) build_C(C, ldp, p, d_gx, d_gy, th, nullptr, 0, st);     // for the tail's <dC_k, G_C> contraction only
    GemmArgs a{};  // K_tilde (raw) = acos-kernel(ZU Ug^T)
    a.A = ZU; a.lda = ldp; a.B = Ug; a.ldb = ldug; a.C = Kraw; a.ldc = ldm; a.M = m; a.N = m; a.K = p;
    a.rowvec = X2; a.colvec = X2; a.C2 = cosU; a.s02 = th.s02;
    CU(gemm(st, 0, 0, kAcos, smallU, 1, a));
    symmetrize(Kraw, ldm, m, Kt, ldm, st);                                              // U:1024-1025
    CU(mm(st, 0, 1, Kt, ldm, Bw, ldq, tmq1, ldq, m, q, m));                              // Kt B
    CU(mm(st, 1, 1, Bw, ldq, tmq1, ldq, Ktb_raw, ldq, q, q, m));                         // B^T Kt B   U:2047
    symmetrize(Ktb_raw, ldq, q, Ktb, ldq, st);                                          // U:2048
    copy_mat(Ktb, ldq, q, q, Lk, ldq, st);
    CU(factor_any(h, Lk, ldq, q, Linv, tri_tmp, winv, kFlagKtNotPD, kFlagKtSmallPivot, kSlotLogdetK, st));
    CU(mm(st, 1, 1, Linv, ldq, Linv, ldq, Kinv_raw, ldq, q, q, q));                      // K_tilde_b^-1  (U:2067)
    symmetrize(Kinv_raw, ldq, q, Kinv, ldq, st);
    gemv(Kinv, ldq, q, q, 0, in.m_b, bq, st);                                           // b = Kinv m_b
    gemv(Bw, ldq, m, q, 0, bq, bb, st);                                                 // bb = B b
    CU(mm(st, 0, 1, Kinv, ldq, Vw, ldq, KV, ldq, q, q, q));                              // Kinv V
    CU(mm(st, 0, 1, KV, ldq, Kinv, ldq, KVK, ldq, q, q, q));                             // Kinv V Kinv
    lincomb(Mg, ldq, q, q, 2.0, KVK, ldq, -2.0, Kinv, ldq, 0.0, st);                    // Mg = 2 KVK - 2 Kinv
    CU(mm(st, 0, 1, Bw, ldq, Mg, ldq, tmq2, ldq, m, q, q));                              // B Mg
    CU(mm(st, 0, 0, tmq2, ldq, Bw, ldq, Mfull, ldm, m, m, q));                           // Mfull = B Mg B^T
    kl_scalars(in.m_b, bq, Vw, ldq, Kinv, ldq, q, h->d_scal, kSlotMb, kSlotTrVK, qq[12], st);   // qq[12]: free until the tail
    if (proj) {
      CU(oz_slice_rows(Mg, ldq, q, q, ozM, mgq, oz_ldk(q), mge, st));                    // B operand of T_b = K_b Mg^T
      CU(oz_slice_rows(Bw, ldq, m, q, ozM, brq, oz_ldk(q), bre, st));                    // B operand of T = T_b B^T
      CU(oz_slice_cols(Bw, ldq, m, q, ozM, btq, oz_ldk(m), bte, ozmx, st));              // B operand of K_b = K B
    } else if (ozM) CU(oz_slice_rows(Mfull, ldm, m, m, ozM, mq, oz_ldk(m), me, st));
    if (ozS) CU(oz_slice_cols(Ug, ldug, m, p, ozS, utq, oz_ldk(m), ute, ozmx, st));     // B operand of GC = D Ug (tail)
    if (chain_overlap) CU(cudaEventRecord(h->ev_chain_join, st));
    return SGP_OK;
  };
  if (!chain_overlap) { rc = enqueue_chain(); if (rc) return rc; rc = enqueue_side(); if (rc) return rc; }

  // ---- split-K plans for the two reductions over images ----------------------------------------------
  const int nzGC = use_sep ? 1 : pick_split(gemm_tiles(p, p), slab), nzS2 = pick_split(gemm_tiles(m, m, true), slab);
  const int nzD = pick_split(gemm_tiles(p, m), slab);
  const long zGC = (long)p * ldp, zS2 = (long)m * ldm, zD = (long)p * ldm;
  if (use_sep) CU(h->dpart.reserve((size_t)std::max(nzD, nzO) * zD * sizeof(double)));
  // persistent int8 slices of the images: per slab, [pairs][rows][ldkP] for X and [pairs][p][ldkN] for X^T.
  // Device-resident images are sliced once and reused while (X, N, ld, mask, slab, slice count) are unchanged --
  // sgp_invalidate_cache() after modifying the images in place; host-streamed slabs are sliced every time.
  const long nslabs_all = (N + slab - 1) / slab;
  bool oz_reslice = true;
  if (ozS) {
    const long keep = streaming ? 1 : nslabs_all;
    CU(h->xq.reserve(streaming ? oz_slice_bytes(slab, p, ozI) : oz_slice_bytes(N, p, ozI)));
    CU(h->xe.reserve((size_t)(streaming ? slab : N) * sizeof(int)));
    CU(h->xtq.reserve((size_t)keep * oz_pairsI * p * ldkN));
    CU(h->xte.reserve((size_t)keep * p * sizeof(int)));
    oz_reslice = streaming || regather || !oz_was_valid || h->oz_S != ozI || h->oz_slab != slab;
    h->oz_S = ozI; h->oz_slab = slab;
    CU(oz_slice_rows(ZU, ldp, m, p, ozI, zuq, ldkP, zue, st));      // slices of U C (this evaluation's theta)
  }
  double* Dpart = (double*)h->dpart.p;
  CU(h->gcpart.reserve((size_t)nzGC * zGC * sizeof(double)));
  CU(h->s2part.reserve((size_t)std::max(nzS2, nzK) * zS2 * sizeof(double)));
  double* GCpart = (double*)h->gcpart.p; double* S2part = (double*)h->s2part.p;

  const double* rdev = streaming ? (const double*)h->hr.p : in.r;

  // ---- image pass, slab by slab ---------------------------------------------------------------------------
  int slab_i = 0;
  for (long i0 = 0; i0 < N; i0 += slab, ++slab_i) {
    const long ns = std::min(slab, N - i0);
    const int acc_flag = slab_i > 0 ? 1 : 0;
    const double* Xs; long ldxs;
    if (streaming) {
      if (slab_i + 1 < nslabs) CU(issue_copy(slab_i + 1));
      const int bsel = slab_i & 1;
      CU(cudaStreamWaitEvent(st, h->ev_copied[bsel], 0));
      if (full_mask && P % 2 == 0) { Xs = (const double*)h->hx[bsel].p; ldxs = P; }
      else { gather_cols((const double*)h->hx[bsel].p, P, ns, d_idx, p, Xslab_stream, ldp, st); Xs = Xslab_stream; ldxs = ldp; }
    } else {
      Xs = Xg + i0 * ldxg; ldxs = ldxg;
    }
    mark(1);
    GemmArgs g2{};
    if (use_sep) {
      // s_i = x_i^T C x_i + sigma_0^2 and the five x_i^T dC_k x_i from the Kronecker structure of C (sep_quad.cu)
      SepQuadArgs sq{};
      sq.X = Xs; sq.ldx = ldxs; sq.nrows = ns; sq.p = p; sq.off = d_off; sq.alpha = d_alpha; sq.n = n;
      sq.er = th.er; sq.eb = th.eb; sq.ex = th.ex; sq.ey = th.ey; sq.amp = th.amp; sq.s02 = th.s02;
      sq.s = s_s; sq.nrm = X1; sq.Q = Qs; sq.full = (p_true == (int)P) ? 1 : 0;
      if (chain_overlap && slab_i == 0) sq.max_ctas = num_sms() - 20;   // leave SMs to the inducing-point chain
      CU(sep_quad(sq, st));
      mark(2);
      // G2: K = acos-kernel(X (U C)^T)   -- x_i^T C u_j with the m x p matrix ZU = U C
      g2.A = Xs; g2.lda = ldxs; g2.B = ZU; g2.ldb = ldp;
    } else {
      // G1: Z = X C, partial s_i = x_i^T C x_i                                  (U:978, U:1029)
      GemmArgs g1{};
      g1.A = Xs; g1.lda = ldxs; g1.B = C; g1.ldb = ldp; g1.C = Zs; g1.ldc = ldp; g1.M = (int)ns; g1.N = p; g1.K = p;
      g1.aux = Xs; g1.ldaux = ldxs; g1.part1 = partN; g1.ldpart = ldpartN;
      CU(gemm(st, 0, 1, kRowdot, false, 1, g1));
      mark(2);
      norms_finalize(partN, ldpartN, ntp, ns, th.s02, s_s, X1, st);
      // G2: K = acos-kernel(Z Ug^T)                                             (U:982-990)
      g2.A = Zs; g2.lda = ldp; g2.B = Ug; g2.ldb = ldug;
    }
    g2.C = Ks; g2.ldc = ldm; g2.M = (int)ns; g2.N = m; g2.K = p;
    const long oz_si = streaming ? 0 : slab_i;                     // which persistent slice block this slab uses
    int8_t* xq_s = nullptr; int* xe_s = nullptr; int8_t* xtq_s = nullptr; int* xte_s = nullptr;
    if (ozS) {
      xq_s = (int8_t*)h->xq.p + oz_slice_bytes(oz_si * slab, p, ozI);
      xe_s = (int*)h->xe.p + oz_si * slab;
      xtq_s = (int8_t*)h->xtq.p + (size_t)oz_si * oz_pairsI * p * ldkN;
      xte_s = (int*)h->xte.p + oz_si * p;
      if (oz_reslice) CU(oz_slice_rows(Xs, ldxs, ns, p, ozI, xq_s, ldkP, xe_s, st));
      // The persistent product leaves no room for the chain's kernels on an SM it occupies, so when the chain is longer
      // than the quadratic forms it hides behind (large q: its q x q x m GEMMs grow as q^2 m; or a short first slab) the
      // product of the first slab leaves the same 20 SMs free: +15 % on this launch instead of a stall of the whole
      // main stream at the join.  Estimates from the measured config-2 times (chain 1.3 ms at q = m = 500, quadratic
      // forms 1.3 ms per 65 536 images of 40 x 40).  m = 1000, N = 50 000: 11.06 -> 10.15 ms per evaluation.
      // SGP_CHAIN_XUC=0/1 overrides (A/B measurements).
      static const int env_xuc = getenv("SGP_CHAIN_XUC") ? atoi(getenv("SGP_CHAIN_XUC")) : -1;
      bool reserve = false;
      if (chain_overlap && slab_i == 0) {
        const double chain_ms = 1.3 * (q / 500.0) * (q / 500.0) * (m / 500.0);
        const double quad_ms = 1.3 * ((double)ns / 65536.0) * (n / 40.0) * (n / 40.0) * (n / 40.0);
        reserve = env_xuc >= 0 ? env_xuc != 0 : chain_ms > 1.15 * quad_ms;
      }
      CU(oz_gemm(xq_s, xe_s, ns, zuq, zue, m, p, ldkP, ozI, Ks, ldm, 1, 0, 0, st, nullptr, reserve ? num_sms() - 20 : 0));
    } else {
      CU(gemm(st, 0, 0, kStore, false, 1, g2));
    }
    mark(9);
    acos_forward(Ks, cos_s, ldm, ns, m, X1, X2, th.s02, st);
    mark(3);
    if (chain_overlap && slab_i == 0) {
      rc = enqueue_chain(); if (rc) return rc;
      rc = enqueue_side(); if (rc) return rc;
      CU(cudaStreamWaitEvent(st, h->ev_chain_join, 0));
    }
    // G3: T = K Mfull^T ; partial lam_var / lam_m row sums                    (U:1090-1101)
    GemmArgs g3{};
    g3.A = Ks; g3.lda = ldm; g3.B = Mfull; g3.ldb = ldm; g3.C = Ts; g3.ldc = ldm; g3.M = (int)ns; g3.N = m; g3.K = m;
    g3.aux = Ks; g3.ldaux = ldm; g3.colvec = bb; g3.part1 = pm1; g3.part2 = pm2; g3.ldpart = ldpartN;
    if (proj) {
      CU(oz_slice_rows(Ks, ldm, ns, m, ozM, kq, oz_ldk(m), ke, st));
      CU(oz_gemm(kq, ke, ns, btq, bte, q, m, oz_ldk(m), ozM, Kb, ldq, 1, 0, 0, st));          // K_b = K B   (U:2049)
      CU(oz_slice_rows(Kb, ldq, ns, q, ozM, kbq, oz_ldk(q), kbe, st, bq, pm2));              // pm2 = K_b b = lam_m
      const OzMoments mom{Kb, ldq, pm1, ldpartN};
      CU(oz_gemm(kbq, kbe, ns, mgq, mge, q, q, oz_ldk(q), ozM, Tb, ldq, 1, 0, 0, st, &mom));  // T_b = K_b Mg^T + row sums
      CU(oz_slice_rows(Tb, ldq, ns, q, ozM, tbq, oz_ldk(q), tbe, st));
      CU(oz_gemm(tbq, tbe, ns, brq, bre, m, q, oz_ldk(q), ozM, Ts, ldm, 1, 0, 0, st));        // T = T_b B^T
    } else if (ozM) {
      CU(oz_slice_rows(Ks, ldm, ns, m, ozM, kq, oz_ldk(m), ke, st, bb, pm2));   // pm2 = K bb on the way
      const OzMoments mom{Ks, ldm, pm1, ldpartN};
      CU(oz_gemm(kq, ke, ns, mq, me, m, m, oz_ldk(m), ozM, Ts, ldm, 1, 0, 0, st, &mom));
    } else {
      CU(gemm(st, 0, 0, kMoments, false, 1, g3));
    }
    mark(4);
    moments_rows(s_s, pm1, pm2, ldpartN, proj ? (q + 15) / 16 : (ozM ? ng16 : ntm), ozM ? 1 : ntm, rdev + i0, ns, A, in.lambda0, lam_m + i0, lam_var + i0, fvec + i0, u_s,
                 v_s, blockpart, st);
    const int nbm = moments_rows_blocks(ns);
    static_assert(kSlotSumF == kSlotRLam + 1 && kSlotSumR == kSlotRLam + 2 && kSlotNaN == kSlotRLam + 3, "slot order");
    sum_to_slots(blockpart, nbm, 4, 4, h->d_scal, kSlotRLam, acc_flag, st);
    // adjoint of the arc-cosine map: R (in place over T), row weights w, column sums kappa / y
    // ... and, for the int8-slice products that follow, the column maxima of R and of diag(sqrt(-v)) K
    acos_backward(0, Ts, cos_s, Ks, ldm, ns, m, u_s, v_s, bb, X1, X2, w_s, colk, coly, ldm, sumpartR, st,
                  (ozS || (ozM && !proj)) ? cmaxR : nullptr, cmaxW);
    const int nbb = acos_backward_blocks(ns);
    col_reduce(colk, nbb, m, ldm, kappa, acc_flag, st);
    col_reduce(coly, nbb, m, ldm, yv, acc_flag, st);
    sum_to_slot(sumpartR, nbb, 1, 1.0, h->d_scal, kSlotSig, 1, st);
    sum_to_slot(w_s, ns, 1, 1.0, h->d_scal, kSlotSig, 1, st);
    // G6: Sig2 += K^T diag(v) K
    auto sig2_product = [&]() -> cudaError_t {
      if (proj) {     // S2b += K_b^T diag(v) K_b directly (q x q): no B^T Sig2 B in the tail
        const int ldk_s = oz_ldk((int)ns);
        cudaError_t e = oz_slice_cols(Kb, ldq, ns, q, ozM, kbtq, ldk_s, kbte, ozmx, st, v_s);
        if (e != cudaSuccess) return e;
        return oz_gemm(kbtq, kbte, q, kbtq, kbte, q, (int)ns, ldk_s, ozM, S2part, ldq, nzK, (long)q * ldq, acc_flag, st,
                       nullptr, 0, -1.0, 1);
      }
      if (ozM) {
        // v_i = -A^2 f_i / 2 <= 0, so K^T diag(v) K = -W^T W with W = diag(sqrt(-v)) K: ONE transposed slicing serves
        // both operands of the int8 product; only the tiles on and below the diagonal are computed
        const int ldk_s = oz_ldk((int)ns);
        oz_exps_from_colmax(cmaxW, acos_backward_blocks(ns), m, ldm, 1, rte, st);
        cudaError_t e = oz_slice_cols(Ks, ldm, ns, m, ozW, rtq, ldk_s, rte, ozmx, st, v_s, true);
        if (e != cudaSuccess) return e;
        return oz_gemm(rtq, rte, m, rtq, rte, m, (int)ns, ldk_s, ozW, S2part, ldm, nzK, zS2, acc_flag, st, nullptr, 0, -1.0, 1);
      }
      GemmArgs g6{};
      g6.A = Ks; g6.lda = ldm; g6.B = Ks; g6.ldb = ldm; g6.C = S2part; g6.ldc = ldm; g6.M = m; g6.N = m; g6.K = (int)ns;
      g6.k_chunk = (int)rup((ns + nzS2 - 1) / nzS2, 16); g6.zstride = zS2; g6.accumulate = acc_flag; g6.kscale = v_s;
      g6.lower_only = 1;   // K^T diag(v) K is symmetric: only the CTA tiles on and below the diagonal are computed
      return gemm(st, 1, 1, kStore, false, nzS2, g6);
    };
    mark(5);
    if (use_sep) {
      // sum_i w_i x_i^T dC_k x_i from the stored quadratic forms; D += X^T R (split-K over the images of the slab)
      wq_reduce(w_s, Qs, ns, wqpart, st);
      sum_to_slots(wqpart, wq_reduce_blocks(ns), 5, 5, h->d_scal, kSlotGW, acc_flag, st);
      mark(7);
      // G6: Sig2 += K^T diag(v) K (before X^T R: the q x q tail depends on it, not on D)
      CU(sig2_product());
      const bool last = i0 + ns >= N;
      if (tail_overlap && last) {
        // the q x q tail and the inducing-point backward pass start now on the chain stream; this slab's X^T R
        // (which they do not need) runs next to them on a slightly smaller grid
        CU(cudaEventRecord(h->ev_chain_fork, st));
        CU(cudaStreamWaitEvent(h->chain_stream, h->ev_chain_fork, 0));
      }
      mark(6);
      if (ozS) {
        // the contraction runs over the images: slices of X^T (persistent) and of R^T (this evaluation)
        const int ldk_s = oz_ldk((int)ns);
        if (oz_reslice) CU(oz_slice_cols(Xs, ldxs, ns, p, ozI, xtq_s, ldk_s, xte_s, ozmx, st));
        oz_exps_from_colmax(cmaxR, acos_backward_blocks(ns), m, ldm, 0, rte, st);
        CU(oz_slice_cols(Ts, ldm, ns, m, ozX, rtq, ldk_s, rte, ozmx, st, nullptr, true));
        mark(10);
        CU(oz_gemm(xtq_s, xte_s, p, rtq, rte, m, (int)ns, ldk_s, ozX, Dpart, ldm, nzO, zD, acc_flag, st, nullptr,
                   (tail_overlap && last) ? num_sms() - 20 : 0));
      } else {
        GemmArgs g5{};
        g5.A = Xs; g5.lda = ldxs; g5.B = Ts; g5.ldb = ldm; g5.C = Dpart; g5.ldc = ldm; g5.M = p; g5.N = m; g5.K = (int)ns;
        g5.k_chunk = (int)rup((ns + nzD - 1) / nzD, 16); g5.zstride = zD; g5.accumulate = acc_flag;
        CU(gemm(st, 1, 1, kStore, false, nzD, g5));
      }
    } else {
      // G4: Y = R Ug + diag(w) X   (into the Z buffer)
      GemmArgs g4{};
      g4.A = Ts; g4.lda = ldm; g4.B = Ug; g4.ldb = ldug; g4.C = Zs; g4.ldc = ldp; g4.M = (int)ns; g4.N = p; g4.K = m;
      g4.rowvec = w_s; g4.aux = Xs; g4.ldaux = ldxs;
      CU(gemm(st, 0, 1, kAxpyRow, false, 1, g4));
      mark(6);
      // G5: GC += X^T Y   (split-K over the images of the slab)
      GemmArgs g5{};
      g5.A = Xs; g5.lda = ldxs; g5.B = Zs; g5.ldb = ldp; g5.C = GCpart; g5.ldc = ldp; g5.M = p; g5.N = p; g5.K = (int)ns;
      g5.k_chunk = (int)rup((ns + nzGC - 1) / nzGC, 16); g5.zstride = zGC; g5.accumulate = acc_flag;
      CU(gemm(st, 1, 1, kStore, false, nzGC, g5));
      mark(7);
      CU(sig2_product());
    }
    mark(8);
    if (streaming) CU(cudaEventRecord(h->ev_used[slab_i & 1], st));
  }

  // ---- adjoint of K_tilde_b (q x q algebra) ---------------------------------------------------------------
  cudaStream_t st_loop = st;
  if (tail_overlap) st = h->chain_stream;     // issued after the fork above; joins before the final contraction
  if (proj) {
    slice_reduce(S2part, oz_eff_split((int)std::min(slab, N), nzK), (long)q * ldq, (long)q * ldq, S2b, 0, st);
    mirror_lower(S2b, ldq, q, st);
  } else {
    slice_reduce(S2part, ozM ? oz_eff_split((int)std::min(slab, N), nzK) : nzS2, zS2, zS2, Sig2, 0, st);
    mirror_lower(Sig2, ldm, m, st);
  }
  col_weights(kappa, nullptr, nullptr, X2, m, beta, st);                                 // beta_j = kappa_j / (2 X2_j)
  sum_to_slot(beta, m, 1, 1.0, h->d_scal, kSlotSig, 1, st);
  gemv(Bw, ldq, m, q, 1, yv, y_b, st);                                                  // y_b = B^T y
  if (!proj) {
    CU(mm(st, 0, 1, Sig2, ldm, Bw, ldq, tmq1, ldq, m, q, m));
    CU(mm(st, 1, 1, Bw, ldq, tmq1, ldq, S2b, ldq, q, q, m));                             // B^T Sig2 B
  }
  CU(mm(st, 0, 0, Kinv, ldq, Vw, ldq, KV, ldq, q, q, q));                                // Kinv V^T
  lincomb(W2, ldq, q, q, 2.0, KV, ldq, 0.0, nullptr, 0, -1.0, st);                      // 2 Kinv V^T - I
  CU(mm(st, 0, 1, S2b, ldq, W2, ldq, inner, ldq, q, q, q));
  rank1_add(inner, ldq, q, q, 1.0, y_b, in.m_b, st);                                    // + y_b m_b^T
  CU(mm(st, 0, 1, Kinv, ldq, inner, ldq, Sq, ldq, q, q, q));                             // S = Kinv (...)
  CU(mm(st, 0, 1, Sq, ldq, Kinv, ldq, GKtb, ldq, q, q, q, -1.0));                        // -S Kinv
  gkt_assemble(GKtb, ldq, q, Kinv, KVK, bq, h->kl_weight, st);                                        // KL part (U:1333)
  CU(mm(st, 0, 1, Bw, ldq, GKtb, ldq, tmq3, ldq, m, q, q));
  CU(mm(st, 0, 0, tmq3, ldq, Bw, ldq, GKt, ldm, m, m, q));                               // GKt = B GKtb B^T

  // ---- inducing-point path, backward -----------------------------------------------------------------------
  acos_backward(1, GKt, cosU, nullptr, ldm, m, m, nullptr, nullptr, nullptr, X2, X2, wr, colkU, nullptr, ldm,
                sumpartU, st);
  col_reduce(colkU, nblk_bwdU, m, ldm, kappaU, 0, st);
  col_weights(kappaU, nullptr, nullptr, X2, m, wc, st);
  sum_to_slot(sumpartU, nblk_bwdU, 1, 1.0, h->d_scal, kSlotSig, 1, st);
  sum_to_slot(wr, m, 1, 1.0, h->d_scal, kSlotSig, 1, st);
  sum_to_slot(wc, m, 1, 1.0, h->d_scal, kSlotSig, 1, st);
  col_weights(nullptr, nullptr, wr, X2, m, wU, st);                                      // wU = wr
  lincomb(wU, ldm, 1, m, 1.0, wU, ldm, 1.0, wc, ldm, 0.0, st);                          // + wc
  lincomb(wU, ldm, 1, m, 1.0, wU, ldm, 1.0, beta, ldm, 0.0, st);                        // + beta
  {
    GemmArgs g{};  // YU = RU Ug + diag(wU) Ug
    g.A = GKt; g.lda = ldm; g.B = Ug; g.ldb = ldug; g.C = ZU; g.ldc = ldp; g.M = m; g.N = p; g.K = m;
    g.rowvec = wU; g.aux = Ug; g.ldaux = ldug;
    CU(gemm(st, 0, 1, kAxpyRow, smallU, 1, g));
    // Ug^T YU enters only through <dC_k, .> and every dC_k is symmetric, so <dC_k, Ug^T YU> = <dC_k, YU^T Ug>: with the
    // int8 engine on, YU^T is added to D = X^T R and ONE product (D + YU^T) Ug serves both terms
    if (!ozS) CU(mm(st, 1, 1, Ug, ldug, ZU, ldp, GCU, ldp, p, p, m));                    // Ug^T YU
  }
  if (tail_overlap) {
    CU(cudaEventRecord(h->ev_chain_join, st));
    st = st_loop;
  }

  // ---- <dC_k, GC> and result assembly -----------------------------------------------------------------------
  if (use_sep) {
    slice_reduce(Dpart, ozS ? oz_eff_split((int)std::min(slab, N), nzO) : nzD, zD, zD, Dm, 0, st);
    if (ozS) {                                                                            // (X^T R + YU^T) U
      if (tail_overlap) CU(cudaStreamWaitEvent(st, h->ev_chain_join, 0));
      add_transposed(Dm, ldm, p, m, ZU, ldp, st);
      CU(oz_slice_rows(Dm, ldm, p, m, ozS, dq, oz_ldk(m), de, st));
      CU(oz_gemm(dq, de, p, utq, ute, p, m, oz_ldk(m), ozS, GCpart, ldp, 1, 0, 0, st));
    } else {
      CU(mm(st, 0, 1, Dm, ldm, Ug, ldug, GCpart, ldp, p, p, m));
    }
  }
  if (tail_overlap) CU(cudaStreamWaitEvent(st, h->ev_chain_join, 0));
  contract_dC(GCpart, nzGC, zGC, ozS ? nullptr : GCU, C, ldp, p, d_gx, d_gy, th, gpart, st);
  sum_to_slots(gpart, p, 5, 5, h->d_scal, kSlotG, 0, st);
  if (use_sep) sum_to_slots(h->d_scal + kSlotGW, 1, 5, 5, h->d_scal, kSlotG, 1, st);
  if (side_pending) CU(cudaStreamWaitEvent(st, h->ev_join, 0));
  const bool async_out = in.out_dev != nullptr;
  if (h->comm_on && in.exchange) {
    h->comm.epoch += 1;           // every rank issues its evaluations in the same order: the epochs agree
    finalize_exchange(h->d_scal, h->d_flags, A, in.lambda0, th, p_true, in.logdetV_host ? *in.logdetV_host : 0.0,
                      in.logdetV_host ? 1 : 0, async_out ? in.out_dev : h->d_out, proj ? 1.0 : 0.0, h->comm, st);
  } else {
    finalize(h->d_scal, h->d_flags, A, in.lambda0, th, p_true, in.logdetV_host ? *in.logdetV_host : 0.0,
             in.logdetV_host ? 1 : 0, async_out ? in.out_dev : h->d_out, proj ? 1.0 : 0.0, st);
  }
  CU(cudaGetLastError());
  if (!async_out) CU(cudaMemcpyAsync(h->h_out, h->d_out, sizeof(double) * 16, cudaMemcpyDeviceToHost, st));
  mark(-1);
  const double t_enq = now_ms();
  // asynchronous calls return here: everything is enqueued in stream order, the caches are valid for later work on this
  // stream.  (With option "profile" the stage events have to be read, so profiled calls always wait.)
  if (!async_out || h->profile) CU(cudaStreamSynchronize(st));
  h->cache_valid = !streaming;           // everything the caches hold is produced by work enqueued (or completed) on st
  h->oz_valid = !streaming && ozS > 0;
  if (dbg_t) fprintf(stderr, "[sgp timing] first copy issued %.3f ms after entry; all work enqueued at %.3f; done at %.3f\n",
                     t_issue - t_entry, t_enq - t_entry, now_ms() - t_entry);
  if (out_host && !async_out) std::memcpy(out_host, h->h_out, sizeof(double) * 16);
  if (h->profile) {
    for (size_t i = 0; i + 1 < h->prof_used; ++i) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, h->prof_ev[i], h->prof_ev[i + 1]);
      const int sg = h->prof_stage[i];
      if (sg >= 0 && sg < 16) h->stage_ms[sg] += ms;
    }
    h->stage_calls++;
  }

  h->dbg.clear();
  h->dbg["C"] = {C, p, p, ldp};   h->dbg["Kt"] = {Kt, m, m, ldm};     h->dbg["Ktb"] = {Ktb, q, q, ldq};
  h->dbg["Kinv"] = {Kinv, q, q, ldq}; h->dbg["Mfull"] = {Mfull, m, m, ldm}; h->dbg["GKt"] = {GKt, m, m, ldm};
  h->dbg["GCU"] = {GCU, p, p, ldp};   h->dbg["Sig2"] = {Sig2, m, m, ldm};   h->dbg["bb"] = {bb, 1, m, ldm};
  h->dbg["y"] = {yv, 1, m, ldm};      h->dbg["beta"] = {beta, 1, m, ldm};   h->dbg["X2"] = {X2, 1, m, ldm};
  h->dbg["GCpart"] = {GCpart, (long)nzGC * p, p, ldp}; h->dbg["scal"] = {h->d_scal, 1, kNumSlots, kNumSlots};
  h->dbg["K_last"] = {Ks, std::min(slab, N - (long)(nslabs - 1) * slab), m, ldm};
  h->dbg["Y_last"] = {Zs, std::min(slab, N - (long)(nslabs - 1) * slab), p, ldp};
  h->dbg["R_last"] = {Ts, std::min(slab, N - (long)(nslabs - 1) * slab), m, ldm};
  h->dbg["w_last"] = {w_s, 1, std::min(slab, N - (long)(nslabs - 1) * slab), slab};
  h->dbg["KtRaw"] = {Kraw, m, m, ldm}; h->dbg["Linv"] = {Linv, q, q, ldq}; h->dbg["GKtb"] = {GKtb, q, q, ldq};
  return SGP_OK;
}

// =====================================================================================================
// ntilde == nt: the closure as the reference evaluates it when every training image is an inducing image
// (U:2032 `K = K_tilde`, U:2068 `a = B`; one_cell_active_training.ipynb cell 17 always runs this way).
//   K = K_tilde (symmetrised, from xtilde), K_b = K B, a = B (frozen), Kvec_i = x_i^T C x_i + sigma_0^2 (from x),
//   lam_m = B m_b, lam_var = Kvec + rowsum(a o (a V_b) - K_b o a), f, loglik, KL as in the general case.
// The gradient is the reference's FORMULA (lambda_moments, U:1113-1120, applied with a = B):
//   da = (dK_b - a dK_tilde_b) K_tilde_b^-1, dlam_m = da m_b,
//   dlam_var = dKvec + 2 rowdot(da, a V_b) - rowdot(dK_b, a) - rowdot(K_b, da)
// (da vanishes only when q = m).  Reverse mode of exactly that, with u = A(r - f), v = -A^2 f / 2:
//   Ga = u m_b^T + v o (2 a V_b - K_b)                      (adjoint of da)
//   H = Ga K_tilde_b^-1;   Gq = -a^T H - kl_weight * dKL/dK_tilde_b       (adjoint of dK_tilde_b)
//   G_Ktilde = (H - diag(v) a + B Gq) B^T                   (dK_b = dK B, dK_tilde_b = B^T dK B, and dK = dK_tilde)
//   then the inducing-point backward pass of the general case on G_Ktilde, plus the dKvec term: X^T diag(v) X against
//   dC_k and 2 sigma_0 sum(v).  Everything is m x m / m x p: dense kernels, one stream.
// =====================================================================================================
int run_closure_same(sgp_handle* h, const ClosureIn& in, double* out_host, cudaStream_t st) {
  int rc = ensure_ready(h); if (rc) return rc;
  const int m = in.m, q = in.q, n = in.n_px;
  if (in.N != m) return fail(h, SGP_E_ARG, "run_closure_same: N != m");
  std::vector<uint8_t> mask; std::vector<int> idx; std::vector<double> gx, gy;
  const int p = compute_mask(in.theta, n, mask, idx, gx, gy);       // exactly the reference's x[:, mask]
  if (p == 0) return fail(h, SGP_E_ARG, "sgp_elbo_grad: empty receptive-field mask");
  const ThetaDev th = make_theta(in.theta);
  const double A = std::exp(in.logA);
  const long ldp = rup(p, 16), ldm = rup(m, 16), ldq = rup(q, 16);
  const int ntp_s = (p + 63) / 64, ntq = (q + 127) / 128;
  const long ldpart = ldm;
  const int nblk_bwd = acos_backward_blocks(m), nbm = moments_rows_blocks(m);
  size_t need = 0;
  auto acc = [&](size_t nelem) { need = ((need + 255) & ~size_t(255)) + nelem * sizeof(double); };
  acc(p); acc(p); acc((size_t)(p + 1) / 2 + 8); acc((size_t)p * ldp);
  for (int i = 0; i < 5; ++i) acc((size_t)m * ldp);                   // Xg, Ug, ZU, ZX, YX
  for (int i = 0; i < 4; ++i) acc((size_t)m * ldm);                   // Kraw, Kt, cosU, GKt
  for (int i = 0; i < 8; ++i) acc((size_t)m * ldq);                   // Bw, Kb, aV, Ga, H, BGq, E, spare
  for (int i = 0; i < 10; ++i) acc((size_t)q * ldq);
  acc(64 * ldq); acc(chol_winv_elems(q)); acc(chol_winv_elems(q));
  for (int i = 0; i < 16; ++i) acc(ldm);
  acc((size_t)std::max(ntp_s, ntq) * ldpart * 2);
  acc((size_t)nbm * 4); acc((size_t)nblk_bwd * ldm); acc(nblk_bwd);
  acc((size_t)p * ldp); acc((size_t)p * ldp); acc((size_t)p * 5);
  for (int i = 0; i < 3; ++i) acc(m);
  need += 4096;
  RESERVE_ARENA(need);
  Arena ar(h->arena.p, h->arena.bytes);
  double* d_gx = ar.take<double>(p); double* d_gy = ar.take<double>(p);
  int* d_idx = (int*)ar.take<double>((size_t)(p + 1) / 2 + 8);
  double* C = ar.take<double>((size_t)p * ldp);
  double* Xg = ar.take<double>((size_t)m * ldp); double* Ug = ar.take<double>((size_t)m * ldp);
  double* ZU = ar.take<double>((size_t)m * ldp); double* ZX = ar.take<double>((size_t)m * ldp);
  double* YX = ar.take<double>((size_t)m * ldp);
  double* Kraw = ar.take<double>((size_t)m * ldm); double* Kt = ar.take<double>((size_t)m * ldm);
  double* cosU = ar.take<double>((size_t)m * ldm); double* GKt = ar.take<double>((size_t)m * ldm);
  double* Bw = ar.take<double>((size_t)m * ldq); double* Kb = ar.take<double>((size_t)m * ldq);
  double* aV = ar.take<double>((size_t)m * ldq); double* Ga = ar.take<double>((size_t)m * ldq);
  double* Hm = ar.take<double>((size_t)m * ldq); double* BGq = ar.take<double>((size_t)m * ldq);
  double* Em = ar.take<double>((size_t)m * ldq); (void)ar.take<double>((size_t)m * ldq);
  double* qq[10]; for (int i = 0; i < 10; ++i) qq[i] = ar.take<double>((size_t)q * ldq);
  double* tri_tmp = ar.take<double>(64 * ldq);
  double* winv = ar.take<double>(chol_winv_elems(q)); double* winv2 = ar.take<double>(chol_winv_elems(q));
  double* vecs[16]; for (int i = 0; i < 16; ++i) vecs[i] = ar.take<double>(ldm);
  double* part1 = ar.take<double>((size_t)std::max(ntp_s, ntq) * ldpart);
  double* part2 = ar.take<double>((size_t)std::max(ntp_s, ntq) * ldpart);
  double* blockpart = ar.take<double>((size_t)nbm * 4);
  double* colkU = ar.take<double>((size_t)nblk_bwd * ldm); double* sumpartU = ar.take<double>(nblk_bwd);
  double* GCU = ar.take<double>((size_t)p * ldp); double* GCX = ar.take<double>((size_t)p * ldp);
  double* gpart = ar.take<double>((size_t)p * 5);
  double* lam_m = in.lam_m ? in.lam_m : ar.take<double>(m);
  double* lam_var = in.lam_var ? in.lam_var : ar.take<double>(m);
  double* fvec = in.f ? in.f : ar.take<double>(m);
  double* Vw = qq[0]; double* Ktb_raw = qq[1]; double* Ktb = qq[2]; double* Lk = qq[3]; double* Linv = qq[4];
  double* Kinv_raw = qq[5]; double* Kinv = qq[6]; double* KV = qq[7]; double* KVK = qq[8]; double* Gq = qq[9];
  double* tU = vecs[0]; double* X2 = vecs[1]; double* sX = vecs[2]; double* nX = vecs[3]; double* bq = vecs[4];
  double* lm_t = vecs[5]; double* lv_t = vecs[6]; double* u_s = vecs[7]; double* v_s = vecs[8]; double* wr = vecs[9];
  double* kappaU = vecs[10]; double* wc = vecs[11]; double* wU = vecs[12]; double* Lv = Ktb_raw;   // reused after K_tilde_b

  {   // small uploads through the pinned ring (pageable uploads would synchronise the stream)
    char* pin = nullptr; int pslot = 0;
    const size_t pb = (size_t)p * 8;
    rc = pin_acquire(h, 3 * pb + 64, &pin, &pslot); if (rc) return rc;
    std::memcpy(pin, gx.data(), pb); std::memcpy(pin + pb, gy.data(), pb); std::memcpy(pin + 2 * pb, idx.data(), (size_t)p * 4);
    CU(cudaMemcpyAsync(d_gx, pin, pb, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d_gy, pin + pb, pb, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d_idx, pin + 2 * pb, (size_t)p * 4, cudaMemcpyHostToDevice, st));
    CU(cudaEventRecord(h->pin_ev[pslot], st));
    h->pin_used[pslot] = true;
  }
  CU(cudaMemsetAsync(h->d_flags, 0, sizeof(int) * 4, st));
  CU(cudaMemsetAsync(h->d_scal, 0, sizeof(double) * kNumSlots, st));
  if (in.X32) gather_cols_f32(in.X32, in.ldx, m, d_idx, p, Xg, ldp, st);
  else gather_cols(in.X, in.ldx, m, d_idx, p, Xg, ldp, st);
  if (in.U32) gather_cols_f32(in.U32, in.ldu, m, d_idx, p, Ug, ldp, st);
  else gather_cols(in.U, in.ldu, m, d_idx, p, Ug, ldp, st);
  copy_mat(in.B, in.ldb, m, q, Bw, ldq, st);
  copy_mat(in.V, in.ldv, q, q, Vw, ldq, st);
  build_C(C, ldp, p, d_gx, d_gy, th, nullptr, 0, st);

  // ---- forward ------------------------------------------------------------------------------------------
  auto rowdot_gemm = [&](const double* Xm, double* Z, double* s_out, double* n_out) -> int {   // Z = Xm C, s = rowdot + s0^2
    GemmArgs g{};
    g.A = Xm; g.lda = ldp; g.B = C; g.ldb = ldp; g.C = Z; g.ldc = ldp; g.M = m; g.N = p; g.K = p;
    g.aux = Xm; g.ldaux = ldp; g.part1 = part1; g.ldpart = ldpart;
    CU(gemm(st, 0, 1, kRowdot, true, 1, g));
    norms_finalize(part1, ldpart, ntp_s, m, th.s02, s_out, n_out, st);
    return SGP_OK;
  };
  rc = rowdot_gemm(Ug, ZU, tU, X2); if (rc) return rc;                                      // U:978-979 on xtilde
  rc = rowdot_gemm(Xg, ZX, sX, nX); if (rc) return rc;                                      // Kvec from x (U:2033)
  {
    GemmArgs a{};   // K_tilde (raw) = acos-kernel(ZU Ug^T), cos kept for the backward pass
    a.A = ZU; a.lda = ldp; a.B = Ug; a.ldb = ldp; a.C = Kraw; a.ldc = ldm; a.M = m; a.N = m; a.K = p;
    a.rowvec = X2; a.colvec = X2; a.C2 = cosU; a.s02 = th.s02;
    CU(gemm(st, 0, 0, kAcos, true, 1, a));
  }
  symmetrize(Kraw, ldm, m, Kt, ldm, st);                                                    // U:1024-1025
  CU(mm(st, 0, 1, Kt, ldm, Bw, ldq, Kb, ldq, m, q, m));                                      // K_b = K B, K = K_tilde (U:2032, 2049)
  CU(mm(st, 1, 1, Bw, ldq, Kb, ldq, Ktb_raw, ldq, q, q, m));                                 // B^T K_tilde B (U:2047)
  symmetrize(Ktb_raw, ldq, q, Ktb, ldq, st);                                                // U:2048
  copy_mat(Ktb, ldq, q, q, Lk, ldq, st);
  CU(factor_any(h, Lk, ldq, q, Linv, tri_tmp, winv, kFlagKtNotPD, kFlagKtSmallPivot, kSlotLogdetK, st));
  CU(mm(st, 1, 1, Linv, ldq, Linv, ldq, Kinv_raw, ldq, q, q, q));                            // K_tilde_b^-1 (U:2067)
  symmetrize(Kinv_raw, ldq, q, Kinv, ldq, st);
  gemv(Kinv, ldq, q, q, 0, in.m_b, bq, st);
  CU(mm(st, 0, 1, Kinv, ldq, Vw, ldq, KV, ldq, q, q, q));
  CU(mm(st, 0, 1, KV, ldq, Kinv, ldq, KVK, ldq, q, q, q));
  kl_scalars(in.m_b, bq, Vw, ldq, Kinv, ldq, q, h->d_scal, kSlotMb, kSlotTrVK, Gq, st);      // Gq: scratch until the backward pass
  if (!in.logdetV_host) {
    copy_mat(Vw, ldq, q, q, Lv, ldq, st);
    CU(factor_any(h, Lv, ldq, q, nullptr, nullptr, winv2, kFlagVNotPD, kFlagVSmallPivot, kSlotLogdetV, st));
  }
  {
    GemmArgs g{};   // aV = a V^T with a = B; part1 = rowdot(aV, a), part2 = a m_b   (U:1090-1101)
    g.A = Bw; g.lda = ldq; g.B = Vw; g.ldb = ldq; g.C = aV; g.ldc = ldq; g.M = m; g.N = q; g.K = q;
    g.aux = Bw; g.ldaux = ldq; g.colvec = in.m_b; g.part1 = part1; g.part2 = part2; g.ldpart = ldpart;
    CU(gemm(st, 0, 0, kMoments, false, 1, g));
  }
  moments_combine(sX, part1, part2, ldpart, ntq, ntq, Bw, ldq, Kb, ldq, m, q, lm_t, lv_t, st);
  moments_rows(lv_t, part1, lm_t, ldpart, 0, 1, in.r, m, A, in.lambda0, lam_m, lam_var, fvec, u_s, v_s, blockpart, st);
  sum_to_slots(blockpart, nbm, 4, 4, h->d_scal, kSlotRLam, 0, st);
  // ---- backward -----------------------------------------------------------------------------------------
  same_ga(Ga, ldq, m, q, u_s, v_s, in.m_b, aV, Kb, st);
  CU(mm(st, 0, 1, Ga, ldq, Kinv, ldq, Hm, ldq, m, q, q));                                    // H = Ga K_tilde_b^-1
  CU(mm(st, 1, 1, Bw, ldq, Hm, ldq, Gq, ldq, q, q, m, -1.0));                                // -a^T H
  gkt_assemble(Gq, ldq, q, Kinv, KVK, bq, h->kl_weight, st);                                // - kl_weight * dKL/dK_tilde_b (U:1333)
  CU(mm(st, 0, 1, Bw, ldq, Gq, ldq, BGq, ldq, m, q, q));
  same_combine(Em, ldq, m, q, Hm, v_s, Bw, BGq, st);
  CU(mm(st, 0, 0, Em, ldq, Bw, ldq, GKt, ldm, m, m, q));                                     // adjoint of the m x m dK_tilde
  acos_backward(1, GKt, cosU, nullptr, ldm, m, m, nullptr, nullptr, nullptr, X2, X2, wr, colkU, nullptr, ldm, sumpartU,
                st);
  col_reduce(colkU, nblk_bwd, m, ldm, kappaU, 0, st);
  col_weights(kappaU, nullptr, nullptr, X2, m, wc, st);
  sum_to_slot(sumpartU, nblk_bwd, 1, 1.0, h->d_scal, kSlotSig, 1, st);
  sum_to_slot(wr, m, 1, 1.0, h->d_scal, kSlotSig, 1, st);
  sum_to_slot(wc, m, 1, 1.0, h->d_scal, kSlotSig, 1, st);
  sum_to_slot(v_s, m, 1, 1.0, h->d_scal, kSlotSig, 1, st);                                  // dKvec[sigma_0] = 2 sigma_0 (U:1036)
  lincomb(wU, ldm, 1, m, 1.0, wr, ldm, 1.0, wc, ldm, 0.0, st);
  {
    GemmArgs g{};   // YU = R Ug + diag(wU) Ug  (into ZU), then Ug^T YU
    g.A = GKt; g.lda = ldm; g.B = Ug; g.ldb = ldp; g.C = ZU; g.ldc = ldp; g.M = m; g.N = p; g.K = m;
    g.rowvec = wU; g.aux = Ug; g.ldaux = ldp;
    CU(gemm(st, 0, 1, kAxpyRow, true, 1, g));
  }
  CU(mm(st, 1, 1, Ug, ldp, ZU, ldp, GCU, ldp, p, p, m));
  scale_rows(YX, Xg, ldp, m, v_s, st);                                                      // dKvec[k] = x_i^T dC_k x_i (U:1042)
  CU(mm(st, 1, 1, Xg, ldp, YX, ldp, GCX, ldp, p, p, m));                                     // X^T diag(v) X
  contract_dC(GCX, 1, (long)p * ldp, GCU, C, ldp, p, d_gx, d_gy, th, gpart, st);
  sum_to_slots(gpart, p, 5, 5, h->d_scal, kSlotG, 0, st);
  const bool async_out = in.out_dev != nullptr;
  finalize(h->d_scal, h->d_flags, A, in.lambda0, th, p, in.logdetV_host ? *in.logdetV_host : 0.0,
           in.logdetV_host ? 1 : 0, async_out ? in.out_dev : h->d_out, 0.0, st);
  CU(cudaGetLastError());
  if (async_out) return SGP_OK;
  CU(cudaMemcpyAsync(h->h_out, h->d_out, sizeof(double) * 16, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (out_host) std::memcpy(out_host, h->h_out, sizeof(double) * 16);
  h->dbg.clear();
  return SGP_OK;
}
}  // namespace

extern "C" {

int sgp_elbo_grad(sgp_handle* h, const double* X, long ldx, const double* r, long N, const double* xtilde, long ldu,
                  int m, int n_px, const double* theta, const double* lo, const double* hi, const double* B, long ldb,
                  int q, const double* m_b, const double* V_b, long ldv, double logA, double lambda0,
                  const double* logdetV_host, double* lam_m, double* lam_var, double* f, double* out_host,
                  void* stream) {
  if (!h || !X || !r || !xtilde || !theta || !B || !m_b || !V_b)
    return fail(h, SGP_E_ARG, "sgp_elbo_grad: null argument");
  if (lo && hi)
    for (int k = 0; k < 6; ++k)
      if (!(lo[k] <= theta[k] && theta[k] <= hi[k]))
        return fail(h, SGP_E_BOUNDS, "sgp_elbo_grad: theta[" + std::to_string(k) + "] outside its limits");
  ClosureIn in{};
  in.X = X; in.ldx = ldx; in.r = r; in.N = N; in.U = xtilde; in.ldu = ldu; in.m = m; in.n_px = n_px;
  in.theta = theta; in.B = B; in.ldb = ldb; in.q = q; in.m_b = m_b; in.V = V_b; in.ldv = ldv;
  in.logA = logA; in.lambda0 = lambda0; in.logdetV_host = logdetV_host;
  in.lam_m = lam_m; in.lam_var = lam_var; in.f = f;
  in.exchange = true;       // no-op unless sgp_comm_import made this handle one shard of an image-sharded evaluation
  return run_closure(h, in, out_host, (cudaStream_t)stream);
}

int sgp_elbo_grad_f32(sgp_handle* h, const float* X, long ldx, const double* r, long N, const float* xtilde, long ldu,
                      int m, int n_px, const double* theta, const double* lo, const double* hi, const double* B, long ldb,
                      int q, const double* m_b, const double* V_b, long ldv, double logA, double lambda0,
                      const double* logdetV_host, double* lam_m, double* lam_var, double* f, double* out_host,
                      void* stream) {
  if (!h || !X || !r || !xtilde || !theta || !B || !m_b || !V_b)
    return fail(h, SGP_E_ARG, "sgp_elbo_grad_f32: null argument");
  if (lo && hi)
    for (int k = 0; k < 6; ++k)
      if (!(lo[k] <= theta[k] && theta[k] <= hi[k]))
        return fail(h, SGP_E_BOUNDS, "sgp_elbo_grad_f32: theta[" + std::to_string(k) + "] outside its limits");
  ClosureIn in{};
  in.X32 = X; in.U32 = xtilde;
  in.X = reinterpret_cast<const double*>(X); in.U = reinterpret_cast<const double*>(xtilde);   // cache key only
  in.ldx = ldx; in.r = r; in.N = N; in.ldu = ldu; in.m = m; in.n_px = n_px;
  in.theta = theta; in.B = B; in.ldb = ldb; in.q = q; in.m_b = m_b; in.V = V_b; in.ldv = ldv;
  in.logA = logA; in.lambda0 = lambda0; in.logdetV_host = logdetV_host;
  in.lam_m = lam_m; in.lam_var = lam_var; in.f = f;
  in.exchange = true;
  // fp32-class engine unless the caller chose one: 5 int8 slices (35 bits per operand >= the 24 of the inputs)
  const int keep = h->i8_slices, keep_i = h->i8_slices_images, keep_w = h->i8_slices_sig2;
  if (h->i8_slices < 0) { h->i8_slices = 5; h->i8_slices_images = 5; h->i8_slices_sig2 = 5; }
  const int rc = run_closure(h, in, out_host, (cudaStream_t)stream);
  h->i8_slices = keep; h->i8_slices_images = keep_i; h->i8_slices_sig2 = keep_w;
  return rc;
}

int sgp_elbo_grad_dev(sgp_handle* h, const double* X, long ldx, const double* r, long N, const double* xtilde, long ldu,
                      int m, int n_px, const double* theta, const double* B, long ldb, int q, const double* m_b,
                      const double* V_b, long ldv, double logA, double lambda0, const double* logdetV_host,
                      double* lam_m, double* lam_var, double* f, double* out_dev, void* stream) {
  if (!h || !X || !r || !xtilde || !theta || !B || !m_b || !V_b || !out_dev)
    return fail(h, SGP_E_ARG, "sgp_elbo_grad_dev: null argument");
  ClosureIn in{};
  in.X = X; in.ldx = ldx; in.r = r; in.N = N; in.U = xtilde; in.ldu = ldu; in.m = m; in.n_px = n_px;
  in.theta = theta; in.B = B; in.ldb = ldb; in.q = q; in.m_b = m_b; in.V = V_b; in.ldv = ldv;
  in.logA = logA; in.lambda0 = lambda0; in.logdetV_host = logdetV_host;
  in.lam_m = lam_m; in.lam_var = lam_var; in.f = f; in.out_dev = out_dev;
  in.exchange = true;
  return run_closure(h, in, nullptr, (cudaStream_t)stream);
}

int sgp_elbo_grad_cells(sgp_handle* h, const double* X, long ldx, long N, const double* xtilde, long ldu, int m,
                        int n_px, int ncells, const double* const* r, const double* theta, const double* const* B,
                        const long* ldb, const int* q, const double* const* m_b, const double* const* V_b,
                        const long* ldv, const double* logA, const double* lambda0, const double* logdetV,
                        double* out_host, void* stream) {
  if (!h || !X || !xtilde || !r || !theta || !B || !ldb || !q || !m_b || !V_b || !ldv || !logA || !lambda0 || !out_host)
    return fail(h, SGP_E_ARG, "sgp_elbo_grad_cells: null argument");
  if (ncells <= 0) return fail(h, SGP_E_ARG, "sgp_elbo_grad_cells: ncells <= 0");
  int rc = ensure_ready(h); if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  CU(h->cells_out.reserve((size_t)ncells * 16 * sizeof(double)));
  double* d_out = (double*)h->cells_out.p;
  // every cell has its own receptive field, hence its own mask; they all run on the whole pixel grid (C = 0 outside a
  // cell's mask) so that ONE resident copy of the images and ONE set of int8 slices serves the population
  const long P = (long)n_px * n_px;
  bool any_cropped = false;
  for (int c = 0; c < ncells && !any_cropped; ++c) {
    int pc = 0;
    sgp_mask_host(theta + 6 * c, n_px, nullptr, &pc);
    any_cropped = pc < P;
  }
  for (int c = 0; c < ncells; ++c) {
    if (!r[c] || !B[c] || !m_b[c] || !V_b[c]) return fail(h, SGP_E_ARG, "sgp_elbo_grad_cells: null per-cell pointer");
    ClosureIn in{};
    in.X = X; in.ldx = ldx; in.r = r[c]; in.N = N; in.U = xtilde; in.ldu = ldu; in.m = m; in.n_px = n_px;
    in.theta = theta + 6 * c; in.B = B[c]; in.ldb = ldb[c]; in.q = q[c]; in.m_b = m_b[c]; in.V = V_b[c]; in.ldv = ldv[c];
    in.logA = logA[c]; in.lambda0 = lambda0[c]; in.logdetV_host = logdetV ? logdetV + c : nullptr;
    in.out_dev = d_out + 16 * c;
    in.force_support = any_cropped ? 1 : -2;
    rc = run_closure(h, in, nullptr, st);
    if (rc) return rc;
  }
  CU(cudaMemcpyAsync(out_host, d_out, (size_t)ncells * 16 * sizeof(double), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return SGP_OK;
}

int sgp_elbo_grad_host(sgp_handle* h, const double* X_host, long ldx, const double* r_host, long N,
                       const double* xtilde_host, long ldu, int m, int n_px, const double* theta,
                       const double* B_host, long ldb, int q, const double* m_b_host, const double* V_b_host,
                       long ldv, double logA, double lambda0, double* out_host) {
  if (!h || !X_host || !r_host || !xtilde_host || !theta || !B_host || !m_b_host || !V_b_host)
    return fail(h, SGP_E_ARG, "sgp_elbo_grad_host: null argument");
  if (N <= 0 || m <= 0 || n_px <= 0 || q <= 0 || q > m) return fail(h, SGP_E_ARG, "sgp_elbo_grad_host: bad sizes");
  if (ldx < (long)n_px * n_px || ldu < (long)n_px * n_px || ldb < q || ldv < q)
    return fail(h, SGP_E_ARG, "sgp_elbo_grad_host: leading dimension smaller than the row length");
  int rc = ensure_ready(h); if (rc) return rc;
  if (!h->compute_stream) CU(cudaStreamCreateWithFlags(&h->compute_stream, cudaStreamNonBlocking));
  cudaStream_t st = h->compute_stream;
  const long P = (long)n_px * n_px;
  // small operands: plain synchronous uploads into a dedicated buffer (a few MB)
  DevBuf& small = h->small;
  const size_t nU = (size_t)m * P, nB = (size_t)m * q, nV = (size_t)q * q;
  CU(small.reserve((nU + nB + nV + q + 64) * sizeof(double)));
  double* dU = (double*)small.p; double* dB = dU + ((nU + 1) & ~size_t(1)); double* dV = dB + ((nB + 1) & ~size_t(1));
  double* dm = dV + ((nV + 1) & ~size_t(1));
  CU(ensure_copy_stream(h));
  cudaStream_t up = h->copy_stream;   // run_closure makes the compute stream wait for these (ev_small)
  CU(cudaMemcpy2DAsync(dU, P * sizeof(double), xtilde_host, ldu * sizeof(double), P * sizeof(double), m,
                       cudaMemcpyHostToDevice, up));
  CU(cudaMemcpy2DAsync(dB, q * sizeof(double), B_host, ldb * sizeof(double), q * sizeof(double), m,
                       cudaMemcpyHostToDevice, up));
  CU(cudaMemcpy2DAsync(dV, q * sizeof(double), V_b_host, ldv * sizeof(double), q * sizeof(double), q,
                       cudaMemcpyHostToDevice, up));
  CU(cudaMemcpyAsync(dm, m_b_host, q * sizeof(double), cudaMemcpyHostToDevice, up));
  ClosureIn in{};
  in.X = nullptr; in.ldx = ldx; in.r = nullptr; in.N = N; in.U = dU; in.ldu = P; in.m = m; in.n_px = n_px;
  in.theta = theta; in.B = dB; in.ldb = q; in.q = q; in.m_b = dm; in.V = dV; in.ldv = q;
  in.logA = logA; in.lambda0 = lambda0; in.logdetV_host = nullptr;
  in.X_host = X_host; in.r_host = r_host;
  return run_closure(h, in, out_host, st);
}

int sgp_comm_export(sgp_handle* h, int world, unsigned char* handle_out) {
  if (!h || !handle_out) return fail(h, SGP_E_ARG, "sgp_comm_export: null argument");
  if (world < 2 || world > kCommMaxRanks) return fail(h, SGP_E_ARG, "sgp_comm_export: world must be 2..16");
  int rc = ensure_ready(h); if (rc) return rc;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  h->comm_on = false;
  CU(h->comm_buf.reserve(comm_buffer_bytes(world)));
  CU(cudaMemset(h->comm_buf.p, 0, comm_buffer_bytes(world)));
  cudaIpcMemHandle_t ih;
  CU(cudaIpcGetMemHandle(&ih, h->comm_buf.p));
  std::memcpy(handle_out, &ih, 64);
  return SGP_OK;
}

int sgp_comm_import(sgp_handle* h, int rank, int world, const unsigned char* handles) {
  if (!h || !handles) return fail(h, SGP_E_ARG, "sgp_comm_import: null argument");
  if (world < 2 || world > kCommMaxRanks || rank < 0 || rank >= world) return fail(h, SGP_E_ARG, "sgp_comm_import: bad rank / world");
  if (!h->comm_buf.p) return fail(h, SGP_E_ARG, "sgp_comm_import: call sgp_comm_export first");
  CU(cudaSetDevice(h->device));
  CommDev c;
  c.rank = rank; c.world = world; c.epoch = 0;
  for (int i = 0; i < world; ++i) {
    void* base = h->comm_buf.p;
    if (i != rank) {
      cudaIpcMemHandle_t ih;
      std::memcpy(&ih, handles + 64 * i, 64);
      if (h->comm_peer[i]) { cudaIpcCloseMemHandle(h->comm_peer[i]); h->comm_peer[i] = nullptr; }
      CU(cudaIpcOpenMemHandle(&base, ih, cudaIpcMemLazyEnablePeerAccess));
      h->comm_peer[i] = base;
    }
    c.slots[i] = (double*)base;
    c.flags[i] = (unsigned long long*)((char*)base + (size_t)2 * world * 16 * sizeof(double));
  }
  h->comm = c;
  h->comm_on = true;
  return SGP_OK;
}

int sgp_comm_close(sgp_handle* h) {
  if (!h) return SGP_E_ARG;
  h->comm_on = false;
  if (h->gemm_ready) {
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    for (int i = 0; i < kCommMaxRanks; ++i) if (h->comm_peer[i]) { cudaIpcCloseMemHandle(h->comm_peer[i]); h->comm_peer[i] = nullptr; }
  }
  return SGP_OK;
}

int sgp_upload_images(sgp_handle* h, const double* X_host, long ldx, const double* r_host, long N,
                      const double* xtilde_host, long ldu, int m, int n_px) {
  if (!h || !X_host || !r_host || !xtilde_host) return fail(h, SGP_E_ARG, "sgp_upload_images: null argument");
  if (N <= 0 || m <= 0 || n_px <= 0) return fail(h, SGP_E_ARG, "sgp_upload_images: bad sizes");
  if (ldx < (long)n_px * n_px || ldu < (long)n_px * n_px)
    return fail(h, SGP_E_ARG, "sgp_upload_images: leading dimension smaller than the row length (n_px^2 pixels per image)");
  int rc = ensure_ready(h); if (rc) return rc;
  if (!h->compute_stream) CU(cudaStreamCreateWithFlags(&h->compute_stream, cudaStreamNonBlocking));
  cudaStream_t st = h->compute_stream;
  const long P = (long)n_px * n_px;
  CU(h->res_x.reserve((size_t)N * P * sizeof(double)));
  CU(h->res_r.reserve((size_t)N * sizeof(double)));
  CU(h->res_u.reserve((size_t)m * P * sizeof(double)));
  CU(cudaMemcpy2DAsync(h->res_x.p, P * sizeof(double), X_host, ldx * sizeof(double), P * sizeof(double), N,
                       cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(h->res_r.p, r_host, (size_t)N * sizeof(double), cudaMemcpyHostToDevice, st));
  CU(cudaMemcpy2DAsync(h->res_u.p, P * sizeof(double), xtilde_host, ldu * sizeof(double), P * sizeof(double), m,
                       cudaMemcpyHostToDevice, st));
  CU(cudaStreamSynchronize(st));
  h->res_N = N; h->res_m = m; h->res_npx = n_px;
  h->cache_valid = false; h->oz_valid = false;      // new content at (possibly) the same addresses
  h->image_token = h->image_token * 2654435761ULL + 0x9e3779b97f4a7c15ULL;   // and a token nobody else has handed out
  return SGP_OK;
}

int sgp_elbo_grad_resident(sgp_handle* h, const double* theta, const double* lo, const double* hi,
                           const double* B_host, long ldb, int q, const double* m_b_host, const double* V_b_host,
                           long ldv, double logA, double lambda0, double* out_host) {
  if (!h || !theta || !B_host || !m_b_host || !V_b_host || !out_host)
    return fail(h, SGP_E_ARG, "sgp_elbo_grad_resident: null argument");
  if (h->res_N <= 0) return fail(h, SGP_E_ARG, "sgp_elbo_grad_resident: call sgp_upload_images first");
  if (lo && hi)
    for (int k = 0; k < 6; ++k)
      if (!(lo[k] <= theta[k] && theta[k] <= hi[k]))
        return fail(h, SGP_E_BOUNDS, "sgp_elbo_grad_resident: theta[" + std::to_string(k) + "] outside its limits");
  int rc = ensure_ready(h); if (rc) return rc;
  cudaStream_t st = h->compute_stream;
  const int m = (int)h->res_m;
  const long P = (long)h->res_npx * h->res_npx;
  if (q <= 0 || q > m || ldb < q || ldv < q) return fail(h, SGP_E_ARG, "sgp_elbo_grad_resident: bad q / leading dimensions");
  const size_t nB = (size_t)m * q, nV = (size_t)q * q;
  CU(h->small.reserve((nB + nV + q + 64) * sizeof(double)));
  double* dB = (double*)h->small.p; double* dV = dB + ((nB + 1) & ~size_t(1)); double* dm = dV + ((nV + 1) & ~size_t(1));
  // the per-evaluation state: B (m x q), V_b (q x q), m_b (q) -- a few MB; the images stay where sgp_upload_images put them
  CU(cudaMemcpy2DAsync(dB, q * sizeof(double), B_host, ldb * sizeof(double), q * sizeof(double), m,
                       cudaMemcpyHostToDevice, st));
  CU(cudaMemcpy2DAsync(dV, q * sizeof(double), V_b_host, ldv * sizeof(double), q * sizeof(double), q,
                       cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(dm, m_b_host, q * sizeof(double), cudaMemcpyHostToDevice, st));
  ClosureIn in{};
  in.X = (const double*)h->res_x.p; in.ldx = P; in.r = (const double*)h->res_r.p; in.N = h->res_N;
  in.U = (const double*)h->res_u.p; in.ldu = P; in.m = m; in.n_px = h->res_npx;
  in.theta = theta; in.B = dB; in.ldb = q; in.q = q; in.m_b = dm; in.V = dV; in.ldv = q;
  in.logA = logA; in.lambda0 = lambda0; in.logdetV_host = nullptr;
  return run_closure(h, in, out_host, st);
}

}  // extern "C"

// =====================================================================================================
// function-level entry points of the reference call surface (a2, a3, a5, a6-a8, a10)
// =====================================================================================================
namespace {

bool aligned2(const void* p, long ld) { return ((uintptr_t)p % 16 == 0) && (ld % 2 == 0); }

}  // namespace

extern "C" {

}  // extern "C"

// K (and, when Kvec_out is given, Kvec_i = x1_i^T C x1_i + sigma_0^2, which the kernel needs anyway as X1^2)
static int acosker_impl(sgp_handle* h, const double* x1, long n1, long ld1, const double* x2, long n2, long ld2, int p,
                        const double* C, long ldc, double sigma_0, double* K, long ldk, int symm, double* Kvec_out,
                        void* stream) {
  if (!h || !x1 || !x2 || !C || !K) return fail(h, SGP_E_ARG, "sgp_acosker: null argument");
  if (symm && n1 != n2) return fail(h, SGP_E_ARG, "sgp_acosker: symmetrize needs n1 == n2");
  if (n1 <= 0 || n2 <= 0 || p <= 0) return fail(h, SGP_E_ARG, "sgp_acosker: sizes must be positive");
  if (ld1 < p || ld2 < p || ldc < p || ldk < n2)
    return fail(h, SGP_E_ARG, "sgp_acosker: leading dimension smaller than the row length");
  int rc = ensure_ready(h); if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const long ldp = rup(p, 16), ldn2 = rup(n2, 16);
  const long slab = std::min<long>(h->slab_rows, rup(n1, 128));
  const int ntp_s = (p + 63) / 64, ntp = (p + 127) / 128;
  const long ldpart = rup(std::max<long>(slab, n2), 16);
  const double s02 = sigma_0 * sigma_0;
  size_t need = 0;
  auto acc = [&](size_t nelem) { need = ((need + 255) & ~size_t(255)) + nelem * sizeof(double); };
  acc((size_t)p * ldp); acc((size_t)n2 * ldp); acc((size_t)n2 * ldp); acc((size_t)slab * ldp); acc((size_t)slab * ldp);
  // image-sized x1: both products on the int8 tensor-core path (csrc/ozaki_i8.cu), x1's digit slices shared by them
  const int ozS = oz_pick(h, n1, std::min<long>(n2, p), p);
  const int ng16 = (p + 15) / 16;
  const int nparts_max = std::max(ntp_s, ozS ? ng16 : 0);
  acc((size_t)nparts_max * ldpart); acc(n2); acc(n2); acc(slab); acc(slab); acc((size_t)slab * ldn2); acc((size_t)slab * ldn2);
  acc(symm ? (size_t)n1 * ldn2 : 1);
  if (ozS) {
    acc((size_t)n2 * ldp); acc(oz_slice_bytes(slab, p, ozS) / 8 + 16); acc(oz_slice_bytes(p, p, ozS) / 8 + 16);
    acc(oz_slice_bytes(n2, p, ozS) / 8 + 16); acc(slab / 2 + 8); acc(p); acc(n2); acc(p);
  }
  need += 4096;
  RESERVE_ARENA(need);
  Arena ar(h->arena.p, h->arena.bytes);
  double* Cw = ar.take<double>((size_t)p * ldp);
  double* x2w = ar.take<double>((size_t)n2 * ldp);
  double* Z2 = ar.take<double>((size_t)n2 * ldp);
  double* x1w = ar.take<double>((size_t)slab * ldp);
  double* Z1 = ar.take<double>((size_t)slab * ldp);
  double* part = ar.take<double>((size_t)nparts_max * ldpart);
  double* t2 = ar.take<double>(n2); double* X2 = ar.take<double>(n2);
  double* s1 = ar.take<double>(slab); double* X1 = ar.take<double>(slab);
  double* Kw = ar.take<double>((size_t)slab * ldn2); double* cosw = ar.take<double>((size_t)slab * ldn2);
  double* Kfull = symm ? ar.take<double>((size_t)n1 * ldn2) : nullptr;
  double* Wt = nullptr; int8_t* xq = nullptr; int8_t* cq = nullptr; int8_t* wq = nullptr;
  int* xe = nullptr; int* ce = nullptr; int* we = nullptr; unsigned long long* cmx = nullptr;
  if (ozS) {
    Wt = ar.take<double>((size_t)n2 * ldp);
    xq = (int8_t*)ar.take<double>(oz_slice_bytes(slab, p, ozS) / 8 + 16);
    cq = (int8_t*)ar.take<double>(oz_slice_bytes(p, p, ozS) / 8 + 16);
    wq = (int8_t*)ar.take<double>(oz_slice_bytes(n2, p, ozS) / 8 + 16);
    xe = (int*)ar.take<double>(slab / 2 + 8); ce = (int*)ar.take<double>(p); we = (int*)ar.take<double>(n2);
    cmx = (unsigned long long*)ar.take<double>(p);
  }

  const double* Cc = C; long ldcc = ldc;
  if (!aligned2(C, ldc)) { copy_mat(C, ldc, p, p, Cw, ldp, st); Cc = Cw; ldcc = ldp; }
  const double* x2c = x2; long ld2c = ld2;
  if (!aligned2(x2, ld2)) { copy_mat(x2, ld2, (int)n2, p, x2w, ldp, st); x2c = x2w; ld2c = ldp; }
  {
    GemmArgs g{};
    g.A = x2c; g.lda = ld2c; g.B = Cc; g.ldb = ldcc; g.C = Z2; g.ldc = ldp; g.M = (int)n2; g.N = p; g.K = p;
    g.aux = x2c; g.ldaux = ld2c; g.part1 = part; g.ldpart = ldpart;
    CU(gemm(st, 0, 1, kRowdot, true, 1, g));
    norms_finalize(part, ldpart, ntp_s, n2, s02, t2, X2, st);
  }
  if (ozS) {
    // x1^T C x2 = x1 . (x2 C^T): the p x p product moves to the n2 side, exact for any C
    CU(mm(st, 0, 0, x2c, ld2c, Cc, ldcc, Wt, ldp, (int)n2, p, p));
    CU(oz_slice_rows(Wt, ldp, n2, p, ozS, wq, oz_ldk(p), we, st));
    CU(oz_slice_cols(Cc, ldcc, p, p, ozS, cq, oz_ldk(p), ce, cmx, st));     // B operand of x1 C: rows = columns of C
  }
  for (long i0 = 0; i0 < n1; i0 += slab) {
    const long ns = std::min(slab, n1 - i0);
    const double* xs = x1 + i0 * ld1; long lds = ld1;
    if (!aligned2(xs, ld1)) { copy_mat(xs, ld1, (int)ns, p, x1w, ldp, st); xs = x1w; lds = ldp; }
    if (ozS) {
      CU(oz_slice_rows(xs, lds, ns, p, ozS, xq, oz_ldk(p), xe, st));
      const OzMoments mom{xs, lds, part, ldpart};                           // rowdot(x1 C, x1)  (U:978)
      CU(oz_gemm(xq, xe, ns, cq, ce, p, p, oz_ldk(p), ozS, Z1, ldp, 1, 0, 0, st, &mom));
      norms_finalize(part, ldpart, ng16, ns, s02, s1, X1, st);
      if (Kvec_out) CU(cudaMemcpyAsync(Kvec_out + i0, s1, sizeof(double) * ns, cudaMemcpyDeviceToDevice, st));
      CU(oz_gemm(xq, xe, ns, wq, we, (int)n2, p, oz_ldk(p), ozS, Kw, ldn2, 1, 0, 0, st));
      acos_forward(Kw, cosw, ldn2, ns, (int)n2, X1, X2, s02, st);           // U:982-990
      if (symm) copy_mat(Kw, ldn2, (int)ns, (int)n2, Kfull + i0 * ldn2, ldn2, st);
      else copy_mat(Kw, ldn2, (int)ns, (int)n2, K + i0 * ldk, ldk, st);
      continue;
    }
    GemmArgs g{};
    g.A = xs; g.lda = lds; g.B = Cc; g.ldb = ldcc; g.C = Z1; g.ldc = ldp; g.M = (int)ns; g.N = p; g.K = p;
    g.aux = xs; g.ldaux = lds; g.part1 = part; g.ldpart = ldpart;
    CU(gemm(st, 0, 1, kRowdot, false, 1, g));
    norms_finalize(part, ldpart, ntp, ns, s02, s1, X1, st);
    if (Kvec_out) CU(cudaMemcpyAsync(Kvec_out + i0, s1, sizeof(double) * ns, cudaMemcpyDeviceToDevice, st));
    GemmArgs a{};  // K slab into scratch (ld = ldn2, shared with the cosine scratch), then out
    a.A = Z1; a.lda = ldp; a.B = x2c; a.ldb = ld2c; a.C = Kw; a.ldc = ldn2; a.M = (int)ns; a.N = (int)n2; a.K = p;
    a.rowvec = X1; a.colvec = X2; a.C2 = cosw; a.s02 = s02;
    CU(gemm(st, 0, 0, kAcos, false, 1, a));
    if (symm) copy_mat(Kw, ldn2, (int)ns, (int)n2, Kfull + i0 * ldn2, ldn2, st);
    else copy_mat(Kw, ldn2, (int)ns, (int)n2, K + i0 * ldk, ldk, st);
  }
  if (symm) symmetrize(Kfull, ldn2, (int)n1, K, ldk, st);   // U:1024-1025
  CU(cudaGetLastError());
  return SGP_OK;
}

extern "C" {

int sgp_acosker(sgp_handle* h, const double* x1, long n1, long ld1, const double* x2, long n2, long ld2, int p,
                const double* C, long ldc, double sigma_0, double* K, long ldk, int symm, void* stream) {
  return acosker_impl(h, x1, n1, ld1, x2, n2, ld2, p, C, ldc, sigma_0, K, ldk, symm, nullptr, stream);
}

int sgp_acosker_kvec(sgp_handle* h, const double* x1, long n1, long ld1, const double* x2, long n2, long ld2, int p,
                     const double* C, long ldc, double sigma_0, double* K, long ldk, int symm, double* Kvec,
                     void* stream) {
  if (!Kvec) return fail(h, SGP_E_ARG, "sgp_acosker_kvec: null argument");
  return acosker_impl(h, x1, n1, ld1, x2, n2, ld2, p, C, ldc, sigma_0, K, ldk, symm, Kvec, stream);
}

// New columns of K against images whose squared norms are already known: kvec1_i = x1_i^T C x1_i + sigma_0^2 is what
// acosker(x1, diag=True) returned at this theta (the model's 'Kvec').  The n1 x p x p product x1 C -- all of the cost of
// sgp_acosker when n2 is a handful of images -- is not repeated: K_lin = x1 (x2 C)^T + sigma_0^2 moves C to the n2 side
// (x2 C is n2 x p x p), the arc-cosine map is the same epilogue.  One pass over x1: 8 n1 p bytes.
int sgp_acosker_known(sgp_handle* h, const double* x1, long n1, long ld1, const double* kvec1, const double* x2, long n2,
                      long ld2, int p, const double* C, long ldc, double sigma_0, double* K, long ldk, void* stream) {
  if (!h || !x1 || !kvec1 || !x2 || !C || !K) return fail(h, SGP_E_ARG, "sgp_acosker_known: null argument");
  if (n1 <= 0 || n2 <= 0 || p <= 0 || ldk < n2) return fail(h, SGP_E_ARG, "sgp_acosker_known: bad shape");
  if (ld1 < p || ld2 < p || ldc < p)
    return fail(h, SGP_E_ARG, "sgp_acosker_known: leading dimension smaller than the row length");
  int rc = ensure_ready(h); if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const long ldp = rup(p, 16), ldn2 = rup(n2, 16);
  const long slab = std::min<long>(h->slab_rows, rup(n1, 128));
  const int ntp_s = (p + 63) / 64;
  const long ldpart = rup(std::max<long>(slab, n2), 16);
  const double s02 = sigma_0 * sigma_0;
  size_t need = 0;
  auto acc = [&](size_t nelem) { need = ((need + 255) & ~size_t(255)) + nelem * sizeof(double); };
  acc((size_t)p * ldp); acc((size_t)n2 * ldp); acc((size_t)n2 * ldp); acc((size_t)n2 * ldp); acc((size_t)slab * ldp);
  acc((size_t)ntp_s * ldpart); acc(n2); acc(n2); acc(slab); acc(slab); acc((size_t)slab * ldn2); acc((size_t)slab * ldn2);
  need += 4096;
  RESERVE_ARENA(need);
  Arena ar(h->arena.p, h->arena.bytes);
  double* Cw = ar.take<double>((size_t)p * ldp);
  double* x2w = ar.take<double>((size_t)n2 * ldp);
  double* Z2 = ar.take<double>((size_t)n2 * ldp);
  double* Wt = ar.take<double>((size_t)n2 * ldp);
  double* x1w = ar.take<double>((size_t)slab * ldp);
  double* part = ar.take<double>((size_t)ntp_s * ldpart);
  double* t2 = ar.take<double>(n2); double* X2 = ar.take<double>(n2);
  double* s1 = ar.take<double>(slab); double* X1 = ar.take<double>(slab);
  double* Kw = ar.take<double>((size_t)slab * ldn2); double* cosw = ar.take<double>((size_t)slab * ldn2);
  const double* Cc = C; long ldcc = ldc;
  if (!aligned2(C, ldc)) { copy_mat(C, ldc, p, p, Cw, ldp, st); Cc = Cw; ldcc = ldp; }
  const double* x2c = x2; long ld2c = ld2;
  if (!aligned2(x2, ld2)) { copy_mat(x2, ld2, (int)n2, p, x2w, ldp, st); x2c = x2w; ld2c = ldp; }
  {
    GemmArgs g{};                                                            // X2^2 = rowdot(x2 C, x2) + sigma_0^2  (U:979)
    g.A = x2c; g.lda = ld2c; g.B = Cc; g.ldb = ldcc; g.C = Z2; g.ldc = ldp; g.M = (int)n2; g.N = p; g.K = p;
    g.aux = x2c; g.ldaux = ld2c; g.part1 = part; g.ldpart = ldpart;
    CU(gemm(st, 0, 1, kRowdot, true, 1, g));
    norms_finalize(part, ldpart, ntp_s, n2, s02, t2, X2, st);
  }
  CU(mm(st, 0, 0, x2c, ld2c, Cc, ldcc, Wt, ldp, (int)n2, p, p));             // x1^T C x2 = x1 . (x2 C^T)
  for (long i0 = 0; i0 < n1; i0 += slab) {
    const long ns = std::min(slab, n1 - i0);
    const double* xs = x1 + i0 * ld1; long lds = ld1;
    if (!aligned2(xs, ld1)) { copy_mat(xs, ld1, (int)ns, p, x1w, ldp, st); xs = x1w; lds = ldp; }
    norms_finalize(kvec1 + i0, ldpart, 1, ns, 0.0, s1, X1, st);             // X1 = sqrt(kvec1)  (U:978)
    GemmArgs a{};
    a.A = xs; a.lda = lds; a.B = Wt; a.ldb = ldp; a.C = Kw; a.ldc = ldn2; a.M = (int)ns; a.N = (int)n2; a.K = p;
    a.rowvec = X1; a.colvec = X2; a.C2 = cosw; a.s02 = s02;
    CU(gemm(st, 0, 0, kAcos, false, 1, a));                                  // U:982-990
    copy_mat(Kw, ldn2, (int)ns, (int)n2, K + i0 * ldk, ldk, st);
  }
  CU(cudaGetLastError());
  return SGP_OK;
}

int sgp_acosker_diag(sgp_handle* h, const double* x1, long n1, long ld1, int p, const double* C, long ldc,
                     double sigma_0, double* Kvec, void* stream) {
  if (!h || !x1 || !C || !Kvec) return fail(h, SGP_E_ARG, "sgp_acosker_diag: null argument");
  if (n1 <= 0 || p <= 0 || ld1 < p || ldc < p) return fail(h, SGP_E_ARG, "sgp_acosker_diag: bad sizes / leading dimensions");
  int rc = ensure_ready(h); if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const long ldp = rup(p, 16);
  const long slab = std::min<long>(h->slab_rows, rup(n1, 128));
  const int ozS = oz_pick(h, n1, p, p);      // image-sized x1: x1 C on the int8 tensor-core path
  const int ntp = ozS ? (p + 15) / 16 : (p + 127) / 128;
  const long ldpart = rup(slab, 16);
  size_t need = 0;
  auto acc = [&](size_t nelem) { need = ((need + 255) & ~size_t(255)) + nelem * sizeof(double); };
  acc((size_t)p * ldp); acc((size_t)slab * ldp); acc((size_t)slab * ldp); acc((size_t)ntp * ldpart); acc(slab);
  if (ozS) { acc(oz_slice_bytes(slab, p, ozS) / 8 + 16); acc(oz_slice_bytes(p, p, ozS) / 8 + 16); acc(slab / 2 + 8); acc(p); acc(p); }
  need += 4096;
  RESERVE_ARENA(need);
  Arena ar(h->arena.p, h->arena.bytes);
  double* Cw = ar.take<double>((size_t)p * ldp);
  double* x1w = ar.take<double>((size_t)slab * ldp);
  double* Z1 = ar.take<double>((size_t)slab * ldp);
  double* part = ar.take<double>((size_t)ntp * ldpart);
  double* nrm = ar.take<double>(slab);
  const double* Cc = C; long ldcc = ldc;
  if (!aligned2(C, ldc)) { copy_mat(C, ldc, p, p, Cw, ldp, st); Cc = Cw; ldcc = ldp; }
  int8_t* xq = nullptr; int8_t* cq = nullptr; int* xe = nullptr; int* ce = nullptr;
  if (ozS) {
    xq = (int8_t*)ar.take<double>(oz_slice_bytes(slab, p, ozS) / 8 + 16);
    cq = (int8_t*)ar.take<double>(oz_slice_bytes(p, p, ozS) / 8 + 16);
    xe = (int*)ar.take<double>(slab / 2 + 8); ce = (int*)ar.take<double>(p);
    unsigned long long* cmx = (unsigned long long*)ar.take<double>(p);
    CU(oz_slice_cols(Cc, ldcc, p, p, ozS, cq, oz_ldk(p), ce, cmx, st));
  }
  for (long i0 = 0; i0 < n1; i0 += slab) {
    const long ns = std::min(slab, n1 - i0);
    const double* xs = x1 + i0 * ld1; long lds = ld1;
    if (!aligned2(xs, ld1)) { copy_mat(xs, ld1, (int)ns, p, x1w, ldp, st); xs = x1w; lds = ldp; }
    if (ozS) {
      CU(oz_slice_rows(xs, lds, ns, p, ozS, xq, oz_ldk(p), xe, st));
      const OzMoments mom{xs, lds, part, ldpart};
      CU(oz_gemm(xq, xe, ns, cq, ce, p, p, oz_ldk(p), ozS, Z1, ldp, 1, 0, 0, st, &mom));
      norms_finalize(part, ldpart, ntp, ns, sigma_0 * sigma_0, Kvec + i0, nrm, st);
      continue;
    }
    GemmArgs g{};
    g.A = xs; g.lda = lds; g.B = Cc; g.ldb = ldcc; g.C = Z1; g.ldc = ldp; g.M = (int)ns; g.N = p; g.K = p;
    g.aux = xs; g.ldaux = lds; g.part1 = part; g.ldpart = ldpart;
    CU(gemm(st, 0, 1, kRowdot, false, 1, g));
    norms_finalize(part, ldpart, ntp, ns, sigma_0 * sigma_0, Kvec + i0, nrm, st);
  }
  CU(cudaGetLastError());
  return SGP_OK;
}

int sgp_nd_utility(sgp_handle* h, const double* sigma2, const double* mu, long nstar, const double* r, int nr,
                   double* U, double* lam, unsigned char* mask, double* p, double* logp, void* stream) {
  if (!h || !sigma2 || !mu || !r) return fail(h, SGP_E_ARG, "sgp_nd_utility: null argument");
  if (nstar < 0 || nr < 0) return fail(h, SGP_E_ARG, "sgp_nd_utility: bad sizes");
  int rc = ensure_ready(h); if (rc) return rc;
  nd_utility(sigma2, mu, r, nr, nstar, U, lam, mask, p, logp, (cudaStream_t)stream);
  CU(cudaGetLastError());
  return SGP_OK;
}

int sgp_lambda_moments(sgp_handle* h, const double* a, long lda, const double* K_b, long ldk, const double* Kvec,
                       long N, int q, const double* m_b, const double* V_b, long ldv, double* lam_m, double* lam_var,
                       void* stream) {
  if (!h || !a || !K_b || !Kvec || !m_b || !V_b || !lam_m || !lam_var)
    return fail(h, SGP_E_ARG, "sgp_lambda_moments: null argument");
  if (N <= 0 || q <= 0 || lda < q || ldk < q || ldv < q)
    return fail(h, SGP_E_ARG, "sgp_lambda_moments: bad sizes / leading dimensions");
  int rc = ensure_ready(h); if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const long ldq = rup(q, 16);
  const long slab = std::min<long>(h->slab_rows, rup(N, 128));
  const int ozS = oz_pick(h, N, q, q);        // T = a V^T on the int8 tensor-core path (csrc/ozaki_i8.cu)
  const int ntq = ozS ? (q + 15) / 16 : (q + 127) / 128;   // partial row sums per 16- / 128-column group
  const long ldpart = rup(slab, 16);
  size_t need = 0;
  auto acc = [&](size_t nelem) { need = ((need + 255) & ~size_t(255)) + nelem * sizeof(double); };
  acc((size_t)q * ldq); acc((size_t)slab * ldq); acc((size_t)slab * ldq); acc((size_t)ntq * ldpart * 2);
  if (ozS) { acc(oz_slice_bytes(slab, q, ozS) / 8 + 16); acc(oz_slice_bytes(q, q, ozS) / 8 + 16); acc(slab / 2 + 8); acc(q); }
  need += 4096;
  RESERVE_ARENA(need);
  Arena ar(h->arena.p, h->arena.bytes);
  double* Vw = ar.take<double>((size_t)q * ldq);
  double* aw = ar.take<double>((size_t)slab * ldq);
  double* T = ar.take<double>((size_t)slab * ldq);
  double* p1 = ar.take<double>((size_t)ntq * ldpart); double* p2 = ar.take<double>((size_t)ntq * ldpart);
  int8_t* aq = nullptr; int8_t* vq = nullptr; int* ae = nullptr; int* ve = nullptr;
  if (ozS) {
    aq = (int8_t*)ar.take<double>(oz_slice_bytes(slab, q, ozS) / 8 + 16);
    vq = (int8_t*)ar.take<double>(oz_slice_bytes(q, q, ozS) / 8 + 16);
    ae = (int*)ar.take<double>(slab / 2 + 8); ve = (int*)ar.take<double>(q);
    CU(oz_slice_rows(V_b, ldv, q, q, ozS, vq, oz_ldk(q), ve, st));
  } else {
    copy_mat(V_b, ldv, q, q, Vw, ldq, st);
  }
  for (long i0 = 0; i0 < N; i0 += slab) {
    const long ns = std::min(slab, N - i0);
    const double* as = a + i0 * lda; long ldas = lda;
    if (!aligned2(as, lda)) { copy_mat(as, lda, (int)ns, q, aw, ldq, st); as = aw; ldas = ldq; }
    if (ozS) {     // T = a V^T ; part1 = rowdot(T, a) ; part2 = a m_b
      CU(oz_slice_rows(as, ldas, ns, q, ozS, aq, oz_ldk(q), ae, st, m_b, p2));   // p2 = a m_b on the way
      const OzMoments mom{as, ldas, p1, ldpart};
      CU(oz_gemm(aq, ae, ns, vq, ve, q, q, oz_ldk(q), ozS, T, ldq, 1, 0, 0, st, &mom));
    } else {
      GemmArgs g{};
      g.A = as; g.lda = ldas; g.B = Vw; g.ldb = ldq; g.C = T; g.ldc = ldq; g.M = (int)ns; g.N = q; g.K = q;
      g.aux = as; g.ldaux = ldas; g.colvec = m_b; g.part1 = p1; g.part2 = p2; g.ldpart = ldpart;
      CU(gemm(st, 0, 0, kMoments, false, 1, g));
    }
    moments_combine(Kvec + i0, p1, p2, ldpart, ntq, ozS ? 1 : ntq, a + i0 * lda, lda, K_b + i0 * ldk, ldk, ns, q, lam_m + i0,
                    lam_var + i0, st);
  }
  CU(cudaGetLastError());
  return SGP_OK;
}

int sgp_fparam_eval(sgp_handle* h, const double* lam_m, const double* lam_var, const double* r, long N, double logA,
                    double lambda0, double* f, double* out_host, void* stream) {
  if (!h || !lam_m || !lam_var || !r || !out_host) return fail(h, SGP_E_ARG, "sgp_fparam_eval: null argument");
  if (N <= 0) return fail(h, SGP_E_ARG, "sgp_fparam_eval: N must be positive");
  int rc = ensure_ready(h); if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int nb = fparam_blocks(N);
  RESERVE_ARENA((size_t)nb * 8 * sizeof(double) + 4096);
  double* bp = (double*)h->arena.p;
  const double A = std::exp(logA);
  fparam_pass(lam_m, lam_var, r, N, A, lambda0, f, bp, h->d_ticket, h->d_scal + kSlotAux0, st);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(h->h_out, h->d_scal + kSlotAux0, sizeof(double) * 6, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  const double rl = h->h_out[0], sf = h->h_out[1], sdf = h->h_out[2], sr = h->h_out[3], se = h->h_out[4];
  out_host[0] = A * rl + lambda0 * sr - sf;          // loglik            U:1243
  out_host[1] = A * (rl - sdf);                      // dloglik/dlogA     U:1253
  out_host[2] = sr - sf;                             // dloglik/dlambda0  U:1255
  out_host[3] = std::log(sr) - std::log(se);         // lambda0 closed form U:1227
  out_host[4] = sr; out_host[5] = sf; out_host[6] = rl; out_host[7] = h->h_out[5];
  out_host[8] = sdf; out_host[9] = se;               // raw sums, for image-sharded callers (all-reduce, then combine)
  return SGP_OK;
}

int sgp_loglik_sums(sgp_handle* h, const double* r, const double* f, const double* lam_m, const double* lam_var, long N,
                    double logA, double* out_host, void* stream) {
  if (!h || !r || !f || !lam_m || !out_host) return fail(h, SGP_E_ARG, "sgp_loglik_sums: null argument");
  if (N <= 0) return fail(h, SGP_E_ARG, "sgp_loglik_sums: N must be positive");
  int rc = ensure_ready(h); if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int nb = fparam_blocks(N);
  RESERVE_ARENA((size_t)nb * 8 * sizeof(double) + 4096);
  double* bp = (double*)h->arena.p;
  loglik_sums_pass(r, f, lam_m, lam_var, N, std::exp(logA), bp, h->d_ticket, h->d_scal + kSlotAux0, st);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(h->h_out, h->d_scal + kSlotAux0, sizeof(double) * 5, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  for (int k = 0; k < 5; ++k) out_host[k] = h->h_out[k];
  return SGP_OK;
}

int sgp_estep_terms(sgp_handle* h, const double* a, long lda, const double* r, const double* f, long N, int q,
                    double logA, double* g, double* G, long ldg, void* stream) {
  if (!h || !a || !r || !f || !g || !G) return fail(h, SGP_E_ARG, "sgp_estep_terms: null argument");
  if (N <= 0 || q <= 0 || lda < q || ldg < q) return fail(h, SGP_E_ARG, "sgp_estep_terms: bad sizes / leading dimensions");
  int rc = ensure_ready(h); if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const double A = std::exp(logA);
  const long ldq = rup(q, 16);
  const long slab = std::min<long>(h->slab_rows, rup(N, 128));
  const int ozS = oz_pick(h, slab, q, q);   // G on the int8 tensor-core path: (sqrt(f) a)^T (sqrt(f) a), lower tiles
  const int nz = ozS ? oz_split(q, q, (int)std::min(slab, N), ozS, 1) : pick_split(gemm_tiles(q, q, true), slab);
  const long zs = (long)q * ldq;
  const int ncb = colsum_blocks(slab);
  const long ldkN = oz_ldk((int)slab);
  size_t need = 0;
  auto acc = [&](size_t nelem) { need = ((need + 255) & ~size_t(255)) + nelem * sizeof(double); };
  acc((size_t)slab * ldq); acc((size_t)nz * zs); acc((size_t)ncb * ldq); acc((size_t)q * ldq);
  if (ozS) { acc((size_t)((ozS + 1) / 2) * q * ldkN / 8 + 16); acc(q); acc(q); }
  need += 4096;
  RESERVE_ARENA(need);
  Arena ar(h->arena.p, h->arena.bytes);
  double* aw = ar.take<double>((size_t)slab * ldq);
  double* Gpart = ar.take<double>((size_t)nz * zs);
  double* colpart = ar.take<double>((size_t)ncb * ldq);
  double* Gw = ar.take<double>((size_t)q * ldq);
  int8_t* atq = nullptr; int* ate = nullptr; unsigned long long* amx = nullptr;
  if (ozS) {
    atq = (int8_t*)ar.take<double>((size_t)((ozS + 1) / 2) * q * ldkN / 8 + 16);
    ate = (int*)ar.take<double>(q); amx = (unsigned long long*)ar.take<double>(q);
  }
  int si = 0;
  for (long i0 = 0; i0 < N; i0 += slab, ++si) {
    const long ns = std::min(slab, N - i0);
    const double* as = a + i0 * lda; long ldas = lda;
    if (!aligned2(as, lda)) { copy_mat(as, lda, (int)ns, q, aw, ldq, st); as = aw; ldas = ldq; }
    if (ozS) {      // G += a^T diag(f) a   (U:1422)
      const int ldk_s = oz_ldk((int)ns);
      CU(oz_slice_cols(as, ldas, ns, q, ozS, atq, ldk_s, ate, amx, st, f + i0));
      CU(oz_gemm(atq, ate, q, atq, ate, q, (int)ns, ldk_s, ozS, Gpart, ldq, nz, zs, si > 0, st, nullptr, 0, 1.0, 1));
    } else {
      GemmArgs gg{};
      gg.A = as; gg.lda = ldas; gg.B = as; gg.ldb = ldas; gg.C = Gpart; gg.ldc = ldq; gg.M = q; gg.N = q; gg.K = (int)ns;
      gg.k_chunk = (int)rup((ns + nz - 1) / nz, 16); gg.zstride = zs; gg.accumulate = si > 0; gg.kscale = f + i0;
      gg.lower_only = 1;
      CU(gemm(st, 1, 1, kStore, false, nz, gg));
    }
    colsum_weighted(as, ldas, ns, q, r + i0, f + i0, colpart, ldq, st);   // g += a^T (r - f)   (U:1421)
    col_reduce(colpart, colsum_blocks(ns), q, ldq, g, si > 0, st);
  }
  slice_reduce(Gpart, ozS ? oz_eff_split((int)std::min(slab, N), nz) : nz, zs, zs, Gw, 0, st);
  mirror_lower(Gw, ldq, q, st);
  lincomb(G, ldg, q, q, A * A, Gw, ldq, 0.0, nullptr, 0, 0.0, st);
  scale_vec(g, q, A, st);
  CU(cudaGetLastError());
  return SGP_OK;
}

}  // extern "C"
