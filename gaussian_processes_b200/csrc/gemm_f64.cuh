// fp64 tensor-core GEMM family for sm_100a (DMMA.8x8x4 via mma.sync.m8n8k4.f64).
//
//   C[M x N] (op)= alpha * sum_k A(m,k) * B(n,k)          k in [k_begin(z), k_end(z))
//
// Operand layouts (template):
//   AL/BL = 0 : "K-major"  -- element (row, k) at ptr[row*ld + k]   (k contiguous)
//   AL/BL = 1 : "MN-major" -- element (row, k) at ptr[k*ld + row]   (row contiguous)
// so X*C, Z*U^T, X^T*Y (reduction over images), K^T diag(v) K ... are all this one kernel.
//
// Tiling: CTA tile BM x BN x kBK (64 x 128 x 16 for the large products, 64 x 64 x 16 for the small ones), 4 warps
// as 2(m) x 2(n); each warp owns (BM/2) x (BN/2) as (BM/16) x (BN/16) DMMA 8x8 accumulators (32 x 64 -> 32 DMMAs
// per 12 fragment loads).  3-stage cp.async pipeline, 90 KB of shared memory -> TWO independent CTAs per SM: while
// one CTA sits at its per-k-tile barrier or in its epilogue the other keeps the fp64 tensor pipe busy (one
// 8-warp CTA per SM measured 82 % DMMA-pipe activity; the library's own d884 kernels use this same shape).
// Shared-memory rows are padded (+4 doubles) so every fragment LDS.64 is bank-conflict free:
//   K-major  tile [rows][kBK+4] : addr = row*20 + k   -> (g*4 + t) mod 16 distinct per half-warp
//   MN-major tile [kBK][rows+4] : addr = k*(BM+4)+row -> (t*4 + g) mod 16 distinct per half-warp
// Fused epilogues (MODE) implement the per-image reductions of the Spatial_GP pass, see below.
#pragma once
#include "sgp_common.cuh"

namespace sgp {

enum GemmMode : int {
  kStore = 0,    // C = alpha*acc (+ C if accumulate); split-K slices go to C + z*zstride
  kRowdot = 1,   // C = acc;  part1[nt][row] = sum_col acc*aux[row][col]              (Z = X C, s = x^T C x)
  kAcos = 2,     // C = arc-cos kernel(acc + s02, rowvec[row]*colvec[col]); C2 = cos    (utils.py:982-990)
  kMoments = 3,  // C = acc;  part1 = sum_col acc*aux ; part2 = sum_col aux*colvec      (T = K Mfull^T)
  kAxpyRow = 4,  // C = acc + rowvec[row]*aux[row][col]                                 (Y = R U + w.X)
};

struct GemmArgs {
  const double* A; long lda;
  const double* B; long ldb;
  double* C; long ldc;
  int M, N, K;
  int k_chunk;            // split-K: slice z covers [z*k_chunk, min(K,(z+1)*k_chunk)); k_chunk % 16 == 0
  long zstride;           // split-K: output slice stride (elements)
  double alpha; int accumulate;
  const double* kscale;   // optional per-k scale applied to A (K^T diag(v) K)
  const double* aux; long ldaux;
  const double* rowvec; const double* colvec;
  double* part1; double* part2; long ldpart;
  double* C2; double s02;
  int lower_only;         // kStore with M == N: skip CTA tiles strictly above the diagonal (symmetric product)
};

#ifndef SGP_BK
#define SGP_BK 16
#endif
#ifndef SGP_STAGES
#define SGP_STAGES 3
#endif
constexpr int kBK = SGP_BK;          // k-depth of one pipeline stage (multiple of 16)
constexpr int kStages = SGP_STAGES;
constexpr int kGemmThreads = 128;

template <int BM, int BN>
constexpr int gemm_smem_bytes() {
  // each operand stage is sized for the larger of its two possible layouts
  constexpr int a = (BM * (kBK + 4) > kBK * (BM + 4)) ? BM * (kBK + 4) : kBK * (BM + 4);
  constexpr int b = (BN * (kBK + 4) > kBK * (BN + 4)) ? BN * (kBK + 4) : kBK * (BN + 4);
  return kStages * (a + b) * 8;
}

template <int ROWS, int LAYOUT>
__device__ __forceinline__ void load_tile(double* smem, const double* __restrict__ base, long ld, int row0,
                                          int nrows, int k0, int kend, int tid) {
  // ROWS x kBK doubles = ROWS*kBK/2 chunks of 16 bytes
  constexpr int kCpk = kBK / 2;       // chunks per row of a K-major tile
  constexpr int kChunks = ROWS * kCpk;
#pragma unroll
  for (int c0 = 0; c0 < kChunks; c0 += kGemmThreads) {
    const int c = c0 + tid;
    if (kChunks % kGemmThreads != 0 && c >= kChunks) break;
    int bytes;
    const double* src;
    double* dst;
    if (LAYOUT == 0) {
      const int r = c / kCpk, ch = c % kCpk;
      const int row = row0 + r, k = k0 + ch * 2;
      int nv = kend - k;
      nv = nv < 0 ? 0 : (nv > 2 ? 2 : nv);
      bytes = (row < nrows) ? nv * 8 : 0;
      src = bytes ? base + (long)row * ld + k : base;
      dst = smem + r * (kBK + 4) + ch * 2;
    } else {
      constexpr int kCpr = ROWS / 2;  // chunks per k-row
      const int kr = c / kCpr, ch = c % kCpr;
      const int k = k0 + kr, row = row0 + ch * 2;
      int nv = nrows - row;
      nv = nv < 0 ? 0 : (nv > 2 ? 2 : nv);
      bytes = (k < kend) ? nv * 8 : 0;
      src = bytes ? base + (long)k * ld + row : base;
      dst = smem + kr * (ROWS + 4) + ch * 2;
    }
    cp_async16(dst, src, bytes);
  }
}

template <int BM, int BN, int AL, int BL, int MODE, bool KSCALE>
__global__ void __launch_bounds__(kGemmThreads, (BN == 128 ? 2 : 3)) gemm_f64_kernel(const GemmArgs g) {
  constexpr int WMI = BM / 16;  // 8x8 accumulator tiles per warp along m
  constexpr int WNI = BN / 16;  // ... along n
  constexpr int A_STAGE = (BM * (kBK + 4) > kBK * (BM + 4)) ? BM * (kBK + 4) : kBK * (BM + 4);
  constexpr int B_STAGE = (BN * (kBK + 4) > kBK * (BN + 4)) ? BN * (kBK + 4) : kBK * (BN + 4);
  extern __shared__ __align__(16) double smem[];
  double* sA = smem;
  double* sB = smem + kStages * A_STAGE;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gq = lane >> 2, t = lane & 3;
  const int wm = warp >> 1, wn = warp & 1;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  if (MODE == kStore && g.lower_only && n0 > m0 + BM - 1) return;
  const int kbeg = blockIdx.z * g.k_chunk;
  const int kend = min(g.K, kbeg + g.k_chunk);
  const int KT = (kend - kbeg + kBK - 1) / kBK;

  double acc[WMI][WNI][2];
#pragma unroll
  for (int i = 0; i < WMI; ++i)
#pragma unroll
    for (int j = 0; j < WNI; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

  // ---- prologue: prefetch kStages-1 tiles
#pragma unroll
  for (int s = 0; s < kStages - 1; ++s) {
    if (s < KT) {
      load_tile<BM, AL>(sA + s * A_STAGE, g.A, g.lda, m0, g.M, kbeg + s * kBK, kend, tid);
      load_tile<BN, BL>(sB + s * B_STAGE, g.B, g.ldb, n0, g.N, kbeg + s * kBK, kend, tid);
    }
    cp_async_commit();
  }

  for (int kt = 0; kt < KT; ++kt) {
    cp_async_wait<kStages - 2>();
    __syncthreads();
    {  // refill the stage consumed in iteration kt-1
      const int nk = kt + kStages - 1;
      if (nk < KT) {
        const int s = nk % kStages;
        load_tile<BM, AL>(sA + s * A_STAGE, g.A, g.lda, m0, g.M, kbeg + nk * kBK, kend, tid);
        load_tile<BN, BL>(sB + s * B_STAGE, g.B, g.ldb, n0, g.N, kbeg + nk * kBK, kend, tid);
      }
      cp_async_commit();
    }
    const double* a_s = sA + (kt % kStages) * A_STAGE;
    const double* b_s = sB + (kt % kStages) * B_STAGE;
    // fragments are double-buffered in registers: the loads of k-step kk+1 are issued before the DMMAs of kk
    double af[2][WMI], bf[2][WNI];
    auto load_frags = [&](int buf, int kk) {
      const int k = kk * 4 + t;
#pragma unroll
      for (int i = 0; i < WMI; ++i) {
        const int r = wm * (BM / 2) + i * 8 + gq;
        af[buf][i] = (AL == 0) ? a_s[r * (kBK + 4) + k] : a_s[k * (BM + 4) + r];
      }
      if (KSCALE) {
        const int kg = kbeg + kt * kBK + k;
        const double sc = (kg < kend) ? __ldg(g.kscale + kg) : 0.0;
#pragma unroll
        for (int i = 0; i < WMI; ++i) af[buf][i] *= sc;
      }
#pragma unroll
      for (int j = 0; j < WNI; ++j) {
        const int c = wn * (BN / 2) + j * 8 + gq;
        bf[buf][j] = (BL == 0) ? b_s[c * (kBK + 4) + k] : b_s[k * (BN + 4) + c];
      }
    };
    load_frags(0, 0);
#pragma unroll
    for (int kk = 0; kk < kBK / 4; ++kk) {
      if (kk + 1 < kBK / 4) load_frags((kk + 1) & 1, kk + 1);
#pragma unroll
      for (int i = 0; i < WMI; ++i)
#pragma unroll
        for (int j = 0; j < WNI; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[kk & 1][i], bf[kk & 1][j]);
    }
  }
  cp_async_wait<0>();

  // ---------------------------------------------------------------- epilogue
  // accumulator (i, j, e): row = m0 + wm*BM/2 + i*8 + g ; col = n0 + wn*BN/2 + j*8 + 2t + e
  double rs1[WMI], rs2[WMI];
  if (MODE == kRowdot || MODE == kMoments) {
#pragma unroll
    for (int i = 0; i < WMI; ++i) rs1[i] = rs2[i] = 0.0;
  }
  double* Cz = g.C + (MODE == kStore ? (long)blockIdx.z * g.zstride : 0L);

#pragma unroll
  for (int i = 0; i < WMI; ++i) {
    const int row = m0 + wm * (BM / 2) + i * 8 + gq;
    if (row >= g.M) continue;
    double rv = 0.0;
    if (MODE == kAcos || MODE == kAxpyRow) rv = __ldg(g.rowvec + row);
#pragma unroll
    for (int j = 0; j < WNI; ++j) {
      const int col = n0 + wn * (BN / 2) + j * 8 + 2 * t;
      if (col >= g.N) continue;
      const bool two = (col + 1 < g.N);
      double v0 = acc[i][j][0], v1 = acc[i][j][1];
      double x0 = 0.0, x1 = 0.0;
      if (MODE == kRowdot || MODE == kMoments || MODE == kAxpyRow) {
        const double* ap = g.aux + (long)row * g.ldaux + col;
        if (two) { const double2 xx = *reinterpret_cast<const double2*>(ap); x0 = xx.x; x1 = xx.y; }
        else x0 = *ap;
      }
      if (MODE == kStore) {
        v0 *= g.alpha; v1 *= g.alpha;
        if (g.accumulate) {
          const double* cp = Cz + (long)row * g.ldc + col;
          v0 += cp[0];
          if (two) v1 += cp[1];
        }
      } else if (MODE == kRowdot) {
        rs1[i] += v0 * x0 + (two ? v1 * x1 : 0.0);
      } else if (MODE == kMoments) {
        rs1[i] += v0 * x0 + (two ? v1 * x1 : 0.0);
        rs2[i] += x0 * __ldg(g.colvec + col) + (two ? x1 * __ldg(g.colvec + col + 1) : 0.0);
      } else if (MODE == kAxpyRow) {
        v0 += rv * x0; v1 += rv * x1;
      } else if (MODE == kAcos) {
        double c0, c1 = 0.0;
        v0 = acos_kernel_value(v0 + g.s02, rv * __ldg(g.colvec + col), &c0);
        if (two) v1 = acos_kernel_value(v1 + g.s02, rv * __ldg(g.colvec + col + 1), &c1);
        double* c2p = g.C2 + (long)row * g.ldc + col;
        if (two) *reinterpret_cast<double2*>(c2p) = make_double2(c0, c1);
        else *c2p = c0;
      }
      double* cp = Cz + (long)row * g.ldc + col;
      if (two) *reinterpret_cast<double2*>(cp) = make_double2(v0, v1);
      else *cp = v0;
    }
  }

  if (MODE == kRowdot || MODE == kMoments) {
    // deterministic row reduction: quad shuffle -> 2 n-warps through smem -> one partial per (n-tile,row)
    __syncthreads();  // pipeline smem is free now
    double* red1 = smem;            // [2][BM]
    double* red2 = smem + 2 * BM;   // [2][BM]
#pragma unroll
    for (int i = 0; i < WMI; ++i) {
      double a = rs1[i], b = rs2[i];
      a += __shfl_xor_sync(0xffffffffu, a, 1);
      a += __shfl_xor_sync(0xffffffffu, a, 2);
      if (MODE == kMoments) {
        b += __shfl_xor_sync(0xffffffffu, b, 1);
        b += __shfl_xor_sync(0xffffffffu, b, 2);
      }
      if (t == 0) {
        const int rl = wm * (BM / 2) + i * 8 + gq;
        red1[wn * BM + rl] = a;
        if (MODE == kMoments) red2[wn * BM + rl] = b;
      }
    }
    __syncthreads();
    if (tid < BM && m0 + tid < g.M) {
      const long o = (long)blockIdx.x * g.ldpart + m0 + tid;
      g.part1[o] = red1[tid] + red1[BM + tid];
      if (MODE == kMoments) g.part2[o] = red2[tid] + red2[BM + tid];
    }
  }
}

// ---- host-side launcher (defined in gemm_f64.cu)
struct GemmLaunch {
  int al, bl, mode;
  bool small_tile;   // 64x64 CTA tile instead of 64x128
  int nsplit;        // grid.z (split-K); requires mode == kStore
};
cudaError_t gemm_f64_launch(const GemmArgs& args, const GemmLaunch& cfg, cudaStream_t stream);
cudaError_t gemm_f64_init();  // opt-in shared memory sizes; call once per device

}  // namespace sgp
