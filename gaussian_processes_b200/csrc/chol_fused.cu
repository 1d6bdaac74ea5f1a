// Cholesky factor, its inverse and log det of a q x q matrix (q <= 1024) in ONE launch: a thread-block cluster of 8 CTAs
// walks the blocked algorithm with cluster barriers where cholesky_lower / tri_inverse_lower (sgp_kernels.cu) launch a
// kernel (40 dependent launches at q = 500).  Measured (scripts/probe_chol.py): a few per cent faster than the stepwise path up
// to q ~ 320 and slower beyond (the trailing updates want more than 8 SMs) -- hence the "chol_fused" threshold in sgp_api.cu.
//
// The K_tilde_b factorisation sits on every evaluation's inducing-point chain (utils.py:2047-2067: solve(K_tilde_b, I),
// log_det) and in every E-step iteration (utils.py:1420-1439); north_star item 2 / SURVEY 7.1-5 ask for a blocked on-chip
// version.  The matrix (2 MB at q = 500) stays in L2; tiles move through shared memory.
//
//   phase A, per 64-column block j:  every CTA factors the diagonal block redundantly (same bits everywhere: no broadcast,
//            no barrier) and keeps W = L_jj^-1 in shared memory; the 64-row tiles of the panel below are dealt round-robin:
//            L21 = A21 W^T;  barrier;  the lower tiles of the trailing matrix are dealt round-robin: A22 -= L21 L21^T;  barrier.
//   phase B: X = L^-1 by forward substitution on the identity.  The columns of X are independent: CTA r owns, of every
//            64-column block, the 8 columns 8r .. 8r+7 -- 64 right-hand sides in all, whatever q is -- so the work is
//            balanced and needs no barrier: X_I = -W_I sum_{K<I} L_IK X_K with X_K kept in shared memory.
//   phase C: log det = 2 sum log L_kk (CTA 0).
// All products are 64 x 64 x 64 tiles on the fp64 tensor pipe (DMMA m8n8k4), eight warps with eight rows each.
#include <cuda_runtime.h>

#include "sgp_common.cuh"
#include "sgp_kernels.cuh"
#include "chol_diag.cuh"

namespace sgp {

namespace {

constexpr int kFB = 64;              // block size
constexpr int kFP = 68;              // shared-memory pitch (doubles): conflict-free DMMA fragment loads
constexpr int kFThreads = 256;
constexpr int kFCluster = 8;         // portable cluster size
constexpr int kFMaxN = 1024;

__device__ __forceinline__ unsigned cluster_rank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// every global write of the cluster's CTAs before the barrier is visible to every CTA of the cluster after it
__device__ __forceinline__ void cluster_sync_all() {
  __threadfence();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// 64 x 64 tile (zero-padded beyond rows x cols) global -> shared
__device__ __forceinline__ void load_tile(double* __restrict__ dst, const double* __restrict__ src, long ld, int rows,
                                          int cols) {
  for (int e = threadIdx.x; e < kFB * kFB; e += kFThreads) {
    const int r = e >> 6, c = e & 63;
    dst[r * kFP + c] = (r < rows && c < cols) ? src[(long)r * ld + c] : 0.0;
  }
}

// acc[j][0..1] (rows 8 warp + g, columns 8 j + 2 t, + 1) += A (64 x 64, shared) * B.
// BNK: B is stored [n][k] (C = A B^T);  otherwise [k][n] (C = A B).  NT: number of 8-column tiles of the product.
template <bool BNK, int NT>
__device__ __forceinline__ void tile_mma(const double* __restrict__ As, const double* __restrict__ Bs, double (&acc)[NT][2]) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
#pragma unroll 4
  for (int kk = 0; kk < kFB / 4; ++kk) {
    const int k = 4 * kk + t;
    const double a = As[(8 * warp + g) * kFP + k];
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const double b = BNK ? Bs[(8 * j + g) * kFP + k] : Bs[k * kFP + 8 * j + g];
      dmma884(acc[j][0], acc[j][1], a, b);
    }
  }
}

struct FusedArgs {
  double* A; long ld; int n;          // in: SPD matrix (lower triangle read); out: L in the lower triangle
  double* X; long ldx;                // out (optional): L^-1, strict upper triangle and padding columns zero
  double* Winv;                       // scratch / out: inverses of the diagonal blocks of L, cdiv(n,64) x 64 x 64
  int* flags; int bit_notpd, bit_small;
  double* scal; int slot;             // out (optional, slot >= 0): scal[slot] = log det A = 2 sum log L_kk
};

__global__ void __launch_bounds__(kFThreads, 1) chol_inv_cluster_kernel(const FusedArgs p) {
  extern __shared__ double fsm[];
  double* T0 = fsm;
  double* T1 = T0 + kFB * kFP;
  double* Ws = T1 + kFB * kFP;
  double* Ti = Ws + kFB * kFP;         // 8 inverses of 8 x 8 sub-blocks
  double* pan = Ti + 512;              // panel scratch of the diagonal-block factorisation
  __shared__ double red[32];
  const int rank = (int)cluster_rank();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int n = p.n, nblk = (n + kFB - 1) / kFB;
  const long ld = p.ld;

  // ---------------- phase A: blocked Cholesky ----------------
  for (int jb = 0; jb < nblk; ++jb) {
    const int j0 = jb * kFB, nb = min(kFB, n - j0), j1 = j0 + nb, rem = n - j1;
    {
      // every CTA factors the diagonal block: L in T0, W = L^-1 straight into Ws
      const int bad = diag_factor_inv<kFP>(p.A + (long)j0 * ld + j0, ld, nb, T0, Ws, Ti, pan, p.bit_notpd, p.bit_small);
      if (rank == 0) {
        if (tid == 0 && bad) atomicOr(p.flags, bad);
        // the factor goes back to A only after the next cluster barrier: the other CTAs may still be READING this
        // diagonal block for their own factorisation.  Parked in T1 (free until the trailing update).
        for (int e = tid; e < kFB * kFB; e += kFThreads) {
          const int r = e >> 6, c = e & 63;
          T1[r * kFP + c] = T0[r * kFP + c];
          p.Winv[(long)jb * kFB * kFB + e] = Ws[r * kFP + c];      // read by nobody before phase B
        }
      }
    }
    __syncthreads();
    const int ntile = (rem + kFB - 1) / kFB;
    // panel: L21 = A21 W^T, 64-row tiles dealt round-robin
    for (int ti = rank; ti < ntile; ti += kFCluster) {
      const int r0 = j1 + kFB * ti, rows = min(kFB, n - r0);
      load_tile(T0, p.A + (long)r0 * ld + j0, ld, rows, nb);
      __syncthreads();
      double acc[8][2];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j][0] = acc[j][1] = 0.0;
      tile_mma<true, 8>(T0, Ws, acc);
      const int r = 8 * warp + g;
      if (r < rows) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int c = 8 * j + 2 * t;
          if (c < nb) p.A[(long)(r0 + r) * ld + j0 + c] = acc[j][0];
          if (c + 1 < nb) p.A[(long)(r0 + r) * ld + j0 + c + 1] = acc[j][1];
        }
      }
      __syncthreads();
    }
    cluster_sync_all();
    if (rank == 0) {
      for (int e = tid; e < kFB * kFB; e += kFThreads) {
        const int r = e >> 6, c = e & 63;
        if (r < nb && c < nb) p.A[(long)(j0 + r) * ld + j0 + c] = T1[r * kFP + c];
      }
      __syncthreads();
    }
    if (rem <= 0) break;
    // trailing update: A22 -= L21 L21^T, lower tiles (i >= k) dealt round-robin
    int idx = 0;
    for (int i = 0; i < ntile; ++i)
      for (int k = 0; k <= i; ++k, ++idx) {
        if (idx % kFCluster != rank) continue;
        const int ri0 = j1 + kFB * i, rows_i = min(kFB, n - ri0);
        const int rk0 = j1 + kFB * k, rows_k = min(kFB, n - rk0);
        load_tile(T0, p.A + (long)ri0 * ld + j0, ld, rows_i, nb);
        load_tile(T1, p.A + (long)rk0 * ld + j0, ld, rows_k, nb);
        __syncthreads();
        double acc[8][2];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j][0] = acc[j][1] = 0.0;
        tile_mma<true, 8>(T0, T1, acc);
        const int r = 8 * warp + g;
        if (r < rows_i) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int c = 8 * j + 2 * t;
            double* dst = p.A + (long)(ri0 + r) * ld + rk0 + c;
            if (c < rows_k) dst[0] -= acc[j][0];
            if (c + 1 < rows_k) dst[1] -= acc[j][1];
          }
        }
        __syncthreads();
      }
    cluster_sync_all();
  }
  cluster_sync_all();      // the last diagonal block and every Winv block are in place

  // ---------------- phase C: log det ----------------
  if (rank == 0 && p.slot >= 0) {
    double s = 0.0;
    for (int k = tid; k < n; k += kFThreads) s += log(p.A[(long)k * ld + k]);
    s = block_sum(s, red);
    if (tid == 0) p.scal[p.slot] = 2.0 * s;
  }
  if (!p.X) return;

  // ---------------- phase B: X = L^-1, this CTA's 8 columns of every 64-column block ----------------
  const long ldx = p.ldx;
  for (long e = (long)rank * kFThreads + tid; e < (long)n * (ldx - n); e += (long)kFCluster * kFThreads) {
    const long r = e / (ldx - n), c = n + e % (ldx - n);
    p.X[r * ldx + c] = 0.0;                                   // padding columns
  }
  for (int ch = 0; ch * 8 < nblk; ++ch) {                    // chunks of 8 column blocks = 64 right-hand sides
    const int jbase = 8 * ch;
    // rows above the chunk's first block are zero for all its right-hand sides
    for (int e = tid; e < jbase * kFB * kFB; e += kFThreads) {
      const int row = e >> 6, cc = e & 63;
      const int col = kFB * (jbase + (cc >> 3)) + 8 * rank + (cc & 7);
      if (col < n) p.X[(long)row * ldx + col] = 0.0;
    }
    for (int I = jbase; I < nblk; ++I) {
      const int i0 = I * kFB, rows_i = min(kFB, n - i0);
      double acc[8][2];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j][0] = acc[j][1] = 0.0;
      for (int K = jbase; K < I; ++K) {
        const int k0 = K * kFB;                               // K < I <= nblk - 1: a full block
        load_tile(T0, p.A + (long)i0 * ld + k0, ld, rows_i, kFB);
        for (int e = tid; e < kFB * kFB; e += kFThreads) {    // X_K, this CTA's right-hand sides
          const int row = e >> 6, cc = e & 63;
          const int col = kFB * (jbase + (cc >> 3)) + 8 * rank + (cc & 7);
          T1[row * kFP + cc] = (col < n) ? p.X[(long)(k0 + row) * ldx + col] : 0.0;
        }
        __syncthreads();
        tile_mma<false, 8>(T0, T1, acc);
        __syncthreads();
      }
      // R = E_I - sum ;  X_I = W_I R
      {
        const int r = 8 * warp + g;
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2) {
            const int cc = 8 * j + 2 * t + h2;
            const bool one = (jbase + (cc >> 3) == I) && (r == 8 * rank + (cc & 7));
            T1[r * kFP + cc] = (one ? 1.0 : 0.0) - acc[j][h2];
          }
      }
      for (int e = tid; e < kFB * kFB; e += kFThreads) {
        const int r = e >> 6, c = e & 63;
        T0[r * kFP + c] = p.Winv[(long)I * kFB * kFB + e];
      }
      __syncthreads();
      double out[8][2];
#pragma unroll
      for (int j = 0; j < 8; ++j) out[j][0] = out[j][1] = 0.0;
      tile_mma<false, 8>(T0, T1, out);
      {
        const int r = 8 * warp + g;
        if (r < rows_i) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
#pragma unroll
            for (int h2 = 0; h2 < 2; ++h2) {
              const int cc = 8 * j + 2 * t + h2;
              const int col = kFB * (jbase + (cc >> 3)) + 8 * rank + (cc & 7);
              if (col < n) p.X[(long)(i0 + r) * ldx + col] = out[j][h2];
            }
        }
      }
      __syncthreads();      // this CTA's writes of X_I are visible to its own loads of the next I
    }
  }
}

}  // namespace

bool chol_inv_fused_supported(int n) { return n >= 1 && n <= kFMaxN; }

// A (n x n, ld): lower triangle in, L out.  X (optional): L^-1.  Winv: chol_winv_elems(n) doubles.  slot < 0: no log det.
cudaError_t chol_inv_fused(double* A, long ld, int n, double* X, long ldx, double* Winv, int* flags, int bit_notpd,
                           int bit_small, double* scal, int slot, cudaStream_t st) {
  if (!chol_inv_fused_supported(n) || ld < n || (X && ldx < n)) return cudaErrorInvalidValue;
  const size_t smem = (size_t)(3 * kFB * kFP + 1024) * sizeof(double);       // 112 640 bytes
  cudaError_t e = cudaFuncSetAttribute(chol_inv_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  FusedArgs a{A, ld, n, X, ldx, Winv, flags, bit_notpd, bit_small, scal, slot};
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(kFCluster, 1, 1);
  cfg.blockDim = dim3(kFThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kFCluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  note_launch();
  return cudaLaunchKernelEx(&cfg, chol_inv_cluster_kernel, a);
}

}  // namespace sgp
