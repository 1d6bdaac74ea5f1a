// Cholesky factor L and inverse factor W = L^-1 of one 64 x 64 diagonal block, by one CTA of 256 threads, in shared memory.
//
// Round 1 did this as 64 rank-1 steps over register tiles with one block barrier per step (every thread waits for the
// barrier, the published column, an fp64 rsqrt and 32 FMAs): 40 us per block as a call, most of a q = 500 factorisation
// whatever is done about launches.  Here the block is itself blocked 8 x 8 (30 us per block, 500 -> 407 us at q = 500):
//   per 8-column sub-block s:  warp 0 factors the 8 x 8 diagonal sub-block in registers with shuffles (no block barrier
//   inside the pivot chain) and inverts it;  all threads solve the panel below (times Ti^T) and apply the rank-8 update;
//   then W is built block row by block row from the eight 8 x 8 inverses:  W_IJ = -Ti_I sum_{J<=K<I} L_IK W_KJ.
// ~10 block barriers per sub-block instead of 64 per 8 columns, and the serial pivot chain runs inside one warp.
#pragma once
#include "sgp_common.cuh"

namespace sgp {

constexpr int kDB = 64;    // block
constexpr int kDS = 8;     // sub-block
constexpr int kPanP = 9;   // pitch of the 56 x 8 panel buffer: odd, so that a warp's column-strided reads hit 32 banks

// D, Wm: shared, kDB rows of pitch P (>= 65 recommended: conflict-free column walks); Ti: shared, 8 x 64 doubles;
// pan: shared, 64 x 8 doubles.  In: D holds nothing.  Out: D = L (upper triangle zero, identity beyond nb),
// Wm = L^-1 (same).  Returns the status bits (valid in every thread of warp 0; other threads return 0).
template <int P>
__device__ __forceinline__ int diag_factor_inv(const double* __restrict__ Ag, long ld, int nb, double* __restrict__ D,
                                               double* __restrict__ Wm, double* __restrict__ Ti, double* __restrict__ pan,
                                               int bit_notpd, int bit_small) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int e = tid; e < kDB * kDB; e += blockDim.x) {
    const int r = e >> 6, c = e & 63;
    double v = (r == c) ? 1.0 : 0.0;                          // identity padding beyond nb
    if (r < nb && c < nb) v = (c <= r) ? Ag[(long)r * ld + c] : 0.0;
    D[r * P + c] = v;
    Wm[r * P + c] = 0.0;
  }
  __syncthreads();
  int bad = 0;
#pragma unroll 1
  for (int s = 0; s < kDB / kDS; ++s) {
    const int o = kDS * s;
    if (warp == 0) {
      // ---- 8 x 8 diagonal sub-block: lanes 0..7 own a row each, pivots go round by shuffle ----
      const int r = lane & 7;
      double row[kDS], invd[kDS];
#pragma unroll
      for (int c = 0; c < kDS; ++c) row[c] = (c <= r) ? D[(o + r) * P + o + c] : 0.0;
#pragma unroll
      for (int k = 0; k < kDS; ++k) {
        const double d = __shfl_sync(0xffffffffu, row[k], k);
        const double ds = d > 0.0 ? d : 1.0;                  // keep going with a harmless pivot; the flag tells the host
        const double inv = rsqrt(ds);
        const double lkk = ds * inv;
        invd[k] = inv;                                        // 1 / l_kk, the same bits in every lane
        if (o + k < nb) {
          if (!(d > 0.0)) bad |= bit_notpd;
          else if (lkk < 1e-10) bad |= bit_small;
        }
        if (r == k) row[k] = lkk;
        else if (r > k) row[k] *= inv;
#pragma unroll
        for (int c = k + 1; c < kDS; ++c) {
          const double lc = __shfl_sync(0xffffffffu, row[k], c);
          if (r >= c) row[c] = fma(-row[k], lc, row[c]);
        }
      }
      if (lane < kDS) {
#pragma unroll
        for (int c = 0; c < kDS; ++c) D[(o + r) * P + o + c] = row[c];
      }
      __syncwarp();
      // ---- its inverse: lane c (< 8) solves T x = e_c by forward substitution ----
      if (lane < kDS) {
        const int c = lane;
        double x[kDS];
#pragma unroll
        for (int i = 0; i < kDS; ++i) {
          double acc = (i == c) ? 1.0 : 0.0;
#pragma unroll
          for (int j = 0; j < i; ++j) acc = fma(-D[(o + i) * P + o + j], x[j], acc);
          x[i] = (i >= c) ? acc * invd[i] : 0.0;
        }
#pragma unroll
        for (int i = 0; i < kDS; ++i) Ti[s * 64 + i * kDS + c] = x[i];
      }
    }
    __syncthreads();
    const int below = kDB - o - kDS;                          // rows under the sub-block
    // ---- panel: L[i][o+c] = sum_{j<=c} D[i][o+j] Ti[c][j] ----
    for (int e = tid; e < below * kDS; e += blockDim.x) {
      const int i = o + kDS + (e >> 3), c = e & 7;
      double acc = 0.0;
#pragma unroll
      for (int j = 0; j < kDS; ++j)
        if (j <= c) acc = fma(D[i * P + o + j], Ti[s * 64 + c * kDS + j], acc);
      pan[(e >> 3) * kPanP + c] = acc;
    }
    __syncthreads();
    for (int e = tid; e < below * kDS; e += blockDim.x)
      D[(o + kDS + (e >> 3)) * P + o + (e & 7)] = pan[(e >> 3) * kPanP + (e & 7)];
    // ---- rank-8 update of the lower triangle below / right of the sub-block: a row per warp, columns over the lanes ----
    for (int ii = warp; ii < below; ii += (int)(blockDim.x >> 5)) {
      double li[kDS];
#pragma unroll
      for (int j = 0; j < kDS; ++j) li[j] = pan[ii * kPanP + j];
      for (int cc = lane; cc <= ii; cc += 32) {
        double acc = D[(o + kDS + ii) * P + o + kDS + cc];
#pragma unroll
        for (int j = 0; j < kDS; ++j) acc = fma(-li[j], pan[cc * kPanP + j], acc);
        D[(o + kDS + ii) * P + o + kDS + cc] = acc;
      }
    }
    __syncthreads();
  }
  // ---- W = L^-1, block row by block row ----
#pragma unroll 1
  for (int I = 0; I < kDB / kDS; ++I) {
    const int oi = kDS * I;
    // S[r][c] = sum_{k < oi} L[oi + r][k] W[k][c]  for c < oi  (W is lower triangular: k >= c only)
    for (int e = tid; e < kDS * oi; e += blockDim.x) {
      const int r = e & 7, c = e >> 3;
      double acc0 = 0.0, acc1 = 0.0;                            // two chains: the sum is latency-bound
      for (int k = c & ~7; k < oi; k += 2) {
        acc0 = fma(D[(oi + r) * P + k], Wm[k * P + c], acc0);
        acc1 = fma(D[(oi + r) * P + k + 1], Wm[(k + 1) * P + c], acc1);
      }
      pan[r * kDB + c] = acc0 + acc1;                           // 8 x (<= 56): fits the 512-double panel buffer
    }
    __syncthreads();
    for (int e = tid; e < kDS * oi; e += blockDim.x) {
      const int r = e & 7, c = e >> 3;
      double acc = 0.0;
#pragma unroll
      for (int j = 0; j < kDS; ++j)
        if (j <= r) acc = fma(-Ti[I * 64 + r * kDS + j], pan[j * kDB + c], acc);
      Wm[(oi + r) * P + c] = acc;
    }
    for (int e = tid; e < kDS * kDS; e += blockDim.x) Wm[(oi + (e >> 3)) * P + oi + (e & 7)] = Ti[I * 64 + e];
    __syncthreads();
  }
  return bad;
}

}  // namespace sgp
