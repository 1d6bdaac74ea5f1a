// Cholesky factor L and inverse factor W = L^-1 of one 64 x 64 diagonal block, by one CTA of 256 threads, in shared memory.
//
// Round 1 did this as 64 rank-1 steps over register tiles with one block barrier per step (every thread waits for the
// barrier, the published column, an fp64 rsqrt and 32 FMAs): 40 us per block as a call, most of a q = 500 factorisation
// whatever is done about launches.  Here the block is itself blocked 8 x 8 (25 us per block, 500 -> 388 us at q = 500):
//   per 8-column sub-block s:  warp 0 factors the 8 x 8 diagonal sub-block in registers with shuffles (no block barrier
//   inside the pivot chain) and inverts it;  all threads solve the panel below (times Ti^T) and apply the rank-8 update;
//   then W is merged from the eight 8 x 8 inverses by divide and conquer (8 -> 16 -> 32 -> 64: W21 = -W22 L21 W11).
// ~4 block barriers per sub-block instead of 64 per 8 columns, and the serial pivot chain runs inside one warp.
// Cycle counters of one block (52 k cycles): pivot chains 20 k (the floor of this design: 64 x ~310 cycles of
// shuffle -> rsqrt -> scale -> shuffle -> FMA), the 8 x 8 inverses 3 k, panels 5 k, rank-8 updates 11 k and the merge of W
// 11 k (both bound by shared-memory reads: one load per FMA), load 2 k.
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
  {   // blockDim.x == 256: 16 elements per thread, all loads in flight before the first use
    double v[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int e = tid + 256 * i, r = e >> 6, c = e & 63;
      v[i] = (r < nb && c < nb && c <= r) ? Ag[(long)r * ld + c] : ((r == c && r >= nb) ? 1.0 : 0.0);   // identity beyond nb
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int e = tid + 256 * i, r = e >> 6, c = e & 63;
      D[r * P + c] = v[i];
      Wm[r * P + c] = 0.0;
    }
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
    {
      const int cc = tid & 63, rg = tid >> 6;                 // 4 row groups x 64 columns
      if (cc < below) {
        double lc[kDS];
#pragma unroll
        for (int j = 0; j < kDS; ++j) lc[j] = pan[cc * kPanP + j];
        const int first = cc + ((rg - cc) & 3);               // smallest row >= cc in this thread's residue class
        // up to 14 rows per thread, all chains independent: the update is latency-bound (8 warps, dependent fp64 FMAs)
        double acc[14];
#pragma unroll
        for (int t = 0; t < 14; ++t) {
          const int ii = first + 4 * t;
          acc[t] = (ii < below) ? D[(o + kDS + ii) * P + o + kDS + cc] : 0.0;
        }
#pragma unroll
        for (int j = 0; j < kDS; ++j) {
#pragma unroll
          for (int t = 0; t < 14; ++t) {
            const int ii = first + 4 * t;
            if (ii < below) acc[t] = fma(-pan[ii * kPanP + j], lc[j], acc[t]);
          }
        }
#pragma unroll
        for (int t = 0; t < 14; ++t) {
          const int ii = first + 4 * t;
          if (ii < below) D[(o + kDS + ii) * P + o + kDS + cc] = acc[t];
        }
      }
    }
    __syncthreads();
  }
  // ---- W = L^-1 by divide and conquer: [L11 0; L21 L22]^-1 = [W11 0; -W22 L21 W11, W22], merging 8 -> 16 -> 32 -> 64.
  //      Three levels of two fully parallel products each instead of eight sequential block rows.  T = L21 W11 is parked
  //      in the (zero) upper-right block of its pair inside Wm and cleared afterwards. ----
  for (int e = tid; e < kDB * kDS; e += blockDim.x) {          // level 0: the eight 8 x 8 inverses on the diagonal
    const int I = e >> 6, r = (e >> 3) & 7, c = e & 7;
    Wm[(kDS * I + r) * P + kDS * I + c] = Ti[I * 64 + r * kDS + c];
  }
  __syncthreads();
#pragma unroll 1
  for (int h = kDS; h < kDB; h *= 2) {
    const int sh = (h == 8) ? 3 : ((h == 16) ? 4 : 5);         // log2 h
    // T[r][c] = sum_{k >= c} L21[r][k] W11[k][c]
    for (int e = tid; e < 32 * h; e += blockDim.x) {           // 64 / (2h) pairs x h x h elements
      const int pr = e >> (2 * sh), r = (e >> sh) & (h - 1), c = e & (h - 1), b = 2 * h * pr;
      // full range of k (W11 is lower triangular: the terms k < c are exact zeros) so that the loop unrolls: the sum is
      // latency-bound, four independent chains
      double a4[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll 2
      for (int k = 0; k < h; k += 4) {
#pragma unroll
        for (int u = 0; u < 4; ++u) a4[u] = fma(D[(b + h + r) * P + b + k + u], Wm[(b + k + u) * P + b + c], a4[u]);
      }
      Wm[(b + r) * P + b + h + c] = (a4[0] + a4[1]) + (a4[2] + a4[3]);   // parked in the pair's upper-right block
    }
    __syncthreads();
    // W21[r][c] = -sum_{j <= r} W22[r][j] T[j][c]
    for (int e = tid; e < 32 * h; e += blockDim.x) {
      const int pr = e >> (2 * sh), r = (e >> sh) & (h - 1), c = e & (h - 1), b = 2 * h * pr;
      double a4[4] = {0.0, 0.0, 0.0, 0.0};                     // W22[r][j] = 0 for j > r (exact zeros): full range again
#pragma unroll 2
      for (int j = 0; j < h; j += 4) {
#pragma unroll
        for (int u = 0; u < 4; ++u) a4[u] = fma(Wm[(b + h + r) * P + b + h + j + u], Wm[(b + j + u) * P + b + h + c], a4[u]);
      }
      Wm[(b + h + r) * P + b + c] = -((a4[0] + a4[1]) + (a4[2] + a4[3]));
    }
    __syncthreads();
    for (int e = tid; e < 32 * h; e += blockDim.x) {           // clear the parking block
      const int pr = e >> (2 * sh), r = (e >> sh) & (h - 1), c = e & (h - 1), b = 2 * h * pr;
      Wm[(b + r) * P + b + h + c] = 0.0;
    }
    __syncthreads();
  }
  return bad;
}

}  // namespace sgp
