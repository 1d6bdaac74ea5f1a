// Per-image quadratic forms x_i^T C x_i and x_i^T dC_k x_i through the SEPARABLE structure of localker's C.
//
// localker (utils.py:880-898) builds  C = Amp * D_a (G (x) G) D_a  on the pixel grid, with
//   a_k = exp(-e^{theta_beta} |xi_k - eps_0|^2) (zero outside the receptive-field mask) and
//   G_ab = exp(-e^{theta_rho} (lin_a - lin_b)^2), lin = linspace(-1, 1, n)  -- a Kronecker product of two n x n
// Gaussian matrices, because exp(-rho((dx)^2 + (dy)^2)) factorises.  With M = (x o a) reshaped to n x n [iy][ix]:
//   Sy = M G M^T, Sx = M^T G M, Ey = Sy o G, Ex = Sx o G, rsx/rsy = row sums of Ex/Ey, D2_ab = (lin_a - lin_b)^2
//   x^T C x                    = Amp * sum(Ey)
//   x^T dC_{-log2rho2} x       = -e^{theta_rho} Amp (sum(Ex o D2) + sum(Ey o D2))                  (utils.py:909)
//   x^T dC_{eps_0x} x          = 4 e^{theta_beta} Amp sum_ix (lin_ix - eps_0x) rsx[ix]             (utils.py:904)
//   x^T dC_{eps_0y} x          = 4 e^{theta_beta} Amp sum_iy (lin_iy - eps_0y) rsy[iy]             (utils.py:905)
//   x^T dC_{-2log2beta} x      = -2 e^{theta_beta} Amp (sum (lin_ix-eps_0x)^2 rsx + sum (lin_iy-eps_0y)^2 rsy)  (U:907)
// i.e. 8 n^3 flops per image instead of the 2 p^2 (value) + 2 p^2 (adjoint) of the dense products -- 6.4 n^3 with
// the symmetry of Sy / Sx (only the tiles on and below the diagonal of the second product of each chain).
//
// One warp per image; the four n x n x n products run on the fp64 tensor pipe (DMMA m8n8k4).  The first product
// of each chain stays in registers: its C fragments are turned into A fragments with quad shuffles.
#include "sgp_common.cuh"
#include "sgp_kernels.cuh"

namespace sgp {

namespace {

constexpr int kSepWarps = 8;

__device__ __forceinline__ double lin_coord_dev(int i, int n) {   // torch.linspace(-1, 1, n), as on the host
  if (n == 1) return -1.0;
  const double step = 2.0 / (double)(n - 1);
  return (i < n / 2) ? (-1.0 + step * i) : (1.0 - step * (n - 1 - i));
}

// RB row tiles (8 rows each) of one chain.  TR = false: rows are iy (Sy = M G M^T); TR = true: rows are ix
// (Sx = M^T G M).  Adds the per-lane partial sums T0..T3 described at the end of the function.
// Two row tiles share every G / M fragment of the second operand (7 fragment loads per 10 DMMAs instead of 6 per 5).
// Image tile addressing.  SWZ = 0: padded rows (pitch 8 NT + 4 doubles: conflict-free fragment loads).  SWZ = 1 / 2: pitch
// 8 NT with the column XOR-ed by row bits instead of padding (1: NT odd, pitch = 8 mod 16 doubles, bit 2 <- row bit 1;
// 2: NT even, pitch = 0 mod 16, bits 2-3 <- row bits 0-1): the same conflict-free loads -- for the row-major and the
// transposed fragment patterns alike -- from 8 % less shared memory, which is what lets a warp hold TWO image tiles and
// have cp.async write the next image straight into its final layout (16-byte chunks stay whole: the XOR is a multiple
// of two doubles).
template <int NT, int SWZ>
__device__ __forceinline__ int midx(int row, int col) {
  if (SWZ == 0) return row * (8 * NT + 4) + col;
  if (SWZ == 1) return row * (8 * NT) + (col ^ (((row >> 1) & 1) << 2));
  return row * (8 * NT) + (col ^ ((row & 3) << 2));
}

// G: second operand of the first product; GW: the matrix the second product is weighted with (the same matrix when the
// receptive-field weights are already in M; D_x G D_x / D_y G D_y when M is the raw image, see sep_quad_full_kernel).
// I0 (first row tile of the group) is a template parameter: which tiles of the symmetric second product are live is
// then known at compile time, and groups with few live tiles (<= 5: the first pair and a trailing single row tile)
// split their accumulators by k-half (h) -- twice as many independent DMMA chains of half the length where the
// dependency latency of the chain, not the pipe, was the limiter (ncu: 44 % of the samples on fixed-latency waits).
template <int NT, bool TR, int RB, int SWZ, int I0>
__device__ __forceinline__ void chain_rows(const double* __restrict__ M, const double* __restrict__ G,
                                           const double* __restrict__ GW,
                                           const double* __restrict__ lin, const double* __restrict__ lin2, int lane,
                                           double& T0, double& T1, double& T2, double& T3) {
  constexpr int i0 = I0;
  constexpr int LIVE = (RB == 2) ? (2 * I0 + 3) : (I0 + 1);
  constexpr int NS = (LIVE <= 3) ? 4 : ((LIVE <= 7) ? 2 : 1);
  const int g = lane >> 2, t = lane & 3;
  double P[RB][NT][2];
#pragma unroll
  for (int r = 0; r < RB; ++r)
#pragma unroll
    for (int j = 0; j < NT; ++j) P[r][j][0] = P[r][j][1] = 0.0;
#pragma unroll 2
  for (int kk = 0; kk < 2 * NT; ++kk) {
    const int k = 4 * kk + t;
    double af[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r)
      af[r] = TR ? M[midx<NT, SWZ>(k, 8 * (i0 + r) + g)] : M[midx<NT, SWZ>(8 * (i0 + r) + g, k)];
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const double bf = G[midx<NT, SWZ>(k, 8 * j + g)];
#pragma unroll
      for (int r = 0; r < RB; ++r) dmma884(P[r][j][0], P[r][j][1], af[r], bf);
    }
  }
  // S = P M^T (resp. P M) is symmetric: only the tiles on and below the diagonal are formed (15 of 25 at n = 40);
  // a tile below the diagonal stands for its mirror image too, through the weights of the sums further down.
  double S[NS][RB][NT][2];
#pragma unroll
  for (int z = 0; z < NS; ++z)
#pragma unroll
    for (int r = 0; r < RB; ++r)
#pragma unroll
      for (int j = 0; j < NT; ++j) S[z][r][j][0] = S[z][r][j][1] = 0.0;
#pragma unroll
  for (int jk = 0; jk < NT; ++jk) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      // A fragment (row g, column 4h+t of tile jk) out of the C fragments (row g, columns 2t', 2t'+1)
      const int src = (lane & ~3) | (2 * h + (t >> 1));
      double af[RB];
#pragma unroll
      for (int r = 0; r < RB; ++r) {
        const double v0 = __shfl_sync(0xffffffffu, P[r][jk][0], src);
        const double v1 = __shfl_sync(0xffffffffu, P[r][jk][1], src);
        af[r] = (t & 1) ? v1 : v0;
      }
      const int k = 8 * jk + 4 * h + t;
#pragma unroll
      for (int jn = 0; jn < NT; ++jn) {
        if (jn <= i0 + RB - 1) {
          const double bf = TR ? M[midx<NT, SWZ>(k, 8 * jn + g)] : M[midx<NT, SWZ>(8 * jn + g, k)];
#pragma unroll
          for (int r = 0; r < RB; ++r)
            if (jn <= i0 + r) dmma884(S[(2 * jk + h) % NS][r][jn][0], S[(2 * jk + h) % NS][r][jn][1], af[r], bf);
        }
      }
    }
  }
  // With w = 1/2 on the diagonal tiles and 1 below (a tile below the diagonal also stands for its mirror image, a
  // diagonal tile holds both (a,b) and (b,a)) and E = S o G:
  //   T0 += w E ; T1 += w E (l_a - l_b)^2 ; T2 += w E (l_a + l_b) ; T3 += w E (l_a^2 + l_b^2)
  // so that over the whole matrix  sum E = 2 T0,  sum E o D2 = 2 T1,  sum_a (l_a - c) rs_a = T2 - 2 c T0,
  // sum_a (l_a - c)^2 rs_a = T3 - 2 c T2 + 2 c^2 T0   (rs = row sums of E; c = the receptive-field centre).
#pragma unroll
  for (int r = 0; r < RB; ++r) {
    const int a = 8 * (i0 + r) + g;
    const double la = lin[a], la2 = lin2[a];
#pragma unroll
    for (int jn = 0; jn < NT; ++jn) {
      if (jn <= i0 + r) {
        const double w = (jn == i0 + r) ? 0.5 : 1.0;
        const int b = 8 * jn + 2 * t;
        const double2 gg = *reinterpret_cast<const double2*>(GW + midx<NT, SWZ>(a, b));
        const double2 lb = *reinterpret_cast<const double2*>(lin + b);
        const double2 lb2 = *reinterpret_cast<const double2*>(lin2 + b);
        double s0 = S[0][r][jn][0], s1 = S[0][r][jn][1];
#pragma unroll
        for (int z = 1; z < NS; ++z) { s0 += S[z][r][jn][0]; s1 += S[z][r][jn][1]; }
        const double e0 = s0 * (gg.x * w), e1 = s1 * (gg.y * w);
        const double d0 = la - lb.x, d1 = la - lb.y;
        T0 += e0 + e1;
        T1 = fma(e0 * d0, d0, fma(e1 * d1, d1, T1));
        T2 = fma(e0, la + lb.x, fma(e1, la + lb.y, T2));
        T3 = fma(e0, la2 + lb2.x, fma(e1, la2 + lb2.y, T3));
      }
    }
  }
}

template <int NT, bool TR, int SWZ, int I0>
__device__ __forceinline__ void chain_from(const double* __restrict__ M, const double* __restrict__ G,
                                           const double* __restrict__ GW, const double* __restrict__ lin,
                                           const double* __restrict__ lin2, int lane, double& T0, double& T1,
                                           double& T2, double& T3) {
  if constexpr (I0 + 1 < NT) {
    chain_rows<NT, TR, 2, SWZ, I0>(M, G, GW, lin, lin2, lane, T0, T1, T2, T3);
    chain_from<NT, TR, SWZ, I0 + 2>(M, G, GW, lin, lin2, lane, T0, T1, T2, T3);
  } else if constexpr (I0 < NT) {
    chain_rows<NT, TR, 1, SWZ, I0>(M, G, GW, lin, lin2, lane, T0, T1, T2, T3);
  }
}

template <int NT, bool TR, int SWZ = 0>
__device__ __forceinline__ void chain(const double* __restrict__ M, const double* __restrict__ G,
                                      const double* __restrict__ GW,
                                      const double* __restrict__ lin, const double* __restrict__ lin2, int lane,
                                      double& T0, double& T1, double& T2, double& T3) {
  chain_from<NT, TR, SWZ, 0>(M, G, GW, lin, lin2, lane, T0, T1, T2, T3);
}

// Per-image results from the eight per-lane sums of the two chains (see chain_rows).
__device__ __forceinline__ void sep_store(const SepQuadArgs& a, long img, int lane, double y0, double y1, double y2,
                                          double y3, double x0, double x1, double x2, double x3) {
  const double base = warp_sum(2.0 * y0);
  const double dd = warp_sum(2.0 * (x1 + y1));
  const double ax1 = warp_sum(x2 - 2.0 * a.ex * x0);
  const double ay1 = warp_sum(y2 - 2.0 * a.ey * y0);
  const double a2 = warp_sum((x3 - 2.0 * a.ex * x2 + 2.0 * a.ex * a.ex * x0) +
                             (y3 - 2.0 * a.ey * y2 + 2.0 * a.ey * a.ey * y0));
  if (lane == 0) {
    const double sq = a.amp * base;          // x^T C x
    const double s = sq + a.s02;
    a.s[img] = s;
    a.nrm[img] = sqrt(s);
    if (a.Q) {
      double* q = a.Q + img * 5;             // units of the closure's gradient slots (see finalize_kernel)
      q[0] = sq;                             // Amp        : <C, x x^T>
      q[1] = -2.0 * a.eb * a.amp * a2;       // -2log2beta
      q[2] = -a.er * a.amp * dd;             // -log2rho2
      q[3] = 2.0 * a.amp * ax1;              // eps_0x / (2 e^{theta_beta})
      q[4] = 2.0 * a.amp * ay1;              // eps_0y / (2 e^{theta_beta})
    }
  }
}

// pstage > 0: every warp owns a staging buffer of pstage doubles and the NEXT image's row is brought in with cp.async
// while the current image is being multiplied (the plain loads otherwise show up as long-scoreboard stalls).
template <int NT>
__global__ void __launch_bounds__(kSepWarps * 32, 1) sep_quad_kernel(const SepQuadArgs a, const int pstage) {
  constexpr int NP = 8 * NT, LD = NP + 4;
  extern __shared__ __align__(16) double sm[];
  double* sG = sm;                 // NP x LD
  double* sLin = sG + NP * LD;     // NP coordinates, then NP squares
  double* sLin2 = sLin + NP;
  double* sAx = sLin + 2 * NP;     // separable receptive-field weights: alpha[iy][ix] = sAy[iy] * sAx[ix]
  double* sAy = sLin + 3 * NP;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* sM = sLin + 4 * NP + (size_t)warp * NP * LD;

  for (int e = tid; e < NP; e += blockDim.x) {
    const double l = (e < a.n) ? lin_coord_dev(e, a.n) : 0.0;
    sLin[e] = l;
    sLin2[e] = l * l;
    sAx[e] = (e < a.n) ? exp(-a.eb * (l - a.ex) * (l - a.ex)) : 0.0;
    sAy[e] = (e < a.n) ? exp(-a.eb * (l - a.ey) * (l - a.ey)) : 0.0;
  }
  for (int e = tid; e < NP * LD; e += blockDim.x) {
    const int r = e / LD, c = e % LD;
    double v = 0.0;
    if (r < a.n && c < a.n) {
      const double d = lin_coord_dev(c, a.n) - lin_coord_dev(r, a.n);
      v = exp(-a.er * (d * d));
    }
    sG[e] = v;
  }
  for (int e = lane; e < NP * LD; e += 32) sM[e] = 0.0;   // pixels outside the mask / padding stay zero
  __syncthreads();

  double* sStage = sLin + 4 * NP + (size_t)kSepWarps * NP * LD + (size_t)warp * pstage;
  const long stride = (long)gridDim.x * kSepWarps;
  auto prefetch = [&](long img) {
    const double* xr = a.X + img * a.ldx;
    for (int c = 2 * lane; c < a.p; c += 64) cp_async16(sStage + c, xr + c, (a.p - c >= 2) ? 16 : 8);
    cp_async_commit();
  };
  long img = (long)blockIdx.x * kSepWarps + warp;
  if (pstage > 0 && img < a.nrows) prefetch(img);
  for (; img < a.nrows; img += stride) {
    if (pstage > 0) {
      cp_async_wait<0>();
      __syncwarp();
      if (a.full && (a.n & 1) == 0) {
        // every pixel kept: a row-by-row copy (two pixels per lane) with the separable weights from shared memory --
        // no per-pixel offset / weight loads from global memory (they were 15 % of the kernel's stall samples)
        const int half = a.n >> 1;
        for (int e = lane; e < a.n * half; e += 32) {
          const int r = e / half, c2 = 2 * (e - r * half);
          const double2 xv = *reinterpret_cast<const double2*>(sStage + r * a.n + c2);
          const double2 ax = *reinterpret_cast<const double2*>(sAx + c2);
          const double ay = sAy[r];
          *reinterpret_cast<double2*>(sM + r * LD + c2) = make_double2(xv.x * (ay * ax.x), xv.y * (ay * ax.y));
        }
      } else {
        // scatter into the n x n layout with the receptive-field weights; five pixels per lane in flight
        int c = lane;
        for (; c + 128 < a.p; c += 160) {
          int o[5]; double v[5];
#pragma unroll
          for (int u = 0; u < 5; ++u) { o[u] = __ldg(a.off + c + 32 * u); v[u] = sStage[c + 32 * u] * __ldg(a.alpha + c + 32 * u); }
#pragma unroll
          for (int u = 0; u < 5; ++u) sM[o[u]] = v[u];
        }
        for (; c < a.p; c += 32) sM[a.off[c]] = sStage[c] * a.alpha[c];
      }
      __syncwarp();
      if (img + stride < a.nrows) prefetch(img + stride);
    } else {
      const double* xr = a.X + img * a.ldx;
      for (int c = lane; c < a.p; c += 32) sM[a.off[c]] = xr[c] * a.alpha[c];
    }
    __syncwarp();
    double y0 = 0.0, y1 = 0.0, y2 = 0.0, y3 = 0.0, x0 = 0.0, x1 = 0.0, x2 = 0.0, x3 = 0.0;
    chain<NT, false>(sM, sG, sG, sLin, sLin2, lane, y0, y1, y2, y3);
    chain<NT, true>(sM, sG, sG, sLin, sLin2, lane, x0, x1, x2, x3);
    __syncwarp();
    sep_store(a, img, lane, y0, y1, y2, y3, x0, x1, x2, x3);
  }
}

// Every pixel kept (dense receptive field), n even, rows 16-byte aligned: the receptive-field weights are separable,
// alpha[iy][ix] = ay[iy] ax[ix], so they move out of the image and into the two Gaussian matrices:
//   Sy o G = (X Gx' X^T) o Gy',  Sx o G = (X^T Gy' X) o Gx',   Gx' = D_ax G D_ax,  Gy' = D_ay G D_ay
// and M is the RAW image.  cp.async writes it straight into the swizzled tile (no staging buffer, no scatter pass: the
// scatter and its offset / weight loads were 13 % of the other kernel), two tiles per warp so that the next image
// streams in under the whole of the current image's 800 DMMAs.
template <int NT, int SWZ>
__global__ void __launch_bounds__(kSepWarps * 32, 1) sep_quad_full_kernel(const SepQuadArgs a) {
  constexpr int NP = 8 * NT;
  extern __shared__ __align__(16) double sm[];
  double* sGx = sm;                // NP x NP each, same swizzled layout as the image tiles
  double* sGy = sGx + NP * NP;
  double* sLin = sGy + NP * NP;
  double* sLin2 = sLin + NP;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* sMw = sLin + 2 * NP + (size_t)warp * 2 * NP * NP;    // two NP x NP tiles

  for (int e = tid; e < NP; e += blockDim.x) {
    const double l = (e < a.n) ? lin_coord_dev(e, a.n) : 0.0;
    sLin[e] = l;
    sLin2[e] = l * l;
  }
  for (int e = tid; e < NP * NP; e += blockDim.x) {
    const int r = e / NP, c = e % NP;
    double vx = 0.0, vy = 0.0;
    if (r < a.n && c < a.n) {
      const double lr = lin_coord_dev(r, a.n), lc = lin_coord_dev(c, a.n);
      const double d = lc - lr;
      const double gv = exp(-a.er * (d * d));
      vx = gv * (exp(-a.eb * (lr - a.ex) * (lr - a.ex)) * exp(-a.eb * (lc - a.ex) * (lc - a.ex)));
      vy = gv * (exp(-a.eb * (lr - a.ey) * (lr - a.ey)) * exp(-a.eb * (lc - a.ey) * (lc - a.ey)));
    }
    sGx[midx<NT, SWZ>(r, c)] = vx;
    sGy[midx<NT, SWZ>(r, c)] = vy;
  }
  for (int e = lane; e < 2 * NP * NP; e += 32) sMw[e] = 0.0;   // padding rows / columns stay zero
  __syncthreads();

  const long stride = (long)gridDim.x * kSepWarps;
  const int half = a.n >> 1;
  auto prefetch = [&](long img, double* dst) {
    const double* xr = a.X + img * a.ldx;
    int r = lane / half, c = lane - r * half;                  // 16-byte chunk c of row r
    for (int e = lane; e < a.n * half; e += 32) {
      cp_async16(dst + midx<NT, SWZ>(r, 2 * c), xr + 2 * e, 16);
      c += 32;
      while (c >= half) { c -= half; ++r; }
    }
    cp_async_commit();
  };
  long img = (long)blockIdx.x * kSepWarps + warp;
  int buf = 0;
  if (img < a.nrows) prefetch(img, sMw);
  for (; img < a.nrows; img += stride, buf ^= 1) {
    cp_async_wait<0>();
    __syncwarp();
    if (img + stride < a.nrows) prefetch(img + stride, sMw + (buf ^ 1) * NP * NP);
    const double* sM = sMw + buf * NP * NP;
    double y0 = 0.0, y1 = 0.0, y2 = 0.0, y3 = 0.0, x0 = 0.0, x1 = 0.0, x2 = 0.0, x3 = 0.0;
    chain<NT, false, SWZ>(sM, sGx, sGy, sLin, sLin2, lane, y0, y1, y2, y3);
    chain<NT, true, SWZ>(sM, sGy, sGx, sLin, sLin2, lane, x0, x1, x2, x3);
    __syncwarp();
    sep_store(a, img, lane, y0, y1, y2, y3, x0, x1, x2, x3);
  }
}

template <int NT>
cudaError_t launch_sep(const SepQuadArgs& a, cudaStream_t st) {
  constexpr int NP = 8 * NT, LD = NP + 4;
  if (a.full && a.p == a.n * a.n && (a.n & 1) == 0 && (a.ldx % 2 == 0) && (reinterpret_cast<uintptr_t>(a.X) % 16 == 0)) {
    constexpr int SWZ = (NT & 1) ? 1 : 2;
    const size_t smem_f = (size_t)(2 * NP * NP + 2 * NP + kSepWarps * 2 * NP * NP) * sizeof(double);
    if (smem_f <= 232448) {
      cudaError_t e = cudaFuncSetAttribute(sep_quad_full_kernel<NT, SWZ>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)smem_f);
      if (e != cudaSuccess) return e;
      long nb = (a.nrows + kSepWarps - 1) / kSepWarps;
      const long cap = a.max_ctas > 0 ? a.max_ctas : num_sms();
      if (nb > cap) nb = cap;
      note_launch();
      sep_quad_full_kernel<NT, SWZ><<<(unsigned)nb, kSepWarps * 32, smem_f, st>>>(a);
      return cudaGetLastError();
    }
  }
  size_t smem = (size_t)(NP * LD + 4 * NP + kSepWarps * NP * LD) * sizeof(double);
  // staging buffers for the cp.async prefetch, if the rows are 16-byte aligned and everything fits in 227 KB
  int pstage = (a.p + 1) & ~1;
  const bool aligned = (a.ldx % 2 == 0) && (reinterpret_cast<uintptr_t>(a.X) % 16 == 0);
  if (!aligned || smem + (size_t)kSepWarps * pstage * sizeof(double) > 232448) pstage = 0;
  smem += (size_t)kSepWarps * pstage * sizeof(double);
  cudaError_t e = cudaFuncSetAttribute(sep_quad_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  long nb = (a.nrows + kSepWarps - 1) / kSepWarps;
  const long cap = a.max_ctas > 0 ? a.max_ctas : num_sms();   // fewer CTAs leave SMs to a concurrent stream
  if (nb > cap) nb = cap;
  note_launch();
  sep_quad_kernel<NT><<<(unsigned)nb, kSepWarps * 32, smem, st>>>(a, pstage);
  return cudaGetLastError();
}

}  // namespace

int sep_quad_tiles(int n_px) { return (n_px + 7) / 8; }
bool sep_quad_supported(int n_px) { return n_px >= 1 && n_px <= 8 * kSepMaxTiles; }
int sep_quad_ld(int n_px) { return 8 * sep_quad_tiles(n_px) + 4; }

cudaError_t sep_quad(const SepQuadArgs& a, cudaStream_t st) {
  if (a.nrows <= 0) return cudaSuccess;
  switch (sep_quad_tiles(a.n)) {
    case 1: return launch_sep<1>(a, st);
    case 2: return launch_sep<2>(a, st);
    case 3: return launch_sep<3>(a, st);
    case 4: return launch_sep<4>(a, st);
    case 5: return launch_sep<5>(a, st);
    case 6: return launch_sep<6>(a, st);
    default: return cudaErrorInvalidValue;
  }
}

// ZU = U C and t_j = u_j^T C u_j + sigma_0^2 for the m inducing images through the same Kronecker structure, without C:
//   (C u)_a = Amp a_a [G (u o a as an n x n image) G]_a        (C = Amp D_a (G (x) G) D_a, G symmetric)
// -- 4 n^3 flops per image instead of the 2 p^2 of the dense product, and, more to the point, ONE launch in front of the
// image pass instead of build_C -> slices of U -> slices of C -> U C product -> norms (five dependent launches, ~0.15 ms of
// every evaluation's critical path; C itself is only needed again by the tail's <dC, G_C> contraction and is built on the
// chain stream).  One block per inducing image, plain fp64 FMAs (m images: ~0.1 GFLOP in all).
__global__ void __launch_bounds__(256) sep_apply_C_kernel(const double* __restrict__ U, long ldu, int m, int p,
                                                          const int* __restrict__ off, int off_ld,
                                                          const double* __restrict__ alpha, int n, double er, double amp,
                                                          double s02, double* __restrict__ ZU, long ldz,
                                                          double* __restrict__ t_out, double* __restrict__ nrm_out) {
  extern __shared__ double sm_apply[];
  const int pitch = n + 1;
  double* G = sm_apply;                 // n x n Gaussian matrix of the grid coordinates
  double* M = G + n * pitch;            // the weighted image, later Z = G T
  double* T = M + n * pitch;            // T = M G
  __shared__ double red[8];
  const int j = blockIdx.x;
  const double* u = U + (long)j * ldu;
  for (int e = threadIdx.x; e < n * n; e += blockDim.x) {
    const int a = e / n, b = e % n;
    const double d = lin_coord_dev(a, n) - lin_coord_dev(b, n);
    G[a * pitch + b] = exp(-er * d * d);
    M[a * pitch + b] = 0.0;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < p; k += blockDim.x) {
    const int o = off[k], iy = o / off_ld, ix = o % off_ld;
    M[iy * pitch + ix] = alpha[k] * u[k];
  }
  __syncthreads();
  for (int e = threadIdx.x; e < n * n; e += blockDim.x) {      // T = M G
    const int iy = e / n, jx = e % n;
    double acc = 0.0;
    for (int k = 0; k < n; ++k) acc = fma(M[iy * pitch + k], G[k * pitch + jx], acc);
    T[iy * pitch + jx] = acc;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < n * n; e += blockDim.x) {      // Z = G T  (into M)
    const int iy = e / n, jx = e % n;
    double acc = 0.0;
    for (int k = 0; k < n; ++k) acc = fma(G[iy * pitch + k], T[k * pitch + jx], acc);
    M[iy * pitch + jx] = acc;
  }
  __syncthreads();
  double t = 0.0;
  for (int k = threadIdx.x; k < ldz; k += blockDim.x) {
    double z = 0.0;
    if (k < p) {
      const int o = off[k], iy = o / off_ld, ix = o % off_ld;
      z = amp * alpha[k] * M[iy * pitch + ix];
      t = fma(u[k], z, t);
    }
    ZU[(long)j * ldz + k] = z;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = t;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];
    s += s02;
    t_out[j] = s;
    nrm_out[j] = sqrt(s);
  }
}

cudaError_t sep_apply_C(const double* U, long ldu, int m, int p, const int* off, const double* alpha, int n, double er,
                        double amp, double s02, double* ZU, long ldz, double* t_out, double* nrm_out, cudaStream_t st) {
  if (m <= 0) return cudaSuccess;
  if (!sep_quad_supported(n)) return cudaErrorInvalidValue;
  const size_t smem = (size_t)3 * n * (n + 1) * sizeof(double);
  if (smem > 48 * 1024) {            // 3 * 48 * 49 * 8 = 56 KB at the largest n: above the default limit (per device)
    cudaError_t e = cudaFuncSetAttribute(sep_apply_C_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  note_launch();
  sep_apply_C_kernel<<<m, 256, smem, st>>>(U, ldu, m, p, off, sep_quad_ld(n), alpha, n, er, amp, s02, ZU, ldz, t_out,
                                            nrm_out);
  return cudaGetLastError();
}

// blockpart[blk*5 + k] = sum_{i in blk} w_i Q[i][k]
__global__ void __launch_bounds__(256) wq_reduce_kernel(const double* __restrict__ w, const double* __restrict__ Q,
                                                        long n, double* __restrict__ blockpart) {
  __shared__ double red[32];
  double acc[5] = {0, 0, 0, 0, 0};
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const double wi = w[i];
#pragma unroll
    for (int k = 0; k < 5; ++k) acc[k] += wi * Q[i * 5 + k];
  }
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const double sres = block_sum(acc[k], red);
    if (threadIdx.x == 0) blockpart[(long)blockIdx.x * 5 + k] = sres;
  }
}
int wq_reduce_blocks(long n) {
  long nb = (n + 1023) / 1024;
  return (int)(nb < 1 ? 1 : (nb > 296 ? 296 : nb));
}
void wq_reduce(const double* w, const double* Q, long n, double* blockpart, cudaStream_t st) {
  note_launch();
  wq_reduce_kernel<<<wq_reduce_blocks(n), 256, 0, st>>>(w, Q, n, blockpart);
}

}  // namespace sgp
