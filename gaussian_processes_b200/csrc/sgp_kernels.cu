// Non-GEMM kernels of the Spatial_GP ELBO+gradient pass (sm_100a).  See sgp_kernels.cuh.
#include "sgp_kernels.cuh"
#include "chol_diag.cuh"

#include <atomic>

#include "gemm_f64.cuh"
#include "sgp_common.cuh"

namespace sgp {

static inline int cdiv(long a, long b) { return (int)((a + b - 1) / b); }

static std::atomic<long> g_launches{0};
void note_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
long launch_counter() { return g_launches.load(std::memory_order_relaxed); }
int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// =====================================================================================================
// a1: localker (reference utils.py:880-910)
// =====================================================================================================
__global__ void build_C_kernel(double* __restrict__ C, long ldc, int p, const double* __restrict__ gx,
                               const double* __restrict__ gy, ThetaDev th, double* __restrict__ dC, long dcs) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y;
  if (j >= ldc) return;
  double c = 0.0;
  double d[5] = {0, 0, 0, 0, 0};
  if (i < p && j < p) {
    const double xi = gx[i], yi = gy[i], xj = gx[j], yj = gy[j];
    const double lai = -th.eb * ((xi - th.ex) * (xi - th.ex) + (yi - th.ey) * (yi - th.ey));
    const double laj = -th.eb * ((xj - th.ex) * (xj - th.ex) + (yj - th.ey) * (yj - th.ey));
    const double ai = exp(lai), aj = exp(laj);
    const double ls = -th.er * ((xj - xi) * (xj - xi) + (yj - yi) * (yj - yi));
    const double e = exp(ls);
    const double cij = ((th.amp * ai) * e) * aj;   // U:895 evaluation order
    const double cji = ((th.amp * aj) * e) * ai;
    c = (cij + cji) / 2;                            // U:898
    if (dC) {
      d[0] = c / th.amp;                                   // Amp        U:902
      d[1] = c * (lai + laj);                              // -2log2beta U:907
      d[2] = c * ls;                                       // -log2rho2  U:909
      d[3] = 2.0 * th.eb * c * (xi + xj - 2 * th.ex);      // eps_0x     U:904
      d[4] = 2.0 * th.eb * c * (yi + yj - 2 * th.ey);      // eps_0y     U:905
    }
  }
  if (i < p) {
    C[(long)i * ldc + j] = c;
    if (dC) {
#pragma unroll
      for (int k = 0; k < 5; ++k) dC[k * dcs + (long)i * ldc + j] = d[k];
    }
  }
}

void build_C(double* C, long ldc, int p, const double* gx, const double* gy, const ThetaDev& th, double* dC,
             long dcs, cudaStream_t st) {
  dim3 grid(cdiv(ldc, 128), (unsigned)p);   // p rows x ldc columns (column padding zeroed)
  note_launch(); build_C_kernel<<<grid, 128, 0, st>>>(C, ldc, p, gx, gy, th, dC, dcs);
}

__global__ void gather_cols_kernel(const double* __restrict__ X, long ldx, long nrows, const int* __restrict__ idx,
                                   int p, double* __restrict__ out, long ldo) {
  const long row = blockIdx.y;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < ldo; c += gridDim.x * blockDim.x)
    out[row * ldo + c] = (c < p) ? X[row * ldx + idx[c]] : 0.0;
}

// same gather from FLOAT32 images (sgp_elbo_grad_f32): the fp64 working copy is made here, once per image set
__global__ void gather_cols_f32_kernel(const float* __restrict__ X, long ldx, long nrows, const int* __restrict__ idx,
                                       int p, double* __restrict__ out, long ldo) {
  const long row = blockIdx.y;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < ldo; c += gridDim.x * blockDim.x)
    out[row * ldo + c] = (c < p) ? (double)X[row * ldx + idx[c]] : 0.0;
}

void gather_cols_f32(const float* X, long ldx, long nrows, const int* idx, int p, double* out, long ldo,
                     cudaStream_t st) {
  if (nrows <= 0) return;
  const int gx = cdiv(ldo, 256) > 8 ? 8 : (int)cdiv(ldo, 256);
  for (long r0 = 0; r0 < nrows; r0 += 65535) {
    const long nr = (nrows - r0) < 65535 ? (nrows - r0) : 65535;
    note_launch();
    gather_cols_f32_kernel<<<dim3(gx, (unsigned)nr), 256, 0, st>>>(X + r0 * ldx, ldx, nr, idx, p, out + r0 * ldo, ldo);
  }
}

void gather_cols(const double* X, long ldx, long nrows, const int* idx, int p, double* out, long ldo,
                 cudaStream_t st) {
  if (nrows <= 0) return;
  const long maxy = 65535;
  for (long r0 = 0; r0 < nrows; r0 += maxy) {
    const long nr = (nrows - r0) < maxy ? (nrows - r0) : maxy;
    dim3 grid(cdiv(ldo, 256) > 8 ? 8 : cdiv(ldo, 256), (unsigned)nr);
    note_launch(); gather_cols_kernel<<<grid, 256, 0, st>>>(X + r0 * ldx, ldx, nr, idx, p, out + r0 * ldo, ldo);
  }
}

__global__ void norms_finalize_kernel(const double* __restrict__ part, long ldpart, int nparts, long n, double s02,
                                      double* __restrict__ s, double* __restrict__ nrm) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double a = 0.0;
  for (int t = 0; t < nparts; ++t) a += part[t * ldpart + i];
  a += s02;
  s[i] = a;
  nrm[i] = sqrt(a);
}

void norms_finalize(const double* part, long ldpart, int nparts, long n, double s02, double* s, double* nrm,
                    cudaStream_t st) {
  if (n <= 0) return;
  note_launch(); norms_finalize_kernel<<<cdiv(n, 256), 256, 0, st>>>(part, ldpart, nparts, n, s02, s, nrm);
}

// =====================================================================================================
// a5-a7: moments / rate / adjoint weights per image (reference utils.py:1090-1101, 1138, 1243, 1266)
// =====================================================================================================
__global__ void moments_rows_kernel(const double* __restrict__ s, const double* __restrict__ part1,
                                    const double* __restrict__ part2, long ldpart, int nparts, int nparts2,
                                    const double* __restrict__ r, long n, double A, double lambda0,
                                    double* __restrict__ lam_m, double* __restrict__ lam_var, double* __restrict__ f,
                                    double* __restrict__ u, double* __restrict__ v, double* __restrict__ blockpart) {
  __shared__ double red[32];
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  double rl = 0.0, fs = 0.0, rs = 0.0, nn = 0.0;
  if (i < n) {
    double q1 = 0.0, q2 = 0.0;
    for (int t = 0; t < nparts; ++t) q1 += part1[t * ldpart + i];
    for (int t = 0; t < nparts2; ++t) q2 += part2[t * ldpart + i];
    const double lv = s[i] + 0.5 * q1;
    const double lm = q2;
    const double fi = exp(A * lm + 0.5 * A * A * lv + lambda0);   // U:1138
    const double ri = r[i];
    lam_m[i] = lm;
    lam_var[i] = lv;
    f[i] = fi;
    u[i] = A * (ri - fi);          // d loglik / d lam_m   (U:1266)
    v[i] = -0.5 * A * A * fi;      // d loglik / d lam_var
    rl = ri * lm; fs = fi; rs = ri; nn = (fi != fi) ? 1.0 : 0.0;
  }
  rl = block_sum(rl, red);
  fs = block_sum(fs, red);
  rs = block_sum(rs, red);
  nn = block_sum(nn, red);
  if (threadIdx.x == 0) {
    blockpart[blockIdx.x * 4 + 0] = rl;
    blockpart[blockIdx.x * 4 + 1] = fs;
    blockpart[blockIdx.x * 4 + 2] = rs;
    blockpart[blockIdx.x * 4 + 3] = nn;
  }
}

int moments_rows_blocks(long n) { return cdiv(n, 256); }

void moments_rows(const double* s, const double* part1, const double* part2, long ldpart, int nparts, int nparts2,
                  const double* r, long n, double A, double lambda0, double* lam_m, double* lam_var, double* f,
                  double* u, double* v, double* blockpart, cudaStream_t st) {
  if (n <= 0) return;
  note_launch(); moments_rows_kernel<<<moments_rows_blocks(n), 256, 0, st>>>(s, part1, part2, ldpart, nparts, nparts2, r, n, A, lambda0,
                                                              lam_m, lam_var, f, u, v, blockpart);
}

// =====================================================================================================
// arc-cosine map as its own streaming pass (utils.py:982-990): K <- X1X2 * J(cos), cos stored beside it.
// Run at full occupancy so the long fp64 acos/sqrt/div dependency chains overlap across many warps
// (as a GEMM epilogue they ran on 8 warps per SM with nothing to hide their latency).
// =====================================================================================================
__global__ void __launch_bounds__(256) acos_forward_kernel(double* __restrict__ K, double* __restrict__ cosm, long ld,
                                                           long n, int m, const double* __restrict__ na,
                                                           const double* __restrict__ nb, double s02) {
  // Two rows per block iteration, one column pair per thread (no 64-bit index division per element); the NEXT two rows
  // are loaded before the current ones are worked on: ncu showed 41 % of the warp samples waiting on the single 16-byte
  // load each thread had in flight in front of ~280 dependent instructions.
  const int m2 = (m + 1) >> 1;
  const long step = 2L * gridDim.x;
  for (int j2 = threadIdx.x; j2 < m2; j2 += blockDim.x) {
    const int j = 2 * j2;
    const bool two = j + 1 < m;
    const double nb0 = __ldg(nb + j), nb1 = two ? __ldg(nb + j + 1) : 1.0;
    long i = 2L * blockIdx.x;
    double2 cur[2];
#pragma unroll
    for (int q = 0; q < 2; ++q)
      cur[q] = (i + q < n) ? *reinterpret_cast<const double2*>(K + (i + q) * ld + j) : make_double2(0.0, 0.0);
    for (; i < n; i += step) {
      double2 nxt[2];
#pragma unroll
      for (int q = 0; q < 2; ++q)
        nxt[q] = (i + step + q < n) ? *reinterpret_cast<const double2*>(K + (i + step + q) * ld + j)
                                    : make_double2(0.0, 0.0);
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        if (i + q < n) {
          const double nai = __ldg(na + i + q);
          double c0, c1 = 0.0;
          const double k0 = acos_kernel_value(cur[q].x + s02, nai * nb0, &c0);
          const double k1 = two ? acos_kernel_value(cur[q].y + s02, nai * nb1, &c1) : 0.0;
          const long o = (i + q) * ld + j;                              // ld is even, padding columns are in bounds
          *reinterpret_cast<double2*>(K + o) = make_double2(k0, k1);
          *reinterpret_cast<double2*>(cosm + o) = make_double2(c0, c1);
        }
      }
      cur[0] = nxt[0]; cur[1] = nxt[1];
    }
  }
}
void acos_forward(double* K, double* cosm, long ld, long n, int m, const double* na, const double* nb, double s02,
                  cudaStream_t st) {
  if (n <= 0 || m <= 0) return;
  const int m2 = (m + 1) >> 1;
  const int threads = m2 >= 192 ? 256 : (m2 >= 96 ? 128 : 64);
  const long nb_ = std::min<long>((n + 1) / 2, (long)num_sms() * (2048 / threads));
  note_launch(); acos_forward_kernel<<<(unsigned)nb_, threads, 0, st>>>(K, cosm, ld, n, m, na, nb, s02);
}

// =====================================================================================================
// adjoint of the arc-cosine nonlinearity (reference local partials utils.py:996-1021)
// =====================================================================================================
// CM = column slots per thread (m <= 256*CM); rows are processed RU at a time with all their loads issued first,
// so enough memory requests are in flight per SM to approach the HBM rate (the pass reads T, cos, K and writes R).
template <int MODE, int CM>
__global__ void __launch_bounds__(256, (CM <= 4) ? 2 : 1) acos_backward_kernel(double* __restrict__ G, const double* __restrict__ cosm,
                                                            const double* __restrict__ Kmat, long ld, long n, int m,
                                                            const double* __restrict__ u, const double* __restrict__ v,
                                                            const double* __restrict__ bb,
                                                            const double* __restrict__ na,
                                                            const double* __restrict__ nb, double* __restrict__ rowout,
                                                            double* __restrict__ colpart_k,
                                                            double* __restrict__ colpart_y, long ldcol,
                                                            double* __restrict__ sumpart,
                                                            double* __restrict__ colmax_r,
                                                            double* __restrict__ colmax_w) {
  // RU rows per batch; the loads of batch b+1 are issued before batch b is worked on (register double buffer), so
  // every warp keeps RU*CM*3 loads in flight all the time: the pass reads T, cos, K and writes R, and with the loads
  // of a batch only issued after the previous batch's arithmetic, ncu showed 48 % of the samples on long-scoreboard.
  constexpr int RU = (CM <= 2) ? 2 : 1;
  __shared__ double rowred[8][kBwdRows];
  __shared__ double red[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // persistent blocks: each walks row chunks blockIdx.x, blockIdx.x + gridDim.x, ... and keeps its column sums / maxima
  // in registers across them, so the per-block column partials that the follow-up reductions read are <= 296 rows
  // instead of one per 64 images (those reductions were ~20 us each, eight per evaluation)
  const long nchunks = (n + kBwdRows - 1) / kBwdRows;
  long r0 = (long)blockIdx.x * kBwdRows;
  double ck[CM], cy[CM], bbj[CM], nbj[CM], mr[CM], mw[CM];
#pragma unroll
  for (int c = 0; c < CM; ++c) {
    ck[c] = cy[c] = mr[c] = mw[c] = 0.0;
    const int j = tid + c * 256;
    bbj[c] = (MODE == 0 && j < m) ? bb[j] : 0.0;
    nbj[c] = (j < m) ? nb[j] : 0.0;
  }
  double sumR = 0.0;
  double gv[RU][CM], cv[RU][CM], kv[RU][CM], ui[RU], vi[RU], nai[RU];
  auto load_batch = [&](int rr0, double (&g_)[RU][CM], double (&c_)[RU][CM], double (&k_)[RU][CM], double (&u_)[RU],
                        double (&v_)[RU], double (&n_)[RU]) {
#pragma unroll
    for (int q = 0; q < RU; ++q) {
      const long i = r0 + rr0 + q;
      const bool rv = rr0 + q < kBwdRows && i < n;
      u_[q] = (MODE == 0 && rv) ? u[i] : 0.0;
      v_[q] = (MODE == 0 && rv) ? v[i] : 0.0;
      n_[q] = rv ? na[i] : 1.0;
#pragma unroll
      for (int c = 0; c < CM; ++c) {
        const int j = tid + c * 256;
        const bool ok = rv && j < m;
        const long o = i * ld + j;
        g_[q][c] = ok ? G[o] : 0.0;
        c_[q][c] = ok ? cosm[o] : 0.0;
        k_[q][c] = (MODE == 0 && ok) ? Kmat[o] : 0.0;
      }
    }
  };
  for (long chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x) {
  r0 = chunk * kBwdRows;
  load_batch(0, gv, cv, kv, ui, vi, nai);
  for (int rr0 = 0; rr0 < kBwdRows; rr0 += RU) {
    if (r0 + rr0 >= n) break;
    double gn[RU][CM], cn[RU][CM], kn[RU][CM], un[RU], vn[RU], nn[RU];
    load_batch(rr0 + RU, gn, cn, kn, un, vn, nn);
#pragma unroll
    for (int q = 0; q < RU; ++q) {
      const long i = r0 + rr0 + q;
      double rho = 0.0;
#pragma unroll
      for (int c = 0; c < CM; ++c) {
        const int j = tid + c * 256;
        const double gk = (MODE == 0) ? (ui[q] * bbj[c] + vi[q] * gv[q][c]) : gv[q][c];
        double P, Q;
        acos_partials(cv[q][c], P, Q);
        const double Rv = gk * P;
        const double rq = gk * Q;
        if (i < n && j < m) G[i * ld + j] = Rv;
        sumR += Rv;
        rho += rq * nbj[c];
        ck[c] += rq * nai[q];
        if (MODE == 0) {
          cy[c] += kv[q][c] * ui[q];
          mr[c] = fmax(mr[c], fabs(Rv));
          mw[c] = fmax(mw[c], -vi[q] * kv[q][c] * kv[q][c]);
        }
      }
      rho = warp_sum(rho);
      if (lane == 0) rowred[warp][rr0 + q] = rho;
    }
#pragma unroll
    for (int q = 0; q < RU; ++q) {
      ui[q] = un[q]; vi[q] = vn[q]; nai[q] = nn[q];
#pragma unroll
      for (int c = 0; c < CM; ++c) { gv[q][c] = gn[q][c]; cv[q][c] = cn[q][c]; kv[q][c] = kn[q][c]; }
    }
  }
  __syncthreads();
  if (tid < kBwdRows) {
    const long i = r0 + tid;
    if (i < n) {
      double rho = 0.0;
#pragma unroll
      for (int w = 0; w < 8; ++w) rho += rowred[w][tid];
      rowout[i] = ((MODE == 0) ? v[i] : 0.0) + 0.5 * rho / na[i];
    }
  }
  __syncthreads();    // rowred is reused by the next chunk
  }
#pragma unroll
  for (int c = 0; c < CM; ++c) {
    const int j = tid + c * 256;
    if (j < m) {
      colpart_k[(long)blockIdx.x * ldcol + j] = ck[c];
      if (MODE == 0) colpart_y[(long)blockIdx.x * ldcol + j] = cy[c];
      if (MODE == 0 && colmax_r) {
        colmax_r[(long)blockIdx.x * ldcol + j] = mr[c];
        colmax_w[(long)blockIdx.x * ldcol + j] = mw[c];
      }
    }
  }
  sumR = block_sum(sumR, red);
  if (tid == 0) sumpart[blockIdx.x] = sumR;
}

int acos_backward_blocks(long n) { const long c = cdiv(n, kBwdRows); return (int)(c < 296 ? c : 296); }

template <int MODE>
static void acos_backward_launch(int nblk, double* G, const double* cosm, const double* Kmat, long ld, long n, int m,
                                 const double* u, const double* v, const double* bb, const double* na,
                                 const double* nb, double* rowout, double* colpart_k, double* colpart_y, long ldcol,
                                 double* sumpart, double* colmax_r, double* colmax_w, cudaStream_t st) {
#define SGP_BWD(CM_)                                                                                              \
  acos_backward_kernel<MODE, CM_><<<nblk, 256, 0, st>>>(G, cosm, Kmat, ld, n, m, u, v, bb, na, nb, rowout,        \
                                                        colpart_k, colpart_y, ldcol, sumpart, colmax_r, colmax_w)
  note_launch();
  if (m <= 256) SGP_BWD(1);
  else if (m <= 512) SGP_BWD(2);
  else if (m <= 1024) SGP_BWD(4);
  else SGP_BWD(8);
#undef SGP_BWD
}

void acos_backward(int mode, double* G, const double* cosm, const double* Kmat, long ld, long n, int m,
                   const double* u, const double* v, const double* bb, const double* na, const double* nb,
                   double* rowout, double* colpart_k, double* colpart_y, long ldcol, double* sumpart,
                   cudaStream_t st, double* colmax_r, double* colmax_w) {
  if (n <= 0) return;
  const int nblk = acos_backward_blocks(n);
  if (mode == 0)
    acos_backward_launch<0>(nblk, G, cosm, Kmat, ld, n, m, u, v, bb, na, nb, rowout, colpart_k, colpart_y, ldcol,
                            sumpart, colmax_r, colmax_r ? colmax_w : nullptr, st);
  else
    acos_backward_launch<1>(nblk, G, cosm, Kmat, ld, n, m, u, v, bb, na, nb, rowout, colpart_k, colpart_y, ldcol,
                            sumpart, nullptr, nullptr, st);
}

// =====================================================================================================
// deterministic reductions
// =====================================================================================================
// block = 32 columns x 8 row groups; each thread sums rows rg, rg+8, ... then the 8 groups are added in fixed order
__global__ void __launch_bounds__(256) col_reduce_kernel(const double* __restrict__ part, int nblk, int m, long ld,
                                                         const double* __restrict__ w, double* __restrict__ out,
                                                         int accumulate) {
  __shared__ double red[8][33];
  const int lane = threadIdx.x & 31, rg = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + lane;
  double a = 0.0;
  if (j < m) {
    if (w) for (int b = rg; b < nblk; b += 8) a += part[(long)b * ld + j] * w[b];
    else   for (int b = rg; b < nblk; b += 8) a += part[(long)b * ld + j];
  }
  red[rg][lane] = a;
  __syncthreads();
  if (rg == 0 && j < m) {
    double t = 0.0;
#pragma unroll
    for (int g = 0; g < 8; ++g) t += red[g][lane];
    out[j] = accumulate ? out[j] + t : t;
  }
}
void col_reduce(const double* part, int nblk, int m, long ld, double* out, int accumulate, cudaStream_t st) {
  if (m <= 0) return;
  note_launch(); col_reduce_kernel<<<cdiv(m, 32), 256, 0, st>>>(part, nblk, m, ld, nullptr, out, accumulate);
}

__global__ void slice_reduce_kernel(const double* __restrict__ part, int nz, long zstride, long n,
                                    double* __restrict__ out, int accumulate) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    double a = 0.0;
    for (int z = 0; z < nz; ++z) a += part[z * zstride + i];
    out[i] = accumulate ? out[i] + a : a;
  }
}
void slice_reduce(const double* part, int nz, long zstride, long n, double* out, int accumulate, cudaStream_t st) {
  if (n <= 0) return;
  int nb = cdiv(n, 256);
  if (nb > num_sms() * 8) nb = num_sms() * 8;
  note_launch(); slice_reduce_kernel<<<nb, 256, 0, st>>>(part, nz, zstride, n, out, accumulate);
}

__global__ void sum_to_slot_kernel(const double* __restrict__ x, long n, long stride, double scale,
                                   double* __restrict__ scal, int slot, int accumulate) {
  __shared__ double red[32];
  double a = 0.0;
  for (long i = threadIdx.x; i < n; i += blockDim.x) a += x[i * stride];
  a = block_sum(a, red);
  if (threadIdx.x == 0) scal[slot] = (accumulate ? scal[slot] : 0.0) + scale * a;
}
void sum_to_slot(const double* x, long n, long stride, double scale, double* scal, int slot, int accumulate,
                 cudaStream_t st) {
  note_launch(); sum_to_slot_kernel<<<1, 1024, 0, st>>>(x, n < 0 ? 0 : n, stride, scale, scal, slot, accumulate);
}
// nslots interleaved columns at once: scal[slot0 + k] (+)= sum_i x[i*stride + k]; block k sums column k in the
// same fixed order as sum_to_slot (identical results), one launch instead of nslots
__global__ void sum_to_slots_kernel(const double* __restrict__ x, long n, long stride, double* __restrict__ scal,
                                    int slot0, int accumulate) {
  __shared__ double red[32];
  const int k = blockIdx.x;
  double a = 0.0;
  for (long i = threadIdx.x; i < n; i += blockDim.x) a += x[i * stride + k];
  a = block_sum(a, red);
  if (threadIdx.x == 0) scal[slot0 + k] = (accumulate ? scal[slot0 + k] : 0.0) + a;
}
void sum_to_slots(const double* x, long n, long stride, int nslots, double* scal, int slot0, int accumulate,
                  cudaStream_t st) {
  note_launch(); sum_to_slots_kernel<<<nslots, 1024, 0, st>>>(x, n < 0 ? 0 : n, stride, scal, slot0, accumulate);
}

// =====================================================================================================
// small dense helpers
// =====================================================================================================
__global__ void symmetrize_kernel(const double* __restrict__ A, long lda, int n, double* __restrict__ out, long ldo) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (i < n && j < n) out[(long)i * ldo + j] = (A[(long)i * lda + j] + A[(long)j * lda + i]) * 0.5;
}
void symmetrize(const double* A, long lda, int n, double* out, long ldo, cudaStream_t st) {
  if (n <= 0) return;
  note_launch(); symmetrize_kernel<<<dim3(cdiv(n, 128), n), 128, 0, st>>>(A, lda, n, out, ldo);
}

// A (rows x cols) += B^T, B is cols x rows: 32 x 32 tiles through shared memory (both sides coalesced)
__global__ void __launch_bounds__(256) add_transposed_kernel(double* __restrict__ A, long lda, int rows, int cols,
                                                             const double* __restrict__ B, long ldb) {
  __shared__ double tile[32][33];
  const int i0 = blockIdx.y * 32, j0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int r = ty; r < 32; r += 8)                      // B[j0 + r][i0 + tx]
    tile[r][tx] = (j0 + r < cols && i0 + tx < rows) ? B[(long)(j0 + r) * ldb + i0 + tx] : 0.0;
  __syncthreads();
  for (int r = ty; r < 32; r += 8)
    if (i0 + r < rows && j0 + tx < cols) A[(long)(i0 + r) * lda + j0 + tx] += tile[tx][r];
}
void add_transposed(double* A, long lda, int rows, int cols, const double* B, long ldb, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return;
  note_launch(); add_transposed_kernel<<<dim3(cdiv(cols, 32), cdiv(rows, 32)), 256, 0, st>>>(A, lda, rows, cols, B, ldb);
}

// A[i][j] = A[j][i] for j > i (completes a product whose upper CTA tiles were skipped)
__global__ void mirror_lower_kernel(double* __restrict__ A, long ld, int n) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (i < n && j < n && j > i) A[(long)i * ld + j] = A[(long)j * ld + i];
}
void mirror_lower(double* A, long ld, int n, cudaStream_t st) {
  if (n <= 0) return;
  note_launch(); mirror_lower_kernel<<<dim3(cdiv(n, 128), n), 128, 0, st>>>(A, ld, n);
}

__global__ void lincomb_kernel(double* __restrict__ out, long ldo, int rows, int cols, double a,
                               const double* __restrict__ A, long lda, double b, const double* __restrict__ B, long ldb,
                               double diag) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (i >= rows || j >= cols) return;
  double v = a * A[(long)i * lda + j];
  if (B) v += b * B[(long)i * ldb + j];
  if (i == j) v += diag;
  out[(long)i * ldo + j] = v;
}
void lincomb(double* out, long ldo, int rows, int cols, double a, const double* A, long lda, double b,
             const double* B, long ldb, double diag, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return;
  note_launch(); lincomb_kernel<<<dim3(cdiv(cols, 128), rows), 128, 0, st>>>(out, ldo, rows, cols, a, A, lda, b, B, ldb, diag);
}

__global__ void rank1_add_kernel(double* __restrict__ out, long ldo, int rows, int cols, double alpha,
                                 const double* __restrict__ x, const double* __restrict__ y) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (i < rows && j < cols) out[(long)i * ldo + j] += alpha * x[i] * y[j];
}
void rank1_add(double* out, long ldo, int rows, int cols, double alpha, const double* x, const double* y,
               cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return;
  note_launch(); rank1_add_kernel<<<dim3(cdiv(cols, 128), rows), 128, 0, st>>>(out, ldo, rows, cols, alpha, x, y);
}

// y = A x (trans = 0, one warp per row) or y = A^T x (trans = 1, one thread per column)
__global__ void gemv_n_kernel(const double* __restrict__ A, long lda, int rows, int cols,
                              const double* __restrict__ x, double* __restrict__ y) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  double a = 0.0;
  for (int j = lane; j < cols; j += 32) a += A[(long)row * lda + j] * x[j];
  a = warp_sum(a);
  if (lane == 0) y[row] = a;
}
void gemv(const double* A, long lda, int rows, int cols, int trans, const double* x, double* y, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return;
  if (!trans) {
    note_launch(); gemv_n_kernel<<<cdiv(rows, 8), 256, 0, st>>>(A, lda, rows, cols, x, y);
  } else {
    note_launch(); col_reduce_kernel<<<cdiv(cols, 32), 256, 0, st>>>(A, rows, cols, lda, x, y, 0);
  }
}

__global__ void copy_mat_kernel(const double* __restrict__ A, long lda, int rows, int cols, double* __restrict__ out,
                                long ldo) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= ldo) return;
  for (int i = blockIdx.y; i < rows; i += gridDim.y) out[(long)i * ldo + j] = (j < cols) ? A[(long)i * lda + j] : 0.0;
}
void copy_mat(const double* A, long lda, int rows, int cols, double* out, long ldo, cudaStream_t st) {
  if (rows <= 0) return;
  note_launch(); copy_mat_kernel<<<dim3(cdiv(ldo, 128), rows < 32768 ? rows : 32768), 128, 0, st>>>(A, lda, rows, cols, out, ldo);
}

__global__ void gkt_assemble_kernel(double* __restrict__ G, long ld, int q, const double* __restrict__ Kinv,
                                    const double* __restrict__ KVK, const double* __restrict__ b, double klw) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (i < q && j < q)
    G[(long)i * ld + j] += klw * (-0.5 * Kinv[(long)i * ld + j] + 0.5 * KVK[(long)j * ld + i] + 0.5 * b[i] * b[j]);
}
void gkt_assemble(double* G, long ld, int q, const double* Kinv, const double* KVK, const double* b, double klw,
                  cudaStream_t st) {
  note_launch(); gkt_assemble_kernel<<<dim3(cdiv(q, 128), q), 128, 0, st>>>(G, ld, q, Kinv, KVK, b, klw);
}

__global__ void col_weights_kernel(const double* __restrict__ ka, const double* __restrict__ kb,
                                   const double* __restrict__ extra, const double* __restrict__ nrm, int m,
                                   double* __restrict__ out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  double k = ka ? ka[j] : 0.0;
  if (kb) k += kb[j];
  out[j] = 0.5 * k / nrm[j] + (extra ? extra[j] : 0.0);
}
void col_weights(const double* ka, const double* kb, const double* extra, const double* nrm, int m, double* out,
                 cudaStream_t st) {
  if (m <= 0) return;
  note_launch(); col_weights_kernel<<<cdiv(m, 128), 128, 0, st>>>(ka, kb, extra, nrm, m, out);
}

// =====================================================================================================
// blocked Cholesky (panel 64): [factor 64x64 diagonal block + TRSM of the rows below] ; DMMA trailing update
// =====================================================================================================
constexpr int kNB = 64;

// One CTA factors a 64 x 64 diagonal block D = L L^T and inverts the factor, W = L^-1 (chol_diag.cuh: the block is itself
// blocked 8 x 8, the pivot chain runs inside one warp).  W turns the panel solve below the block and the triangular
// inverse into plain DMMA GEMMs.
constexpr int kCholThreads = 256;
constexpr int kCholPitch = 65;
constexpr size_t kCholSmem = (size_t)(2 * kNB * kCholPitch + 512 + 512) * sizeof(double);   // 74 752 bytes
__global__ void __launch_bounds__(kCholThreads) chol_diag_kernel(double* __restrict__ A, long ld, int nb,
                                                                 double* __restrict__ Wout, int* __restrict__ flags,
                                                                 int bit_notpd, int bit_small) {
  extern __shared__ double chol_sm[];
  double* D = chol_sm;
  double* Wm = D + kNB * kCholPitch;
  double* Ti = Wm + kNB * kCholPitch;
  double* pan = Ti + 512;
  const int bad = diag_factor_inv<kCholPitch>(A, ld, nb, D, Wm, Ti, pan, bit_notpd, bit_small);
  if (threadIdx.x == 0 && bad) atomicOr(flags, bad);
  for (int e = threadIdx.x; e < kNB * kNB; e += kCholThreads) {
    const int r = e >> 6, c = e & 63;
    if (r < nb && c < nb) A[(long)r * ld + c] = D[r * kCholPitch + c];
    Wout[e] = Wm[r * kCholPitch + c];
  }
}

// once per device, before the first factorisation: the diagonal-block kernel's 73 KB of shared memory are above the default
cudaError_t chol_init() {
  return cudaFuncSetAttribute(chol_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCholSmem);
}

// In-place blocked Cholesky; Winv (cdiv(n,64) blocks of 64 x 64) receives the inverses of the diagonal blocks of L.
cudaError_t cholesky_lower(double* A, long ld, int n, int* flags, int bit_notpd, int bit_small, double* Winv,
                           cudaStream_t st) {
  for (int j0 = 0; j0 < n; j0 += kNB) {
    const int nb = (n - j0) < kNB ? (n - j0) : kNB;
    const int rem = n - j0 - nb;
    double* W = Winv + (long)(j0 / kNB) * kNB * kNB;
    note_launch();
    chol_diag_kernel<<<1, kCholThreads, kCholSmem, st>>>(A + (long)j0 * ld + j0, ld, nb, W, flags, bit_notpd, bit_small);
    if (rem > 0) {
      const int j1 = j0 + nb;
      GemmArgs t{};   // panel below the block, in place: L21 = A21 W^T   (one n-tile: a CTA reads only the rows it writes)
      t.A = A + (long)j1 * ld + j0; t.lda = ld;
      t.B = W; t.ldb = kNB;
      t.C = A + (long)j1 * ld + j0; t.ldc = ld;
      t.M = rem; t.N = nb; t.K = nb; t.k_chunk = kNB; t.alpha = 1.0;
      cudaError_t e = gemm_f64_launch(t, GemmLaunch{0, 0, kStore, true, 1}, st);
      if (e != cudaSuccess) return e;
      GemmArgs g{};   // trailing update A22 -= L21 L21^T
      g.A = A + (long)j1 * ld + j0; g.lda = ld;
      g.B = g.A; g.ldb = ld;
      g.C = A + (long)j1 * ld + j1; g.ldc = ld;
      g.M = rem; g.N = rem; g.K = nb; g.k_chunk = kNB; g.alpha = -1.0; g.accumulate = 1;
      e = gemm_f64_launch(g, GemmLaunch{0, 0, kStore, true, 1}, st);
      if (e != cudaSuccess) return e;
    }
  }
  return cudaGetLastError();
}

// X (zeroed) <- the diagonal blocks of L^-1 from the Cholesky sweep
__global__ void place_diag_blocks_kernel(const double* __restrict__ Winv, double* __restrict__ X, long ld, int n) {
  const int j0 = blockIdx.x * kNB;
  for (int e = threadIdx.x; e < kNB * kNB; e += blockDim.x) {
    const int r = e >> 6, c = e & 63;
    if (j0 + r < n && j0 + c < n) X[(long)(j0 + r) * ld + j0 + c] = Winv[(long)blockIdx.x * kNB * kNB + e];
  }
}

__global__ void zero_mat_kernel(double* __restrict__ X, long ld, int rows) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (i < rows && j < ld) X[(long)i * ld + j] = 0.0;
}

cudaError_t tri_inverse_lower(const double* L, long ld, int n, double* X, double* tmp, const double* Winv,
                              cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  note_launch(); zero_mat_kernel<<<dim3(cdiv(ld, 128), n), 128, 0, st>>>(X, ld, n);
  const int nblk = cdiv(n, kNB);
  note_launch(); place_diag_blocks_kernel<<<nblk, 256, 0, st>>>(Winv, X, ld, n);
  for (int I = 1; I < nblk; ++I) {
    const int i0 = I * kNB;
    const int nb = (n - i0) < kNB ? (n - i0) : kNB;
    // tmp[nb x i0] = L[i0:i0+nb, 0:i0] * X[0:i0, 0:i0]
    GemmArgs g{};
    g.A = L + (long)i0 * ld; g.lda = ld;
    g.B = X; g.ldb = ld;
    g.C = tmp; g.ldc = ld;
    g.M = nb; g.N = i0; g.K = i0; g.k_chunk = (i0 + 15) / 16 * 16; g.alpha = 1.0;
    cudaError_t e = gemm_f64_launch(g, GemmLaunch{0, 1, kStore, true, 1}, st);
    if (e != cudaSuccess) return e;
    // X[i0:i0+nb, 0:i0] = -Dinv_I * tmp
    GemmArgs h{};
    h.A = X + (long)i0 * ld + i0; h.lda = ld;
    h.B = tmp; h.ldb = ld;
    h.C = X + (long)i0 * ld; h.ldc = ld;
    h.M = nb; h.N = i0; h.K = nb; h.k_chunk = kNB; h.alpha = -1.0;
    e = gemm_f64_launch(h, GemmLaunch{0, 1, kStore, true, 1}, st);
    if (e != cudaSuccess) return e;
  }
  return cudaGetLastError();
}

__global__ void logdet_kernel(const double* __restrict__ L, long ld, int n, double* __restrict__ scal, int slot) {
  __shared__ double red[32];
  double a = 0.0;
  for (int k = threadIdx.x; k < n; k += blockDim.x) a += log(L[(long)k * ld + k]);
  a = block_sum(a, red);
  if (threadIdx.x == 0) scal[slot] = 2.0 * a;
}
void logdet_from_chol(const double* L, long ld, int n, double* scal, int slot, cudaStream_t st) {
  note_launch(); logdet_kernel<<<1, 256, 0, st>>>(L, ld, n, scal, slot);
}

// one block per row: part[i] = sum_j V_ij Kinv_ij (Kinv is exactly symmetric: symmetrised by the caller), part[q+i] = m_i b_i
__global__ void __launch_bounds__(128) kl_rows_kernel(const double* __restrict__ m_b, const double* __restrict__ b,
                                                      const double* __restrict__ V, long ldv,
                                                      const double* __restrict__ Kinv, long ldk, int q,
                                                      double* __restrict__ part) {
  __shared__ double red[32];
  const int i = blockIdx.x;
  double tr = 0.0;
  for (int j = threadIdx.x; j < q; j += blockDim.x) tr += V[(long)i * ldv + j] * Kinv[(long)i * ldk + j];
  tr = block_sum(tr, red);
  if (threadIdx.x == 0) {
    part[i] = tr;
    part[q + i] = m_b[i] * b[i];
  }
}
void kl_scalars(const double* m_b, const double* b, const double* V, long ldv, const double* Kinv, long ldk, int q,
                double* scal, int slot_mb, int slot_tr, double* scratch, cudaStream_t st) {
  note_launch(); kl_rows_kernel<<<q, 128, 0, st>>>(m_b, b, V, ldv, Kinv, ldk, q, scratch);
  sum_to_slot(scratch, q, 1, 1.0, scal, slot_tr, 0, st);
  sum_to_slot(scratch + q, q, 1, 1.0, scal, slot_mb, 0, st);
}

// =====================================================================================================
// <dC_k, GC> for the 5 localker hyper-parameters; dC_k is regenerated from C and the pixel coordinates
// (reference utils.py:902-909), it is never stored.
// =====================================================================================================
__global__ void __launch_bounds__(256) contract_dC_kernel(const double* __restrict__ GCpart, int nz, long zstride,
                                                          const double* __restrict__ GCextra,
                                                          const double* __restrict__ C, long ld, int p,
                                                          const double* __restrict__ gx,
                                                          const double* __restrict__ gy, ThetaDev th,
                                                          double* __restrict__ gpart) {
  __shared__ double red[32];
  const int i = blockIdx.x;
  double acc[5] = {0, 0, 0, 0, 0};
  if (i < p) {
    const double xi = gx[i], yi = gy[i];
    const double lai = -th.eb * ((xi - th.ex) * (xi - th.ex) + (yi - th.ey) * (yi - th.ey));
    for (int j = threadIdx.x; j < p; j += blockDim.x) {
      const long o = (long)i * ld + j;
      double gsum = GCextra ? GCextra[o] : 0.0;
      for (int z = 0; z < nz; ++z) gsum += GCpart[z * zstride + o];
      const double xj = gx[j], yj = gy[j];
      const double laj = -th.eb * ((xj - th.ex) * (xj - th.ex) + (yj - th.ey) * (yj - th.ey));
      const double ls = -th.er * ((xj - xi) * (xj - xi) + (yj - yi) * (yj - yi));
      const double gc = gsum * C[o];
      acc[0] += gc;
      acc[1] += gc * (lai + laj);
      acc[2] += gc * ls;
      acc[3] += gc * (xi + xj - 2 * th.ex);
      acc[4] += gc * (yi + yj - 2 * th.ey);
    }
  }
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const double sres = block_sum(acc[k], red);
    if (threadIdx.x == 0) gpart[(long)blockIdx.x * 5 + k] = sres;
  }
}
int contract_blocks(int p) { return p; }
void contract_dC(const double* GCpart, int nz, long zstride, const double* GCextra, const double* C, long ld, int p,
                 const double* gx, const double* gy, const ThetaDev& th, double* gpart, cudaStream_t st) {
  if (p <= 0) return;
  note_launch(); contract_dC_kernel<<<p, 256, 0, st>>>(GCpart, nz, zstride, GCextra, C, ld, p, gx, gy, th, gpart);
}

__device__ __forceinline__ void assemble_result(const double* __restrict__ scal, const int* __restrict__ flags, double A,
                                                double lambda0, const ThetaDev& th, int p, double logdetV_override,
                                                int use_override, double out15, double* v) {
  const double s0 = th.s0, amp = th.amp, eb = th.eb;
  const double logdetV = use_override ? logdetV_override : scal[kSlotLogdetV];
  const double loglik = A * scal[kSlotRLam] + lambda0 * scal[kSlotSumR] - scal[kSlotSumF];      // U:1243
  const double KL = -0.5 * logdetV + 0.5 * scal[kSlotLogdetK] + 0.5 * scal[kSlotMb] + 0.5 * scal[kSlotTrVK];  // U:1326
  int fl = flags[0];
  if (scal[kSlotNaN] > 0.0) fl |= kFlagNaN;
  v[0] = loglik;
  v[1] = KL;
  v[2] = 2.0 * s0 * scal[kSlotSig];          // sigma_0
  v[3] = 2.0 * eb * scal[kSlotG + 3];        // eps_0x
  v[4] = 2.0 * eb * scal[kSlotG + 4];        // eps_0y
  v[5] = scal[kSlotG + 1];                   // -2log2beta
  v[6] = scal[kSlotG + 2];                   // -log2rho2
  v[7] = scal[kSlotG + 0] / amp;             // Amp
  v[8] = (double)fl;
  v[9] = (double)p;
  v[10] = scal[kSlotSumR];
  v[11] = scal[kSlotSumF];
  v[12] = scal[kSlotRLam];
  v[13] = scal[kSlotLogdetK];
  v[14] = logdetV;
  v[15] = out15;
}

__global__ void finalize_kernel(const double* __restrict__ scal, const int* __restrict__ flags, double A,
                                double lambda0, ThetaDev th, int p, double logdetV_override, int use_override,
                                double* __restrict__ out, double out15) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double v[16];
  assemble_result(scal, flags, A, lambda0, th, p, logdetV_override, use_override, out15, v);
#pragma unroll
  for (int k = 0; k < 16; ++k) out[k] = v[k];
}

void finalize(const double* scal, const int* flags, double A, double lambda0, const ThetaDev& th, int p,
              double logdetV_override, int use_override, double* out, double out15, cudaStream_t st) {
  note_launch(); finalize_kernel<<<1, 32, 0, st>>>(scal, flags, A, lambda0, th, p, logdetV_override, use_override, out, out15);
}

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// One warp.  Lane k < 16 owns result element k; lane i < world polls rank i's flag.
__global__ void finalize_exchange_kernel(const double* __restrict__ scal, const int* __restrict__ flags, double A,
                                         double lambda0, ThetaDev th, int p, double logdetV_override, int use_override,
                                         double* __restrict__ out, double out15, CommDev comm) {
  __shared__ double v[16];
  __shared__ int timed_out;
  const int lane = threadIdx.x;
  if (lane == 0) {
    double t[16];
    assemble_result(scal, flags, A, lambda0, th, p, logdetV_override, use_override, out15, t);
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = t[k];
    timed_out = 0;
  }
  __syncwarp();
  const int par = (int)(comm.epoch & 1ULL), W = comm.world;
  // 1. my 16 doubles into slot [par][my rank] of every rank's buffer (my own included)
  if (lane < 16)
    for (int j = 0; j < W; ++j) comm.slots[j][((size_t)par * W + comm.rank) * 16 + lane] = v[lane];
  __threadfence_system();
  __syncwarp();
  // 2. publish: flag [par][my rank] of every rank = epoch (release: the slot stores above are visible before it)
  if (lane < W) st_release_sys(comm.flags[lane] + (size_t)par * W + comm.rank, comm.epoch);
  // 3. wait for every rank's flag in MY buffer
  if (lane < W) {
    const unsigned long long* f = comm.flags[comm.rank] + (size_t)par * W + lane;
    const unsigned long long t0 = global_ns();
    while (ld_acquire_sys(f) != comm.epoch) {
      if (global_ns() - t0 > 3000000000ULL) { atomicExch(&timed_out, 1); break; }   // 3 s: a peer never arrived
      __nanosleep(200);
    }
  }
  __threadfence_system();
  __syncwarp();
  // 4. reduce in rank order (identical bits on every rank): sums for loglik / gradients / the image sums, OR for flags
  if (lane < 16) {
    const volatile double* mine = comm.slots[comm.rank] + (size_t)par * W * 16;   // written by the peers: no cached copy
    double r = v[lane];
    if (lane == 0 || (lane >= 2 && lane <= 7) || (lane >= 10 && lane <= 12)) {
      r = 0.0;
      for (int i = 0; i < W; ++i) r += mine[i * 16 + lane];
    } else if (lane == 8) {
      int fl = timed_out ? kFlagCommTimeout : 0;
      for (int i = 0; i < W; ++i) fl |= (int)mine[i * 16 + 8];
      r = (double)fl;
    }
    out[lane] = r;
  }
}

void finalize_exchange(const double* scal, const int* flags, double A, double lambda0, const ThetaDev& th, int p,
                       double logdetV_override, int use_override, double* out, double out15, const CommDev& comm,
                       cudaStream_t st) {
  note_launch();
  finalize_exchange_kernel<<<1, 32, 0, st>>>(scal, flags, A, lambda0, th, p, logdetV_override, use_override, out, out15,
                                             comm);
}


// =====================================================================================================
// E-step / f-param helpers
// =====================================================================================================
__global__ void moments_combine_kernel(const double* __restrict__ Kvec, const double* __restrict__ part1,
                                       const double* __restrict__ part2, long ldpart, int nparts, int nparts2,
                                       const double* __restrict__ a, long lda, const double* __restrict__ K_b,
                                       long ldk, long n, int q, double* __restrict__ lam_m,
                                       double* __restrict__ lam_var) {
  // one warp per image
  const long i = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= n) return;
  double d = 0.0;
  for (int j = lane; j < q; j += 32) d += K_b[i * ldk + j] * a[i * lda + j];
  d = warp_sum(d);
  if (lane == 0) {
    double q1 = 0.0, q2 = 0.0;
    for (int t = 0; t < nparts; ++t) q1 += part1[t * ldpart + i];
    for (int t = 0; t < nparts2; ++t) q2 += part2[t * ldpart + i];
    lam_m[i] = q2;
    lam_var[i] = Kvec[i] + (q1 - d);
  }
}
void moments_combine(const double* Kvec, const double* part1, const double* part2, long ldpart, int nparts,
                     int nparts2, const double* a, long lda, const double* K_b, long ldk, long n, int q, double* lam_m,
                     double* lam_var, cudaStream_t st) {
  if (n <= 0) return;
  note_launch(); moments_combine_kernel<<<cdiv(n, 8), 256, 0, st>>>(Kvec, part1, part2, ldpart, nparts, nparts2, a, lda, K_b,
                                                                    ldk, n, q, lam_m, lam_var);
}

constexpr int kFparamBlocks = 148 * 4;

// Last-block-done finish of NS block-partial sums: every block stores its partials, takes a ticket, and the block that
// draws the last ticket adds all partials in a fixed order (deterministic) into out[0..NS) and re-arms the ticket.
template <int NS>
__device__ __forceinline__ void finish_block_sums(const double (&acc)[NS], double* red, double* __restrict__ blockpart,
                                                  unsigned* __restrict__ ticket, double* __restrict__ out) {
  __shared__ bool is_last;
#pragma unroll
  for (int k = 0; k < NS; ++k) {
    const double sres = block_sum(acc[k], red);
    if (threadIdx.x == 0) blockpart[(long)blockIdx.x * 8 + k] = sres;
  }
  if (threadIdx.x == 0) {
    __threadfence();
    is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
#pragma unroll
  for (int k = 0; k < NS; ++k) {
    double v = 0.0;
    for (unsigned i = threadIdx.x; i < gridDim.x; i += blockDim.x) v += __ldcg(blockpart + (long)i * 8 + k);
    const double sres = block_sum(v, red);
    if (threadIdx.x == 0) out[k] = sres;
  }
  if (threadIdx.x == 0) *ticket = 0u;
}

// f and the sums of the logA closure in one pass (utils.py:1138, 1224, 1243, 1253-1255).  el0 = exp(lambda0) when
// |lambda0| is moderate (el0_ok): then f = exp(x) * el0 shares the exponential with sum exp(x); otherwise, and for
// elements whose exp(x) leaves the normal range, f = exp(x + lambda0) exactly as the reference writes it.
__global__ void __launch_bounds__(256) fparam_kernel(const double* __restrict__ lam_m,
                                                     const double* __restrict__ lam_var,
                                                     const double* __restrict__ r, long n, double A, double lambda0,
                                                     double el0, int el0_ok, double* __restrict__ f,
                                                     double* __restrict__ blockpart, unsigned* __restrict__ ticket,
                                                     double* __restrict__ out) {
  __shared__ double red[32];
  double acc[6] = {0, 0, 0, 0, 0, 0};
  const double hA2 = 0.5 * A * A;
  auto body = [&](double lm, double lv, double ri, double& fout) {
    const double x = A * lm + hA2 * lv;
    const double e = exp(x);                             // U:1224
    double fi = e * el0;                                 // U:1138
    if (!el0_ok || !(e > 1.0e-290 && e < 1.0e290)) fi = exp(x + lambda0);
    fout = fi;
    acc[0] += ri * lm;
    acc[1] += fi;
    acc[2] += (lm + A * lv) * fi;                        // U:1253
    acc[3] += ri;
    acc[4] += e;
    acc[5] += (fi != fi) ? 1.0 : 0.0;
  };
  const long n2 = n >> 1;
  const double2* lm2 = reinterpret_cast<const double2*>(lam_m);
  const double2* lv2 = reinterpret_cast<const double2*>(lam_var);
  const double2* r2 = reinterpret_cast<const double2*>(r);
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(lam_m) | reinterpret_cast<uintptr_t>(lam_var) |
                        reinterpret_cast<uintptr_t>(r) | reinterpret_cast<uintptr_t>(f)) & 15) == 0;
  if (vec_ok) {
    const long stride = (long)gridDim.x * blockDim.x;
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + stride < n2; i += 2 * stride) {           // two independent double2 triples in flight per thread
      const double2 a0 = lm2[i], b0 = lv2[i], c0 = r2[i];
      const double2 a1 = lm2[i + stride], b1 = lv2[i + stride], c1 = r2[i + stride];
      double f0, f1, f2, f3;
      body(a0.x, b0.x, c0.x, f0);
      body(a0.y, b0.y, c0.y, f1);
      body(a1.x, b1.x, c1.x, f2);
      body(a1.y, b1.y, c1.y, f3);
      if (f) {
        reinterpret_cast<double2*>(f)[i] = make_double2(f0, f1);
        reinterpret_cast<double2*>(f)[i + stride] = make_double2(f2, f3);
      }
    }
    for (; i < n2; i += stride) {
      const double2 a = lm2[i], b = lv2[i], c = r2[i];
      double f0, f1;
      body(a.x, b.x, c.x, f0);
      body(a.y, b.y, c.y, f1);
      if (f) reinterpret_cast<double2*>(f)[i] = make_double2(f0, f1);
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
      double f0;
      body(lam_m[n - 1], lam_var[n - 1], r[n - 1], f0);
      if (f) f[n - 1] = f0;
    }
  } else {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
      double f0;
      body(lam_m[i], lam_var[i], r[i], f0);
      if (f) f[i] = f0;
    }
  }
  finish_block_sums<6>(acc, red, blockpart, ticket, out);
}
int fparam_blocks(long n) {
  const int nb = cdiv(n, 512);
  return nb < kFparamBlocks ? (nb < 1 ? 1 : nb) : kFparamBlocks;
}
void fparam_pass(const double* lam_m, const double* lam_var, const double* r, long n, double A, double lambda0,
                 double* f, double* blockpart, unsigned* ticket, double* out, cudaStream_t st) {
  const int ok = std::fabs(lambda0) < 300.0;
  note_launch();
  fparam_kernel<<<fparam_blocks(n), 256, 0, st>>>(lam_m, lam_var, r, n, A, lambda0, ok ? std::exp(lambda0) : 0.0, ok, f,
                                                  blockpart, ticket, out);
}


// sums for compute_loglikelihood with a caller-supplied f (utils.py:1240-1255):
// out[0..5) = sum r*lam_m, sum f, sum (lam_m + A lam_var) f, sum r, nan-count(f)
__global__ void __launch_bounds__(256) loglik_sums_kernel(const double* __restrict__ r, const double* __restrict__ f,
                                                          const double* __restrict__ lam_m,
                                                          const double* __restrict__ lam_var, long n, double A,
                                                          double* __restrict__ blockpart,
                                                          unsigned* __restrict__ ticket, double* __restrict__ out) {
  __shared__ double red[32];
  double acc[5] = {0, 0, 0, 0, 0};
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const double ri = r[i], fi = f[i], lm = lam_m[i];
    const double lv = lam_var ? lam_var[i] : 0.0;
    acc[0] += ri * lm;
    acc[1] += fi;
    acc[2] += (lm + A * lv) * fi;
    acc[3] += ri;
    acc[4] += (fi != fi) ? 1.0 : 0.0;
  }
  finish_block_sums<5>(acc, red, blockpart, ticket, out);
}
void loglik_sums_pass(const double* r, const double* f, const double* lam_m, const double* lam_var, long n, double A,
                      double* blockpart, unsigned* ticket, double* out, cudaStream_t st) {
  note_launch();
  loglik_sums_kernel<<<fparam_blocks(n), 256, 0, st>>>(r, f, lam_m, lam_var, n, A, blockpart, ticket, out);
}

constexpr int kColsumRows = 128;
__global__ void __launch_bounds__(256) colsum_weighted_kernel(const double* __restrict__ a, long lda, long n, int q,
                                                              const double* __restrict__ r,
                                                              const double* __restrict__ f,
                                                              double* __restrict__ colpart, long ldcol) {
  const long r0 = (long)blockIdx.x * kColsumRows;
  for (int j = threadIdx.x; j < q; j += blockDim.x) {
    double acc = 0.0;
    for (int rr = 0; rr < kColsumRows; ++rr) {
      const long i = r0 + rr;
      if (i < n) acc += a[i * lda + j] * (r[i] - f[i]);
    }
    colpart[(long)blockIdx.x * ldcol + j] = acc;
  }
}
int colsum_blocks(long n) { return cdiv(n, kColsumRows); }
void colsum_weighted(const double* a, long lda, long n, int q, const double* r, const double* f, double* colpart,
                     long ldcol, cudaStream_t st) {
  if (n <= 0) return;
  note_launch(); colsum_weighted_kernel<<<colsum_blocks(n), 256, 0, st>>>(a, lda, n, q, r, f, colpart, ldcol);
}

__global__ void scale_vec_kernel(double* __restrict__ x, long n, double alpha) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) x[i] *= alpha;
}
void scale_vec(double* x, long n, double alpha, cudaStream_t st) {
  if (n <= 0) return;
  note_launch(); scale_vec_kernel<<<cdiv(n, 256), 256, 0, st>>>(x, n, alpha);
}

// =====================================================================================================
// ntilde == nt closure (a = B, utils.py:2032 / 2068): the three elementwise pieces of its adjoint (m x q, small)
// =====================================================================================================
// Ga_ij = u_i m_b[j] + v_i (2 aV_ij - Kb_ij): adjoint of `da` in dlam_m = da m, dlam_var = ... + 2 rowdot(da, aV) - rowdot(K_b, da)
__global__ void same_ga_kernel(double* __restrict__ Ga, long ld, int rows, int cols, const double* __restrict__ u,
                               const double* __restrict__ v, const double* __restrict__ m_b,
                               const double* __restrict__ aV, const double* __restrict__ Kb) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (i >= rows || j >= ld) return;
  const long o = (long)i * ld + j;
  Ga[o] = j < cols ? u[i] * m_b[j] + v[i] * (2.0 * aV[o] - Kb[o]) : 0.0;
}
void same_ga(double* Ga, long ld, int rows, int cols, const double* u, const double* v, const double* m_b,
             const double* aV, const double* Kb, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return;
  note_launch(); same_ga_kernel<<<dim3(cdiv(ld, 128), rows), 128, 0, st>>>(Ga, ld, rows, cols, u, v, m_b, aV, Kb);
}
// E_ij = H_ij - v_i a_ij + BG_ij  (H = Ga K^-1: via da -> dK_b; -v a: the -rowdot(dK_b, a) term; BG: via dK_tilde_b)
__global__ void same_combine_kernel(double* __restrict__ E, long ld, int rows, int cols, const double* __restrict__ H,
                                    const double* __restrict__ v, const double* __restrict__ a,
                                    const double* __restrict__ BG) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (i >= rows || j >= ld) return;
  const long o = (long)i * ld + j;
  E[o] = j < cols ? H[o] - v[i] * a[o] + BG[o] : 0.0;
}
void same_combine(double* E, long ld, int rows, int cols, const double* H, const double* v, const double* a,
                  const double* BG, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return;
  note_launch(); same_combine_kernel<<<dim3(cdiv(ld, 128), rows), 128, 0, st>>>(E, ld, rows, cols, H, v, a, BG);
}
// out_ij = w_i in_ij (row scaling; padding columns up to ld copied scaled as well: they are zero)
__global__ void scale_rows_kernel(double* __restrict__ out, const double* __restrict__ in, long ld, int rows,
                                  const double* __restrict__ w) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (i < rows && j < ld) out[(long)i * ld + j] = w[i] * in[(long)i * ld + j];
}
void scale_rows(double* out, const double* in, long ld, int rows, const double* w, cudaStream_t st) {
  if (rows <= 0) return;
  note_launch(); scale_rows_kernel<<<dim3(cdiv(ld, 128), rows), 128, 0, st>>>(out, in, ld, rows, w);
}

}  // namespace sgp
