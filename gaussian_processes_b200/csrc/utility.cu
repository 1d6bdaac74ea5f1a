// Active-learning utility: the reference's nd_utility family, utils.py:416-525, as ONE kernel with
// the Lambert-W function evaluated on the device (the reference round-trips an (r, nstar) array through scipy on the
// CPU, utils.py:464-466).
//
// For every candidate image i (mu_i, sigma2_i = mean / variance of log f at x*) and every count r_j of `r`:
//   z      = exp(r sigma2 + mu) sigma2                         (U:447-448); entries with z = inf are "masked":
//            z -> 0, r sigma2 -> 0 and later r -> 0, log r! -> 0 (U:451-453, U:488-489) -- they STAY in the sums
//   lam    = r sigma2 + mu - W0(z)                              (U:466)
//   log p  = lam r - e^lam - (lam - mu)^2 / (2 sigma2) - log(e^lam sigma2 + 1) / 2 - log r!        (U:492)
//   U_i    = -sum_j p log p + e^(mu + sigma2/2) (mu + sigma2 - 1) - sum_j p log r!                 (U:512-516, U:433)
#include <math_constants.h>

#include "sgp_common.cuh"
#include "sgp_kernels.cuh"

namespace sgp {

namespace {

// Principal branch of the Lambert W function for z >= 0 (W e^W = z), to double precision.
__device__ __forceinline__ double lambert_w0(double z) {
  if (!(z > 0.0)) return 0.0;                      // z == 0 (and the masked entries)
  double w;
  if (z > 1.0e3) {
    // large arguments: Newton on w + log w = log z (no overflow, quadratic convergence from the asymptotic guess)
    const double lz = log(z);
    w = lz - log(lz);
    for (int it = 0; it < 6; ++it) {
      const double g = w + log(w) - lz;
      w -= g * w / (w + 1.0);
    }
    return w;
  }
  // Winitzki-type start, then Halley on w e^w - z
  const double l1 = log1p(z);
  w = l1 * (1.0 - log1p(l1) / (2.0 + l1));
  for (int it = 0; it < 6; ++it) {
    const double ew = exp(w);
    const double f = w * ew - z;
    const double w1 = w + 1.0;
    const double step = f / (ew * w1 - (w + 2.0) * f / (2.0 * w1));
    w -= step;
    if (fabs(step) <= 1.0e-17 * fabs(w)) break;
  }
  return w;
}

// One warp per candidate; lanes stride over the counts.  Optional (r, nstar)-shaped outputs (row-major, ld = nstar).
__global__ void __launch_bounds__(256)
nd_utility_kernel(const double* __restrict__ sigma2, const double* __restrict__ mu, const double* __restrict__ r,
                  int nr, long nstar, double* __restrict__ U, double* __restrict__ lam_out,
                  unsigned char* __restrict__ mask_out, double* __restrict__ p_out, double* __restrict__ logp_out) {
  const int lane = threadIdx.x & 31;
  const long i = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (i >= nstar) return;
  const double s2 = sigma2[i], m = mu[i];
  double h = 0.0, plf = 0.0;
  for (int j = lane; j < nr; j += 32) {
    double rj = r[j];
    double rs = rj * s2;
    double z = exp(rs + m) * s2;
    const bool keep = z != CUDART_INF;
    double lrf = lgamma(rj + 1.0);
    if (!keep) { z = 0.0; rs = 0.0; rj = 0.0; lrf = 0.0; }
    const double lam = rs + m - lambert_w0(z);
    const double el = exp(lam);
    const double d = lam - m;
    const double lp = lam * rj - el - d * d / (2.0 * s2) - 0.5 * log(el * s2 + 1.0) - lrf;
    const double p = exp(lp);
    h += p * lp;
    plf += p * lrf;
    const long o = (long)j * nstar + i;
    if (lam_out) lam_out[o] = lam;
    if (mask_out) mask_out[o] = keep ? 1 : 0;
    if (p_out) p_out[o] = p;
    if (logp_out) logp_out[o] = lp;
  }
  h = warp_sum(h);
  plf = warp_sum(plf);
  if (lane == 0 && U) U[i] = -h - (-exp(m + 0.5 * s2) * (m + s2 - 1.0) + plf);
}

}  // namespace

void nd_utility(const double* sigma2, const double* mu, const double* r, int nr, long nstar, double* U, double* lam,
                unsigned char* mask, double* p, double* logp, cudaStream_t st) {
  if (nstar <= 0 || nr <= 0) return;
  const long blocks = (nstar * 32 + 255) / 256;
  note_launch();
  nd_utility_kernel<<<(unsigned)blocks, 256, 0, st>>>(sigma2, mu, r, nr, nstar, U, lam, mask, p, logp);
}

}  // namespace sgp
