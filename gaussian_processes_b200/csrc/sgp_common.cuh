// Shared device helpers for the Spatial_GP sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace sgp {

// Reference constant: utils.py:25 overwrites torch.pi with a float32-valued pi.
constexpr double kPi32 = 3.1415927410125732;
constexpr double kCosEps = 1.0e-7;     // utils.py:984
constexpr double kMaskThr = 1.0e-3;    // utils.py:883

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide sum of one double per thread; result valid in thread 0. `red` needs >= 32 doubles.
__device__ __forceinline__ double block_sum(double v, double* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  double r = 0.0;
  if (warp == 0) {
    const int nw = (blockDim.x + 31) >> 5;
    r = lane < nw ? red[lane] : 0.0;
    r = warp_sum(r);
  }
  return r;
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, int src_bytes) {
  const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(gsrc), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// D(8x8) += A(8x4) * B(4x8), fp64 tensor core (SASS: DMMA.8x8x4).
// lane = 4*g + t : a = A[g][t], b = B[t][g], c0/c1 = C[g][2t], C[g][2t+1].
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// ---- fp64 reciprocal / square root / arc cosine without libdevice ---------------------------------------------------
// The N x m elementwise passes of the arc-cosine kernel were ISSUE-bound on libdevice's acos / sqrt / division (ncu:
// 181 executed instructions per element, 77 % of the issue slots, half of them moves that materialise 64-bit polynomial
// coefficients and the slow-path guards of division / sqrt).  These versions are branch-free, take their coefficients
// from the constant bank as DFMA operands, and are accurate to 1 ulp (acos: the rational approximation of Sun's fdlibm
// e_acos.c -- restated, |error| <= 1 ulp against numpy.arccos on 2.4e6 points: tests/test_acos_restatement.py keeps the
// numpy restatement honest); the parity bar of the path is 1e-9, not bit-identity.
__device__ __forceinline__ double rcp_nr(double x) {          // x normal and nonzero
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));        // MUFU.RCP64H: ~20 bits
  double e = fma(-x, y, 1.0);
  y = fma(y, e, y);
  e = fma(-x, y, 1.0);
  return fma(y, e, y);
}
__device__ __forceinline__ double div_nr(double a, double b) {  // a / b, b normal and nonzero; faithfully rounded
  const double y = rcp_nr(b);
  const double q = a * y;
  return fma(fma(-b, q, a), y, q);
}
__device__ __forceinline__ double sqrt_nr(double x) {          // x >= 0 and either 0 or normal
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));      // MUFU.RSQ64H
  double g = x * y, h = 0.5 * y;
  double r = fma(-g, h, 0.5);
  g = fma(g, r, g); h = fma(h, r, h);
  r = fma(-g, h, 0.5);
  g = fma(g, r, g); h = fma(h, r, h);
  g = fma(fma(-g, g, x), h, g);
  return x > 0.0 ? g : 0.0;
}

static __constant__ double kAcosP[6] = {1.66666666666666657415e-01, -3.25565818622400915405e-01,
                                        2.01212532134862925881e-01, -4.00555345006794114027e-02,
                                        7.91534994289814532176e-04, 3.47933107596021167570e-05};
static __constant__ double kAcosQ[4] = {-2.40339491173441421878e+00, 2.02094576023350569471e+00,
                                        -6.88283971605453293030e-01, 7.70381505559019352791e-02};
constexpr double kPio2Hi = 1.57079632679489655800e+00, kPio2Lo = 6.12323399573676603587e-17;
constexpr double kPiTrue = 3.14159265358979311600e+00;
constexpr double kInvPi32 = 1.0 / kPi32;

// delta = acos(c) for c in [-1, 1].  |c| < 1/2: pi/2 - (c + c R(c^2));  otherwise with z = (1 - |c|)/2, s = sqrt(z):
// 2 (s + s R(z)) for c > 0 and pi - 2 (s + s R(z)) for c < 0;  R = P/Q of degrees 6/4.
__device__ __forceinline__ double acos_nr(double c) {
  const double a = fabs(c);
  const bool big = a >= 0.5;
  const double z = big ? 0.5 * (1.0 - a) : c * c;
  double pz = fma(z, kAcosP[5], kAcosP[4]);
  pz = fma(z, pz, kAcosP[3]); pz = fma(z, pz, kAcosP[2]); pz = fma(z, pz, kAcosP[1]); pz = fma(z, pz, kAcosP[0]);
  pz *= z;
  double qz = fma(z, kAcosQ[3], kAcosQ[2]);
  qz = fma(z, qz, kAcosQ[1]); qz = fma(z, qz, kAcosQ[0]); qz = fma(z, qz, 1.0);
  const double r = pz * rcp_nr(qz);
  const double s = sqrt_nr(z);
  const double small = kPio2Hi - (c - (kPio2Lo - c * r));
  const double bigp = 2.0 * fma(s, r, s);
  const double bign = kPiTrue - 2.0 * (s + fma(r, s, -kPio2Lo));
  return big ? (c > 0.0 ? bigp : bign) : small;
}

// The arc-cosine nonlinearity of acosker (utils.py:984-990) for one (lin, X1*X2) pair.
// Returns K = X1X2 * J; optionally cos(delta).  1 - c*c is evaluated in the reference's two steps (no FMA contraction):
// near c = 1 (an inducing image that is also a training image) the rounding of c*c is what the reference's sqrt sees.
__device__ __forceinline__ double acos_kernel_value(double lin, double nn, double* cos_out) {
  double c = div_nr(lin, nn + kCosEps);
  c = fmin(fmax(c, -1.0), 1.0);
  const double delta = acos_nr(c);
  const double sq = sqrt_nr(__dsub_rn(1.0, __dmul_rn(c, c)));
  const double J = (sq + kPi32 * c - delta * c) * kInvPi32;
  if (cos_out) *cos_out = c;
  return nn * J;
}

// Local partials used by every derivative in acosker (utils.py:998-1000, 1017-1021):
// P = (pi~ - delta)/pi~ ,  Q = sqrt(1-c^2)/pi~   so that  dK = P*d(lin) + Q*d(X1X2).
__device__ __forceinline__ void acos_partials(double c, double& P, double& Q) {
  const double delta = acos_nr(c);
  P = (kPi32 - delta) * kInvPi32;
  Q = sqrt_nr(__dsub_rn(1.0, __dmul_rn(c, c))) * kInvPi32;
}

}  // namespace sgp
