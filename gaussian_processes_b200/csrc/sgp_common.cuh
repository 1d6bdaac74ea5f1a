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

// The arc-cosine nonlinearity of acosker (utils.py:984-990) for one (lin, X1*X2) pair.
// Returns K = X1X2 * J; optionally cos(delta).  No FMA contraction on 1 - c*c so the value
// matches the reference's two-step evaluation bit for bit in the common case.
__device__ __forceinline__ double acos_kernel_value(double lin, double nn, double* cos_out) {
  double c = lin / (nn + kCosEps);
  c = fmin(fmax(c, -1.0), 1.0);
  const double delta = acos(c);
  const double sq = sqrt(__dsub_rn(1.0, __dmul_rn(c, c)));
  const double J = (sq + kPi32 * c - delta * c) / kPi32;
  if (cos_out) *cos_out = c;
  return nn * J;
}

// Local partials used by every derivative in acosker (utils.py:998-1000, 1017-1021):
// P = (pi~ - delta)/pi~ ,  Q = sqrt(1-c^2)/pi~   so that  dK = P*d(lin) + Q*d(X1X2).
__device__ __forceinline__ void acos_partials(double c, double& P, double& Q) {
  const double delta = acos(c);
  P = (kPi32 - delta) / kPi32;
  Q = sqrt(__dsub_rn(1.0, __dmul_rn(c, c))) / kPi32;
}

}  // namespace sgp
