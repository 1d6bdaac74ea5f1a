// Instantiations + host launcher of the fp64 DMMA GEMM family (see gemm_f64.cuh).
#include "gemm_f64.cuh"
#include "sgp_kernels.cuh"

namespace sgp {

namespace {

template <int BM, int BN, int AL, int BL, int MODE, bool KS>
cudaError_t launch_one(const GemmArgs& a, int nsplit, cudaStream_t st) {
  dim3 grid((a.N + BN - 1) / BN, (a.M + BM - 1) / BM, nsplit);
  note_launch();
  gemm_f64_kernel<BM, BN, AL, BL, MODE, KS><<<grid, kGemmThreads, gemm_smem_bytes<BM, BN>(), st>>>(a);
  return cudaGetLastError();
}

template <int BM, int BN, int AL, int BL, int MODE, bool KS>
cudaError_t optin_one() {
  return cudaFuncSetAttribute(gemm_f64_kernel<BM, BN, AL, BL, MODE, KS>,
                              cudaFuncAttributeMaxDynamicSharedMemorySize, gemm_smem_bytes<BM, BN>());
}

// every (layout, mode) combination the Spatial_GP pass uses, in both tile sizes
#define SGP_FOR_EACH_GEMM(X)      \
  X(0, 0, kStore, false)          \
  X(0, 1, kStore, false)          \
  X(1, 0, kStore, false)          \
  X(1, 1, kStore, false)          \
  X(1, 1, kStore, true)           \
  X(0, 1, kRowdot, false)         \
  X(0, 0, kAcos, false)           \
  X(0, 0, kMoments, false)        \
  X(0, 1, kAxpyRow, false)

}  // namespace

cudaError_t gemm_f64_init() {
  cudaError_t e;
#define X(AL_, BL_, MODE_, KS_)                                                       \
  if ((e = optin_one<64, 128, AL_, BL_, MODE_, KS_>()) != cudaSuccess) return e;      \
  if ((e = optin_one<64, 64, AL_, BL_, MODE_, KS_>()) != cudaSuccess) return e;
  SGP_FOR_EACH_GEMM(X)
#undef X
  return cudaSuccess;
}

cudaError_t gemm_f64_launch(const GemmArgs& a, const GemmLaunch& cfg, cudaStream_t st) {
  if (a.M <= 0 || a.N <= 0) return cudaSuccess;
  const bool ks = a.kscale != nullptr;
  const int nsplit = cfg.nsplit > 0 ? cfg.nsplit : 1;
  if (nsplit > 1 && cfg.mode != kStore) return cudaErrorInvalidValue;
  if ((a.lda & 1) || (a.ldb & 1) || (a.ldc & 1) || (reinterpret_cast<uintptr_t>(a.A) & 15) ||
      (reinterpret_cast<uintptr_t>(a.B) & 15) || (reinterpret_cast<uintptr_t>(a.C) & 15))
    return cudaErrorMisalignedAddress;
#define X(AL_, BL_, MODE_, KS_)                                                          \
  if (cfg.al == AL_ && cfg.bl == BL_ && cfg.mode == MODE_ && ks == KS_) {                \
    return cfg.small_tile ? launch_one<64, 64, AL_, BL_, MODE_, KS_>(a, nsplit, st)      \
                          : launch_one<64, 128, AL_, BL_, MODE_, KS_>(a, nsplit, st);    \
  }
  SGP_FOR_EACH_GEMM(X)
#undef X
  return cudaErrorInvalidValue;
}

}  // namespace sgp
