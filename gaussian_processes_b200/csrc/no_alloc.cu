// sgp_cholesky_noalloc: the blocked lower Cholesky of sgp_cholesky with its scratch in a file-scope device array instead
// of the handle's cudaMalloc'ed arena.  NO counterpart in the reference; added for
// callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// The entry point calls no cudaMalloc / cudaFree / allocating library routine and takes no sgp_handle (creating and
// first using a handle allocates).  Scratch = the inverses of the 64 x 64 diagonal blocks of L that the panel solves
// consume (cholesky_lower's `Winv`) + one status word, for n <= kNoAllocMaxN: 16 x 64 x 64 doubles + 256 bytes
// = 512.25 KiB, held by every GPU that loads the library until the process exits.
//
// In HOST code the array's name below is a host-side stand-in, not a device address: the device address is taken once
// per device, with that device current, by cudaGetSymbolAddress, and that pointer is what the kernels receive.
//
// Every call shares the one array and it keeps the last call's contents: the status word is cleared at the start of
// each call (the block inverses are fully overwritten before they are read), and calls that OVERLAP on the GPU (two
// streams, two threads on one device) race on it -- serialise them.
#include "../../include/sgp_b200.h"

#include <cstdint>
#include <cstdio>
#include <mutex>

#include "gemm_f64.cuh"
#include "sgp_kernels.cuh"

namespace {

constexpr int kNoAllocMaxN = 1024;                                   // stated maximum order of the matrix
constexpr long kWinvElems = (long)(kNoAllocMaxN / 64) * 64 * 64;     // 65 536 doubles = 512 KiB
constexpr long kScratchElems = kWinvElems + 32;                      // + 256 bytes: status word (and padding)

static __device__ __align__(256) double g_noalloc_scratch[kScratchElems];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
double* g_scratch_ptr[kMaxDevices] = {nullptr};    // DEVICE address of g_noalloc_scratch on each device (0 = not set up)

int fail(char* errbuf, int errlen, int code, const char* msg, const char* detail = nullptr) {
  if (errbuf && errlen > 0) std::snprintf(errbuf, (size_t)errlen, "sgp_cholesky_noalloc: %s%s%s", msg, detail ? ": " : "",
                                          detail ? detail : "");
  return code;
}

}  // namespace

extern "C" int sgp_cholesky_noalloc(double* A, long ld, int n, int* flags_dev, void* stream, char* errbuf,
                                    int errbuf_len) {
  if (errbuf && errbuf_len > 0) errbuf[0] = 0;
  if (!A) return fail(errbuf, errbuf_len, SGP_E_ARG, "A is null");
  if (n <= 0) return fail(errbuf, errbuf_len, SGP_E_ARG, "n must be positive");
  if (n > kNoAllocMaxN) {
    char d[96];
    std::snprintf(d, sizeof d, "n = %d exceeds the static scratch (maximum n = %d); use sgp_cholesky", n, kNoAllocMaxN);
    return fail(errbuf, errbuf_len, SGP_E_UNSUPPORTED, "size beyond the maximum", d);
  }
  if (ld < n || (ld & 1) || (reinterpret_cast<uintptr_t>(A) & 15))
    return fail(errbuf, errbuf_len, SGP_E_ARG, "ld must be even and >= n, A 16-byte aligned");
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(errbuf, errbuf_len, SGP_E_CUDA, "cudaGetDevice", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return fail(errbuf, errbuf_len, SGP_E_ARG, "device index out of range");
  double* scratch = nullptr;
  {
    // set-up, once per device, with that device current: device address of the array + the GEMM kernels' opt-in
    // shared-memory attributes (neither allocates)
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_ptr[dev]) {
      double* p = nullptr;
      e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
      if (e != cudaSuccess || !p)
        return fail(errbuf, errbuf_len, SGP_E_CUDA, "cudaGetSymbolAddress", cudaGetErrorString(e));
      e = sgp::gemm_f64_init();
      if (e != cudaSuccess) return fail(errbuf, errbuf_len, SGP_E_CUDA, "gemm_f64_init", cudaGetErrorString(e));
      e = sgp::chol_init();
      if (e != cudaSuccess) return fail(errbuf, errbuf_len, SGP_E_CUDA, "chol_init", cudaGetErrorString(e));
      g_scratch_ptr[dev] = p;
    }
    scratch = g_scratch_ptr[dev];
  }
  cudaStream_t st = (cudaStream_t)stream;
  double* winv = scratch;
  int* flags = reinterpret_cast<int*>(scratch + kWinvElems);
  // the array keeps the last call's contents: the status word must start at zero (the kernels OR bits into it)
  e = cudaMemsetAsync(flags, 0, 4 * sizeof(int), st);
  if (e != cudaSuccess) return fail(errbuf, errbuf_len, SGP_E_CUDA, "cudaMemsetAsync", cudaGetErrorString(e));
  e = sgp::cholesky_lower(A, ld, n, flags, sgp::kFlagKtNotPD, sgp::kFlagKtSmallPivot, winv, st);
  if (e != cudaSuccess) return fail(errbuf, errbuf_len, SGP_E_CUDA, "cholesky_lower", cudaGetErrorString(e));
  if (flags_dev) {
    e = cudaMemcpyAsync(flags_dev, flags, sizeof(int), cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) return fail(errbuf, errbuf_len, SGP_E_CUDA, "cudaMemcpyAsync", cudaGetErrorString(e));
  }
  return SGP_OK;
}
