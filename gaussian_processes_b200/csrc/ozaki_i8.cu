// fp64-accurate GEMM on the 5th-generation tensor cores: C = A * B^T with the operands split into int8 slices
// (Ozaki scheme) and the slice products accumulated EXACTLY in int32 by tcgen05.mma.kind::i8 into TMEM.
//
// There is no tcgen05 kind for f64; the fp64 tensor pipe (DMMA) tops out at 37 TFLOP/s while kind::i8 runs at
// 8192 MAC/clk/SM.  Every row i of A gets a power-of-two scale 2^ea_i > max_k |A_ik| and is written as
//      A_ik = 2^(ea_i - 6) * sum_{s<S} 128^-s * qa_s[i][k],   qa_s in [-64, 64]  (round-to-nearest digits, exact
//      in fp64: the remainder of every step is representable), truncation <= 2^(ea_i - 7S);   B likewise.
// Then   C_ij = 2^(ea_i + eb_j - 12) * sum_{d<S} 128^-d * P_d[i][j],   P_d = sum_{s+t=d} qa_s qb_t^T   (int32, exact
// while K*(d+1)*4096 < 2^31), dropping only the pairs with s+t >= S (each <= 2^-7S relative to the scales).
// With S = 8 the result carries ~2^-53 of |a_i|_max |b_j|_max K -- the accuracy class of a DGEMM -- and is
// deterministic (integer accumulation has no ordering effects).
//
// Kernel shape (one CTA per SM, persistent over 128 x 64 output tiles):
//   * TMEM holds ALL S diagonals of the tile at once: diagonal d lives in columns [64d, 64d+64) (S*64 <= 512), so
//     there is a single epilogue per tile and no fp64 read-modify-write of the output;
//   * slices are stored PAIRWISE INTERLEAVED: for every row and every 64-byte k-block the bytes of slices 2j and
//     2j+1 are adjacent, so a TMA box row is a full 128-byte line ([pair][rows][128 B], 128-byte swizzle) although
//     a stage only advances K by 64; in shared memory slice 2j is k-steps 0,1 of that row and slice 2j+1 k-steps 2,3;
//   * the B slices of one parity are stacked along the MMA's N dimension: one tcgen05.mma with N = 64*c multiplies
//     slice s of A with slices t = pi, pi+2, .. of B and lands in the c diagonals s+t, which are adjacent because
//     TMEM orders the diagonals even-first (0,2,4,6,1,3,5,7) -- S(S+1)/2 slice pairs cost about 2S instructions
//     per 32-deep k-step, each at the full-width rate;
//   * warp 0 = TMA producer, warp 1 = MMA issuer (one thread), warps 2-5 = epilogue (TMEM -> fp64 Horner sum over
//     the diagonals -> scaled store), mbarrier pipelines between them.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>

#include "sgp_common.cuh"
#include "sgp_kernels.cuh"

namespace sgp {

namespace {

constexpr int kOzBM = 128;          // tile rows (TMEM lanes)
constexpr int kOzBN = 64;           // tile columns per diagonal
constexpr int kOzBK = 64;           // K elements per stage; a shared-memory row is 128 B = this k-block of 2 slices
constexpr int kOzRow = 2 * kOzBK;
constexpr int kOzStages = 2;
constexpr int kOzMaxS = 8;
constexpr int kOzThreads = 192;
constexpr int kOzPairA = kOzBM * kOzRow;              // 16 KB: one slice pair of the A tile
constexpr int kOzPairB = kOzBN * kOzRow;              //  8 KB
constexpr int kOzStageA = kOzMaxS / 2 * kOzPairA;     // 64 KB
constexpr int kOzStageB = kOzMaxS / 2 * kOzPairB;     // 32 KB
constexpr int kOzSmem = kOzStages * (kOzStageA + kOzStageB) + 1024 /*alignment*/ + 256 /*barriers*/;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  while (!ok) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
  }
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const void* map, uint32_t bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, int8 x int8 -> int32
__device__ __forceinline__ void mma_i8(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand tile [rows][128 B] with the 128-byte swizzle TMA wrote: 8-row groups are 1024 B apart; a k-step
// inside the row is selected by advancing the start address (the swizzle is a function of the address bits).
// (cute::UMMA::SmemDescriptor: start>>4 | LBO>>4 <<16 | SBO>>4 <<32 | version 1 <<46 | layout SWIZZLE_128B=2 <<61)
__device__ __forceinline__ uint64_t kmajor_sw128_desc(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// TMEM column block of diagonal d (even diagonals first)
__host__ __device__ constexpr int diag_pos(int d) { return (d & 1) * 4 + (d >> 1); }
// cute::UMMA::InstrDescriptor for kind::i8: c = S32 (2<<4), a = b = signed int8 (1<<7, 1<<10), K-major both,
// N>>3 at bit 17, M>>4 at bit 24.
__host__ __device__ constexpr uint32_t i8_idesc(int n) {
  return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kOzBM >> 4) << 24);
}

struct OzArgs {
  const int* ea; const int* eb;   // row exponents of A and of B
  double* C; long ldc;
  int M, N, K, S;
};

__device__ __forceinline__ double pow2d(int e) {   // 2^e for -1022 <= e <= 1023
  return __longlong_as_double((long long)(1023 + e) << 52);
}

__global__ void __launch_bounds__(kOzThreads, 1)
oz_gemm_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, const OzArgs a) {
  extern __shared__ uint8_t oz_smem_raw[];
  const uint32_t raw = smem_u32(oz_smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t bars = base + kOzStages * (kOzStageA + kOzStageB);
  // barrier slots (8 B each): full[0..1], empty[0..1], tmem_full, tmem_empty; then the TMEM base address
  const uint32_t bar_full = bars, bar_empty = bars + 16, bar_tfull = bars + 32, bar_tempty = bars + 40;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(oz_smem_raw + (bars + 48 - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = a.S;
  const int num_nt = (a.N + kOzBN - 1) / kOzBN;
  const int num_tiles = ((a.M + kOzBM - 1) / kOzBM) * num_nt;
  const int nkb = (a.K + kOzBK - 1) / kOzBK;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kOzStages; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
    mbar_init(bar_tfull, 1);
    mbar_init(bar_tempty, 128);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      const uint32_t bytes = (uint32_t)((S + 1) / 2) * (kOzPairA + kOzPairB);
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile / num_nt) * kOzBM, n0 = (tile % num_nt) * kOzBN;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(bar_empty + 8 * stage, phase ^ 1);
          const uint32_t sA = base + stage * (kOzStageA + kOzStageB), sB = sA + kOzStageA;
          mbar_expect_tx(bar_full + 8 * stage, bytes);
          tma_load_3d(sA, &mapA, bar_full + 8 * stage, kb * kOzRow, m0, 0);
          tma_load_3d(sB, &mapB, bar_full + 8 * stage, kb * kOzRow, n0, 0);
          if (++stage == kOzStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0, tphase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(bar_tempty, tphase ^ 1);
        tc_fence_after();
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(bar_full + 8 * stage, phase);
          tc_fence_after();
          const uint32_t sA = base + stage * (kOzStageA + kOzStageB), sB = sA + kOzStageA;
#pragma unroll
          for (int ks = 0; ks < kOzBK / 32; ++ks) {
            for (int s = 0; s < S; ++s) {
              const uint64_t adesc = kmajor_sw128_desc(sA + (s >> 1) * kOzPairA + (s & 1) * kOzBK + ks * 32);
              const uint32_t acc = (kb | ks | s) != 0;
#pragma unroll
              for (int par = 0; par < 2; ++par) {          // B slices t = par, par+2, ..  with s + t <= S-1
                const int last = S - 1 - s - par;
                if (last < 0) continue;
                const int c = (last >> 1) + 1;
                const uint64_t bdesc = kmajor_sw128_desc(sB + par * kOzBK + ks * 32);
                mma_i8(tmem + (uint32_t)(kOzBN * diag_pos(s + par)), adesc, bdesc, i8_idesc(kOzBN * c), acc);
              }
            }
          }
          tc_commit(bar_empty + 8 * stage);
          if (kb == nkb - 1) tc_commit(bar_tfull);
          if (++stage == kOzStages) { stage = 0; phase ^= 1; }
        }
        tphase ^= 1;
      }
    }
  } else {
    // ===== epilogue: TMEM -> fp64 Horner sum over the diagonals -> scaled store =====
    const int q = warp & 3;                       // TMEM lane quarter this warp may read
    uint32_t tphase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m0 = (tile / num_nt) * kOzBM, n0 = (tile % num_nt) * kOzBN;
      const long row = m0 + 32 * q + lane;
      const bool row_ok = row < a.M;
      const double sa = row_ok ? pow2d(a.ea[row] - 12) : 0.0;
      mbar_wait(bar_tfull, tphase);
      tc_fence_after();
      const uint32_t tl = tmem + ((uint32_t)(32 * q) << 16);
#pragma unroll 1
      for (int c = 0; c < kOzBN / 16; ++c) {
        double acc[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) acc[j] = 0.0;
        for (int d = S - 1; d >= 0; --d) {
          uint32_t v[16];
          tmem_ld16(tl + (uint32_t)(kOzBN * diag_pos(d) + 16 * c), v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] = fma(acc[j], 0.0078125, (double)(int)v[j]);
        }
        if (row_ok) {
          double* out = a.C + row * a.ldc + n0 + 16 * c;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int col = n0 + 16 * c + j;
            if (col < a.N) out[j] = acc[j] * sa * pow2d(a.eb[col]);
          }
        }
      }
      tc_fence_before();
      mbar_arrive(bar_tempty);
      tphase ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
  }
}

// One warp per row: power-of-two row scale, then S round-to-nearest base-128 digits per element.
__global__ void __launch_bounds__(256)
oz_slice_rows_kernel(const double* __restrict__ X, long ldx, long rows, int cols, int S, int8_t* __restrict__ q,
                     int ldk, long pair_stride, int* __restrict__ exps) {
  const int lane = threadIdx.x & 31;
  const long warp0 = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long nwarps = ((long)gridDim.x * blockDim.x) >> 5;
  for (long r = warp0; r < rows; r += nwarps) {
    const double* x = X + r * ldx;
    double m = 0.0;
    for (int c = lane; c < cols; c += 32) m = fmax(m, fabs(x[c]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    const int e = (m > 0.0 && m < 1.0e300) ? ilogb(m) + 1 : 0;
    if (lane == 0) exps[r] = e;
    for (int g = lane; g < ldk / 8; g += 32) {
      double t[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int c = 4 * g + j;
        t[j] = c < cols ? scalbn(x[c], 6 - e) : 0.0;
      }
      for (int s = 0; s < S; ++s) {
        uint32_t w = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const double d = rint(t[j]);
          t[j] = (t[j] - d) * 128.0;
          w |= ((uint32_t)(__double2int_rn(d)) & 0xffu) << (8 * j);
        }
        // pair s/2, row r, k-block (4g)/64: [slice 2j: 64 B][slice 2j+1: 64 B]
        const int k = 4 * g;
        *reinterpret_cast<uint32_t*>(q + (s >> 1) * pair_stride + r * ldk + (k >> 6) * 128 + (s & 1) * 64 + (k & 63)) = w;
      }
    }
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// [pairs][rows][ldk] int8 (ldk = 2 * K rounded up to 64), box = 128 B x box_rows rows x pairs, 128-byte swizzle
bool make_slice_map(CUtensorMap* map, const int8_t* q, long rows, int ldk, int S, int box_rows) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return false;
  const int pairs = (S + 1) / 2;
  const cuuint64_t dims[3] = {(cuuint64_t)ldk, (cuuint64_t)rows, (cuuint64_t)pairs};
  const cuuint64_t strides[2] = {(cuuint64_t)ldk, (cuuint64_t)rows * (cuuint64_t)ldk};
  const cuuint32_t box[3] = {(cuuint32_t)kOzRow, (cuuint32_t)box_rows, (cuuint32_t)pairs};
  const cuuint32_t estr[3] = {1, 1, 1};
  return fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<int8_t*>(q), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace

int oz_ldk(int K) { return 2 * ((K + 63) & ~63); }   // bytes per row of one slice PAIR
size_t oz_slice_bytes(long rows, int K, int S) { return (size_t)((S + 1) / 2) * rows * oz_ldk(K); }

cudaError_t oz_init() {
  return cudaFuncSetAttribute(oz_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kOzSmem);
}

cudaError_t oz_slice_rows(const double* X, long ldx, long rows, int cols, int S, int8_t* q, int ldk, int* exps,
                          cudaStream_t st) {
  if (rows <= 0) return cudaSuccess;
  if (S < 1 || S > kOzMaxS || ldk != oz_ldk(cols)) return cudaErrorInvalidValue;
  const long blocks = std::min<long>((rows + 7) / 8, 148L * 8);
  note_launch();
  oz_slice_rows_kernel<<<(unsigned)blocks, 256, 0, st>>>(X, ldx, rows, cols, S, q, ldk, rows * (long)ldk, exps);
  return cudaGetLastError();
}

// C (M x N, ld = ldc) = A * B^T from the int8 slices of A (M rows) and B (N rows); K bytes per row, pitch ldk.
cudaError_t oz_gemm(const int8_t* qa, const int* ea, long M, const int8_t* qb, const int* eb, int N, int K, int ldk,
                    int S, double* C, long ldc, int sm_count, cudaStream_t st) {
  if (M <= 0 || N <= 0) return cudaSuccess;
  if (S < 1 || S > kOzMaxS || (long)K * S * 4096 >= (1L << 31)) return cudaErrorInvalidValue;
  CUtensorMap mapA, mapB;
  if (ldk != oz_ldk(K)) return cudaErrorInvalidValue;
  if (!make_slice_map(&mapA, qa, M, ldk, S, kOzBM) || !make_slice_map(&mapB, qb, N, ldk, S, kOzBN))
    return cudaErrorNotSupported;
  OzArgs a{ea, eb, C, ldc, (int)M, N, K, S};
  const long tiles = ((M + kOzBM - 1) / kOzBM) * ((N + kOzBN - 1) / kOzBN);
  const int grid = (int)std::min<long>(tiles, sm_count > 0 ? sm_count : 148);
  note_launch();
  oz_gemm_kernel<<<grid, kOzThreads, kOzSmem, st>>>(mapA, mapB, a);
  return cudaGetLastError();
}

}  // namespace sgp
