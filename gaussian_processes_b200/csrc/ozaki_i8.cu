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
//   * warp 0 = TMA producer, warp 1 = MMA issuer (one thread), warps 2-9 = epilogue (TMEM -> fp64 Horner sum over
//     the diagonals -> scaled store), mbarrier pipelines between them.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>

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
constexpr int kOzThreads = 320;         // TMA warp, MMA warp, 8 epilogue warps
constexpr int kOzPairA = kOzBM * kOzRow;              // 16 KB: one slice pair of the A tile
constexpr int kOzPairB = kOzBN * kOzRow;              //  8 KB
constexpr int kOzStageA = kOzMaxS / 2 * kOzPairA;     // 64 KB
constexpr int kOzStageB = kOzMaxS / 2 * kOzPairB;     // 32 KB
constexpr int kOzSmem = kOzStages * (kOzStageA + kOzStageB) + 1024 /*alignment*/ + 64 /*barriers*/ +
                        8 * 32 * 16 * 8 /*store staging of the 8 epilogue warps*/;

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
// 16 lanes x 16 columns in the accumulator-fragment layout (cute SM100_TMEM_LOAD_16dp256b2x): thread t of the warp
// holds, for lane rows r = t/4 and r + 8 of the 16 addressed lanes and column pairs 2(t%4), 2(t%4)+1 (+8):
//   v[0..1] = (r, 2c..2c+1)  v[2..3] = (r+8, 2c..)  v[4..5] = (r, 8+2c..)  v[6..7] = (r+8, 8+2c..)
__device__ __forceinline__ void tmem_ld_frag16(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
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

// All MMAs of one stage (two 32-deep k-steps): slice s of A against the B slices t = par, par+2, .. with s + t < S.
// dA / dB: descriptors of the stage's A / B tiles; byte offsets inside the tile go into the 16-byte-unit address field.
template <int S>
__device__ __forceinline__ void issue_stage(uint64_t dA, uint64_t dB, uint32_t tmem, uint32_t first_acc) {
#pragma unroll
  for (int ks = 0; ks < kOzBK / 32; ++ks) {
#pragma unroll
    for (int s = 0; s < S; ++s) {
#pragma unroll
      for (int par = 0; par < 2; ++par) {
        const int last = S - 1 - s - par;
        if (last < 0) continue;
        const int c = (last >> 1) + 1;
        mma_i8(tmem + (uint32_t)(kOzBN * diag_pos(s + par)),
               dA + (uint64_t)(((s >> 1) * kOzPairA + (s & 1) * kOzBK + ks * 32) >> 4),
               dB + (uint64_t)((par * kOzBK + ks * 32) >> 4), i8_idesc(kOzBN * c), (ks | s) ? 1u : first_acc);
      }
    }
  }
}

// SGP_OZ_DEBUG & 64: CTA 0 writes where its MMA-issuing thread and its first epilogue warp spent their cycles
// [0] MMA: waiting for TMEM to be released  [1] MMA: waiting for operand stages  [2] MMA: whole loop
// [3] epilogue: waiting for the tile's MMAs  [4] epilogue: TMEM drain  [5] epilogue: scale + store  [6] tiles of CTA 0
__device__ double g_oz_prof[8];

struct OzArgs {
  const int* ea; const int* eb;   // row exponents of A and of B
  double* C; long ldc;
  int M, N, K, S;
  int nz, kb_per_z; long zstride; int accumulate;   // split-K: slice z covers k-blocks [z*kb_per_z, ..) -> C + z*zstride
  double alpha; int lower_only;   // C = alpha * A B^T; lower_only (M == N): only tiles touching the lower triangle
  int num_tiles;                  // tiles enumerated by tile_origin()
  // optional moments epilogue (T = K Mfull^T, utils.py:1090-1101): per 16-column group g of the output,
  //   part1[g*ldpart + row] = sum over the 16-column group g of C*aux
  const double* aux; long ldaux; double* part1; long ldpart;
  int dbg;   // timing experiments only: 1 = skip A loads, 2 = skip B loads, 4 = every tile loads A rows 0..127
};

// origin of tile t: row-major over (m-tile, n-tile); with lower_only row-tile mt keeps its first min(2mt+2, num_nt)
// n-tiles (the ones that touch the lower triangle: 128-row tiles against 64-column tiles)
__device__ __forceinline__ void tile_origin(int lower_only, int tile, int num_nt, int& m0, int& n0) {
  if (!lower_only) { m0 = (tile / num_nt) * kOzBM; n0 = (tile % num_nt) * kOzBN; return; }
  int mt = 0, base = 0;
  for (;;) {
    const int cnt = min(2 * mt + 2, num_nt);
    if (tile < base + cnt) break;
    base += cnt; ++mt;
  }
  m0 = mt * kOzBM; n0 = (tile - base) * kOzBN;
}

__device__ __forceinline__ double pow2d(int e) {   // 2^e for -1022 <= e <= 1023
  return __longlong_as_double((long long)(1023 + e) << 52);
}

// x * 2^k, exact; the plain multiply whenever 2^k is a normal number (always, except for operands of denormal scale)
__device__ __forceinline__ double scale_pow2(double x, int k) {
  return (k >= -1000 && k <= 1000) ? x * pow2d(k) : scalbn(x, k);
}

// One base-128 digit of the running remainder w (|w| <= 2^(6-7s) on entry): adding 1.5 * 2^(52-7s) rounds w to the
// nearest multiple of 2^(-7s) (ties to even) and leaves that multiple, as a two's-complement integer, in the low
// mantissa bits; the remainder update is exact.  Three fp64 adds per digit, no float<->int conversion instructions.
template <int s>
__device__ __forceinline__ uint32_t take_digit(double& w) {
  const double magic = __longlong_as_double((long long)(1023 + 52 - 7 * s) << 52) * 1.5;
  const double u = w + magic;
  w -= (u - magic);
  return (uint32_t)__double2loint(u) & 0xffu;
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
  const int num_tiles = a.num_tiles;
  const int num_items = num_tiles * a.nz;             // item = z * num_tiles + tile
  const int nkb_all = (a.K + kOzBK - 1) / kOzBK;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kOzStages; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
    mbar_init(bar_tfull, 1);
    mbar_init(bar_tempty, 256);
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
      int stage = 0; uint32_t phase = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
        const int tile = item % num_tiles, z = item / num_tiles;
        int m0, n0;
        tile_origin(a.lower_only, tile, num_nt, m0, n0);
        const int kb0 = z * a.kb_per_z, kb1 = min(nkb_all, kb0 + a.kb_per_z);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(bar_empty + 8 * stage, phase ^ 1);
          const uint32_t sA = base + stage * (kOzStageA + kOzStageB), sB = sA + kOzStageA;
          const uint32_t pairs = (uint32_t)((S + 1) / 2);
          mbar_expect_tx(bar_full + 8 * stage, ((a.dbg & 1) ? 0u : pairs * kOzPairA) + ((a.dbg & 2) ? 0u : pairs * kOzPairB));
          if (!(a.dbg & 1)) tma_load_3d(sA, &mapA, bar_full + 8 * stage, kb * kOzRow, (a.dbg & 4) ? 0 : m0, 0);
          if (!(a.dbg & 2)) tma_load_3d(sB, &mapB, bar_full + 8 * stage, kb * kOzRow, n0, 0);
          if (++stage == kOzStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0, tphase = 0;
      const bool prof = (a.dbg & 64) && blockIdx.x == 0;
      long long c_te = 0, c_full = 0, c_all = prof ? clock64() : 0, ntile = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
        const int z = item / num_tiles;
        const int kb0 = z * a.kb_per_z, kb1 = min(nkb_all, kb0 + a.kb_per_z);
        long long c0 = prof ? clock64() : 0;
        mbar_wait(bar_tempty, tphase ^ 1);
        if (prof) { c_te += clock64() - c0; ++ntile; }
        tc_fence_after();
        for (int kb = kb0; kb < kb1; ++kb) {
          c0 = prof ? clock64() : 0;
          mbar_wait(bar_full + 8 * stage, phase);
          if (prof) c_full += clock64() - c0;
          tc_fence_after();
          const uint32_t sA = base + stage * (kOzStageA + kOzStageB), sB = sA + kOzStageA;
          if (!(a.dbg & 16)) {
            const uint64_t dA = kmajor_sw128_desc(sA), dB = kmajor_sw128_desc(sB);
            const uint32_t first = kb != kb0;
            switch (S) {     // fully unrolled per slice count: the issuing thread must not be the bottleneck
              case 8: issue_stage<8>(dA, dB, tmem, first); break;
              case 7: issue_stage<7>(dA, dB, tmem, first); break;
              case 6: issue_stage<6>(dA, dB, tmem, first); break;
              case 5: issue_stage<5>(dA, dB, tmem, first); break;
              case 4: issue_stage<4>(dA, dB, tmem, first); break;
              case 3: issue_stage<3>(dA, dB, tmem, first); break;
              default: issue_stage<2>(dA, dB, tmem, first); break;
            }
          }
          tc_commit(bar_empty + 8 * stage);
          if (kb == kb1 - 1) tc_commit(bar_tfull);
          if (++stage == kOzStages) { stage = 0; phase ^= 1; }
        }
        tphase ^= 1;
      }
      if (prof) { g_oz_prof[0] = (double)c_te; g_oz_prof[1] = (double)c_full; g_oz_prof[2] = (double)(clock64() - c_all); g_oz_prof[6] = (double)ntile; }
    }
  } else {
    // ===== epilogue (8 warps): TMEM -> fp64 Horner sum over the diagonals in registers -> release TMEM -> scale and
    //       store (overlaps the next tile's MMAs).  The accumulators are read in the FRAGMENT layout (16x256b): a thread
    //       holds 4 rows x (2 + 2) adjacent columns of each 16-column group, so the output goes out as 16-byte stores,
    //       64 contiguous bytes per row and instruction, straight from registers -- no shared-memory transpose, no warp
    //       barriers.  (Round 2: with the row-per-thread layout the staging round trips of the store phase competed with
    //       the tensor core's operand reads for shared-memory bandwidth: MMAs and epilogue needed ~13.5 us per tile each
    //       but 20 us together at K = 500, profiles/r2_ncu_summary.md.) =====
    const int q = warp & 3;                       // TMEM lane quarter this warp may read
    const int h = (warp - 2) >> 2;                // column half of the tile this warp owns
    const int rq = lane >> 2, cq = lane & 3;
    const bool mom = a.part1 != nullptr;
    const bool vecC = ((a.ldc & 1) == 0) && ((reinterpret_cast<uintptr_t>(a.C) & 15) == 0) && ((a.zstride & 1) == 0);
    const bool vecA = mom && ((a.ldaux & 1) == 0) && ((reinterpret_cast<uintptr_t>(a.aux) & 15) == 0);
    uint32_t tphase = 0;
    const bool eprof = (a.dbg & 64) && blockIdx.x == 0 && warp == 2;
    long long e_wait = 0, e_drain = 0, e_store = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      const int tile = item % num_tiles;
      double* Cz = a.C + (long)(item / num_tiles) * a.zstride;
      int m0, n0;
      tile_origin(a.lower_only, tile, num_nt, m0, n0);
      n0 += 32 * h;
      const long long e0 = eprof ? clock64() : 0;
      const long row0 = m0 + 32 * q + rq;         // this thread's rows: row0 + 8 * ai, ai = 0..3
      double sa[4];
#pragma unroll
      for (int ai = 0; ai < 4; ++ai) sa[ai] = (row0 + 8 * ai < a.M) ? pow2d(a.ea[row0 + 8 * ai] - 12) * a.alpha : 0.0;
      mbar_wait(bar_tfull, tphase);
      const long long e1 = eprof ? clock64() : 0;
      tc_fence_after();
      const uint32_t tl = tmem + ((uint32_t)(32 * q) << 16) + 32 * h;
      double acc[2][16];                          // [cg][4 * ai + 2 * j + e]: row row0 + 8 ai, column 8 j + 2 cq + e
#pragma unroll
      for (int cg = 0; cg < 2; ++cg) {
#pragma unroll
        for (int j = 0; j < 16; ++j) acc[cg][j] = 0.0;
        if (a.dbg & 8) continue;
#pragma unroll
        for (int dh = 6; dh >= 0; dh -= 2) {      // diagonals dh+1, dh: four TMEM loads (2 lane halves each) in flight
          uint32_t v[2][2][8];
#pragma unroll
          for (int i = 0; i < 2; ++i)
            if (dh + 1 - i < S) {
#pragma unroll
              for (int h2 = 0; h2 < 2; ++h2)
                tmem_ld_frag16(tl + ((uint32_t)(16 * h2) << 16) + (uint32_t)(kOzBN * diag_pos(dh + 1 - i) + 16 * cg), v[i][h2]);
            }
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 2; ++i)
            if (dh + 1 - i < S) {
#pragma unroll
              for (int h2 = 0; h2 < 2; ++h2)
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                  // register k of the fragment: column block j = k >> 2, second row (+8) = (k >> 1) & 1, parity e = k & 1
                  const int idx = 4 * (2 * h2 + ((k >> 1) & 1)) + 2 * (k >> 2) + (k & 1);
                  // int32 -> double through the 2^52 + 2^31 offset (exact; keeps the conversion off the I2F path)
                  const double dv = __hiloint2double(0x43300000, (int)(v[i][h2][k] ^ 0x80000000u)) - 4503601774854144.0;
                  acc[cg][idx] = fma(acc[cg][idx], 0.0078125, dv);
                }
            }
        }
      }
      tc_fence_before();
      mbar_arrive(bar_tempty);                    // TMEM is free: the next tile's MMAs may start
      tphase ^= 1;
      const long long e2 = eprof ? clock64() : 0;
#pragma unroll
      for (int cg = 0; cg < 2; ++cg) {
        if (a.dbg & 32) break;
        const int c0 = n0 + 16 * cg;
        double sb[2][2];
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int col = c0 + 8 * j + 2 * cq + e;
            sb[j][e] = col < a.N ? pow2d(a.eb[col]) : 0.0;
          }
        double ax[16];
        if (mom) {                                // aux (= K) for the whole group, requested before the stores
#pragma unroll
          for (int ai = 0; ai < 4; ++ai)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const long grow = row0 + 8 * ai;
              const int col = c0 + 8 * j + 2 * cq;
              const double* src = a.aux + grow * a.ldaux + col;
              double x0 = 0.0, x1 = 0.0;
              if (grow < a.M) {
                if (vecA && col + 1 < a.N) { const double2 t = __ldg(reinterpret_cast<const double2*>(src)); x0 = t.x; x1 = t.y; }
                else { if (col < a.N) x0 = __ldg(src); if (col + 1 < a.N) x1 = __ldg(src + 1); }
              }
              ax[4 * ai + 2 * j] = x0; ax[4 * ai + 2 * j + 1] = x1;
            }
        }
#pragma unroll
        for (int ai = 0; ai < 4; ++ai) {
          const long grow = row0 + 8 * ai;
          double t1 = 0.0;
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int col = c0 + 8 * j + 2 * cq;
            const double v0 = acc[cg][4 * ai + 2 * j] * sa[ai] * sb[j][0];
            const double v1 = acc[cg][4 * ai + 2 * j + 1] * sa[ai] * sb[j][1];
            if (grow < a.M) {
              double* dst = Cz + grow * a.ldc + col;
              if (vecC && col + 1 < a.N) {
                double2* d2 = reinterpret_cast<double2*>(dst);
                if (a.accumulate) { const double2 o = *d2; *d2 = make_double2(o.x + v0, o.y + v1); }
                else *d2 = make_double2(v0, v1);
              } else {
                if (col < a.N) dst[0] = a.accumulate ? dst[0] + v0 : v0;
                if (col + 1 < a.N) dst[1] = a.accumulate ? dst[1] + v1 : v1;
              }
            }
            if (mom) t1 = fma(v1, ax[4 * ai + 2 * j + 1], fma(v0, ax[4 * ai + 2 * j], t1));
          }
          if (mom) {                              // row sum over the 16-column group: the 4 threads that share the row
            t1 += __shfl_xor_sync(0xffffffffu, t1, 1);
            t1 += __shfl_xor_sync(0xffffffffu, t1, 2);
            if (cq == 0 && grow < a.M && c0 < a.N) a.part1[(long)(c0 >> 4) * a.ldpart + grow] = t1;
          }
        }
      }
      if (eprof) { const long long e3 = clock64(); e_wait += e1 - e0; e_drain += e2 - e1; e_store += e3 - e2; }
    }
    if (eprof && lane == 0) { g_oz_prof[3] = (double)e_wait; g_oz_prof[4] = (double)e_drain; g_oz_prof[5] = (double)e_store; }
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
                     int ldk, long pair_stride, int* __restrict__ exps, const double* __restrict__ colvec,
                     double* __restrict__ rowdot) {
  const int lane = threadIdx.x & 31;
  const long warp0 = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long nwarps = ((long)gridDim.x * blockDim.x) >> 5;
  for (long r = warp0; r < rows; r += nwarps) {
    const double* x = X + r * ldx;
    double m = 0.0, dot = 0.0;
    if (colvec) {       // the row is in flight anyway: its product with colvec (lam_m = K bb, utils.py:1090)
      for (int c = lane; c < cols; c += 32) { const double xv = x[c]; m = fmax(m, fabs(xv)); dot = fma(xv, __ldg(colvec + c), dot); }
      dot = warp_sum(dot);
      if (lane == 0) rowdot[r] = dot;
    } else {
      for (int c = lane; c < cols; c += 32) m = fmax(m, fabs(x[c]));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    const int e = (m > 0.0 && m < 1.0e300) ? ilogb(m) + 1 : 0;
    if (lane == 0) exps[r] = e;
    for (int g = lane; g < ldk / 8; g += 32) {
      double t[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int c = 4 * g + j;
        t[j] = c < cols ? scale_pow2(x[c], 6 - e) : 0.0;
      }
      // pair s/2, row r, k-block (4g)/64: [slice 2j: 64 B][slice 2j+1: 64 B]
      const int k = 4 * g;
      int8_t* o = q + r * ldk + (k >> 6) * 128 + (k & 63);
      auto put = [&](int sl, uint32_t w) {
        *reinterpret_cast<uint32_t*>(o + (sl >> 1) * pair_stride + (sl & 1) * 64) = w;
      };
#define SGP_OZ_DIGITS4(sl)                                                                                     \
      if (sl < S) put(sl, take_digit<sl>(t[0]) | (take_digit<sl>(t[1]) << 8) | (take_digit<sl>(t[2]) << 16) |   \
                              (take_digit<sl>(t[3]) << 24));
      SGP_OZ_DIGITS4(0) SGP_OZ_DIGITS4(1) SGP_OZ_DIGITS4(2) SGP_OZ_DIGITS4(3)
      SGP_OZ_DIGITS4(4) SGP_OZ_DIGITS4(5) SGP_OZ_DIGITS4(6) SGP_OZ_DIGITS4(7)
#undef SGP_OZ_DIGITS4
    }
  }
}

// Column-wise |max| of a rows x cols matrix as monotone bit patterns (non-negative doubles order like their bits).
__global__ void __launch_bounds__(256)
oz_col_absmax_kernel(const double* __restrict__ X, long ldx, long rows, int cols, const double* __restrict__ rowsq,
                     unsigned long long* __restrict__ mx) {
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  const int rl = threadIdx.x >> 5;
  __shared__ double red[8][33];
  double m = 0.0;
  if (c < cols)
    for (long r = (long)blockIdx.y * 8 + rl; r < rows; r += (long)gridDim.y * 8)
      m = fmax(m, fabs(X[r * ldx + c]) * (rowsq ? sqrt(fabs(rowsq[r])) : 1.0));
  red[rl][threadIdx.x & 31] = m;
  __syncthreads();
  if (rl == 0 && c < cols) {
#pragma unroll
    for (int i = 1; i < 8; ++i) m = fmax(m, red[i][threadIdx.x]);
    atomicMax(mx + c, (unsigned long long)__double_as_longlong(m));
  }
}

__global__ void oz_exps_from_max_kernel(const unsigned long long* __restrict__ mx, int n, int* __restrict__ exps) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double m = __longlong_as_double((long long)mx[i]);
  exps[i] = (m > 0.0 && m < 1.0e300) ? ilogb(m) + 1 : 0;
}

// Column scales from per-block column maxima (written by the pass that produced the matrix): 32 columns x 8 block
// groups per CTA.  squared: the maxima are of the squared entries; the 1e-10 margin keeps 2^e above the true maximum
// whatever the rounding of the square / square root (a scale one bit too large costs one bit of the last digit).
__global__ void __launch_bounds__(256)
oz_exps_from_colmax_kernel(const double* __restrict__ part, int nblk, int cols, long ld, int squared,
                           int* __restrict__ exps) {
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  const int g = threadIdx.x >> 5;
  __shared__ double red[8][33];
  double m = 0.0;
  if (c < cols)
    for (int b = g; b < nblk; b += 8) m = fmax(m, part[(long)b * ld + c]);
  red[g][threadIdx.x & 31] = m;
  __syncthreads();
  if (g == 0 && c < cols) {
#pragma unroll
    for (int i = 1; i < 8; ++i) m = fmax(m, red[i][threadIdx.x]);
    if (squared) m = sqrt(m) * (1.0 + 1.0e-10);
    exps[c] = (m > 0.0 && m < 1.0e300) ? ilogb(m) + 1 : 0;
  }
}

// Slices of X^T: output row = column c of X (rows x cols), K runs over the rows of X.  One block transposes a
// 64 (rows of X) x 64 (columns) tile through shared memory; thread (quarter, c) turns 16 consecutive k of column c
// into one 16-byte store per slice.
// 4 blocks per SM (64 registers): the kernel is latency-bound between its two barriers per tile; at 78 registers / 3
// blocks the two transposed slicings of a slab took 180 us each, with 4 blocks 165 (stage times 0.71 -> 0.69, 0.31 -> 0.29 ms)
__global__ void __launch_bounds__(256, 4)
oz_slice_cols_kernel(const double* __restrict__ X, long ldx, long rows, int cols, int S, int8_t* __restrict__ q,
                     int ldk, long pair_stride, const int* __restrict__ exps, const double* __restrict__ rowsq) {
  __shared__ double tile[64][65];
  __shared__ double rsc[64];
  // column tiles vary fastest across the grid: the blocks resident together read whole rows of X between them
  const int c0 = blockIdx.x * 64;
  const int c = threadIdx.x & 63, quarter = threadIdx.x >> 6;
  const bool live = c0 + c < cols;
  const int e = live ? exps[c0 + c] : 0;
  for (long rt = blockIdx.y; rt * 64 < rows; rt += gridDim.y) {
    const long r0 = rt * 64;
    if (threadIdx.x < 64) rsc[threadIdx.x] = (rowsq && r0 + threadIdx.x < rows) ? sqrt(fabs(rowsq[r0 + threadIdx.x])) : 1.0;
    __syncthreads();
    for (int el = threadIdx.x; el < 64 * 64; el += 256) {
      const int i = el >> 6, cc = el & 63;
      tile[i][cc] = (r0 + i < rows && c0 + cc < cols) ? X[(r0 + i) * ldx + c0 + cc] * rsc[i] : 0.0;
    }
    __syncthreads();
    if (live) {
      double t[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) t[j] = scale_pow2(tile[16 * quarter + j][c], 6 - e);
      int8_t* orow = q + (long)(c0 + c) * ldk + rt * 128 + quarter * 16;
#define SGP_OZ_DIGITS16(sl)                                                                                    \
      if (sl < S) {                                                                                            \
        uint32_t w[4];                                                                                         \
        _Pragma("unroll") for (int j4 = 0; j4 < 4; ++j4)                                                      \
          w[j4] = take_digit<sl>(t[4 * j4]) | (take_digit<sl>(t[4 * j4 + 1]) << 8) |                           \
                  (take_digit<sl>(t[4 * j4 + 2]) << 16) | (take_digit<sl>(t[4 * j4 + 3]) << 24);               \
        *reinterpret_cast<uint4*>(orow + (sl >> 1) * pair_stride + (sl & 1) * 64) =                            \
            make_uint4(w[0], w[1], w[2], w[3]);                                                                \
      }
      SGP_OZ_DIGITS16(0) SGP_OZ_DIGITS16(1) SGP_OZ_DIGITS16(2) SGP_OZ_DIGITS16(3)
      SGP_OZ_DIGITS16(4) SGP_OZ_DIGITS16(5) SGP_OZ_DIGITS16(6) SGP_OZ_DIGITS16(7)
#undef SGP_OZ_DIGITS16
    }
    __syncthreads();
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
                          cudaStream_t st, const double* colvec, double* rowdot) {
  if (rows <= 0) return cudaSuccess;
  if (S < 1 || S > kOzMaxS || ldk != oz_ldk(cols)) return cudaErrorInvalidValue;
  const long blocks = std::min<long>((rows + 7) / 8, (long)num_sms() * 8);
  note_launch();
  oz_slice_rows_kernel<<<(unsigned)blocks, 256, 0, st>>>(X, ldx, rows, cols, S, q, ldk, rows * (long)ldk, exps,
                                                           rowdot ? colvec : nullptr, rowdot);
  return cudaGetLastError();
}

// Slices of X^T for a contraction over the ROWS of X (rows x cols): q is [pairs][cols][oz_ldk(rows)], exps[cols].
// mx: scratch of `cols` 8-byte words.
cudaError_t oz_slice_cols(const double* X, long ldx, long rows, int cols, int S, int8_t* q, int ldk, int* exps,
                          unsigned long long* mx, cudaStream_t st, const double* rowsq, bool exps_ready) {
  if (rows <= 0 || cols <= 0) return cudaSuccess;
  if (S < 1 || S > kOzMaxS || rows > (1L << 30) || ldk != oz_ldk((int)rows)) return cudaErrorInvalidValue;
  if (!exps_ready) {
    cudaError_t e = cudaMemsetAsync(mx, 0, sizeof(unsigned long long) * cols, st);
    if (e != cudaSuccess) return e;
    const int gy = (int)std::min<long>((rows + 63) / 64, (long)num_sms() * 4 / ((cols + 31) / 32) + 1);
    note_launch(2);
    oz_col_absmax_kernel<<<dim3((cols + 31) / 32, gy), 256, 0, st>>>(X, ldx, rows, cols, rowsq, mx);
    oz_exps_from_max_kernel<<<(cols + 255) / 256, 256, 0, st>>>(mx, cols, exps);
  }
  note_launch();
  oz_slice_cols_kernel<<<dim3((cols + 63) / 64, (unsigned)std::min<long>((rows + 63) / 64, 65535)), 256, 0, st>>>(
      X, ldx, rows, cols, S, q, ldk, (long)cols * ldk, exps, rowsq);
  return cudaGetLastError();
}

void oz_exps_from_colmax(const double* part, int nblk, int cols, long ld, int squared, int* exps, cudaStream_t st) {
  if (cols <= 0) return;
  note_launch();
  oz_exps_from_colmax_kernel<<<(cols + 31) / 32, 256, 0, st>>>(part, nblk, cols, ld, squared, exps);
}

// C (M x N, ld = ldc) = A * B^T from the int8 slices of A (M rows) and B (N rows); K bytes per row, pitch ldk.
static long oz_tiles(long M, int N, int lower_only) {
  const long mt = (M + kOzBM - 1) / kOzBM; const int nt = (N + kOzBN - 1) / kOzBN;
  if (!lower_only) return mt * nt;
  long t = 0;
  for (long i = 0; i < mt; ++i) t += std::min<long>(2 * i + 2, nt);
  return t;
}

int oz_split(long M, int N, int K, int S, int lower_only) {
  const long tiles = oz_tiles(M, N, lower_only);
  const int nkb = (K + kOzBK - 1) / kOzBK;
  const int max_kb = (int)(((1L << 31) - 1) / ((long)S * 4096 * kOzBK));   // int32 exactness of one diagonal
  int nz = (nkb + max_kb - 1) / max_kb;
  // fill the SMs: prefer the split whose item count is closest to whole waves, never chunks under 16 k-blocks
  double best = -1.0; int best_nz = nz;
  for (int c = nz; c <= 64 && (c == nz || nkb / c >= 16); ++c) {
    const double waves = (double)tiles * c / (double)num_sms();
    const double eff = waves / std::ceil(waves) - 0.003 * c;
    if (eff > best) { best = eff; best_nz = c; }
    if (tiles >= 4 * num_sms()) break;   // plenty of tiles already: no split beyond what exactness needs
  }
  return best_nz;
}

double* oz_prof_ptr() {
  double* p = nullptr;
  return cudaGetSymbolAddress((void**)&p, g_oz_prof) == cudaSuccess ? p : nullptr;
}

cudaError_t oz_gemm(const int8_t* qa, const int* ea, long M, const int8_t* qb, const int* eb, int N, int K, int ldk,
                    int S, double* C, long ldc, int nz, long zstride, int accumulate, cudaStream_t st,
                    const OzMoments* mom, int max_ctas, double alpha, int lower_only) {
  if (M <= 0 || N <= 0) return cudaSuccess;
  if (S < 1 || S > kOzMaxS || nz < 1) return cudaErrorInvalidValue;
  const int nkb = (K + kOzBK - 1) / kOzBK;
  const int kb_per_z = (nkb + nz - 1) / nz;
  nz = (nkb + kb_per_z - 1) / kb_per_z;          // no empty slices
  if ((long)kb_per_z * kOzBK * S * 4096 >= (1L << 31)) return cudaErrorInvalidValue;
  CUtensorMap mapA, mapB;
  if (ldk != oz_ldk(K)) return cudaErrorInvalidValue;
  if (!make_slice_map(&mapA, qa, M, ldk, S, kOzBM) || !make_slice_map(&mapB, qb, N, ldk, S, kOzBN))
    return cudaErrorNotSupported;
  static const int dbg = getenv("SGP_OZ_DEBUG") ? atoi(getenv("SGP_OZ_DEBUG")) : 0;
  const long tiles = oz_tiles(M, N, lower_only);
  OzArgs a{ea, eb, C, ldc, (int)M, N, K, S, nz, kb_per_z, zstride, accumulate, alpha, lower_only, (int)tiles,
           mom ? mom->aux : nullptr, mom ? mom->ldaux : 0, mom ? mom->part1 : nullptr, mom ? mom->ldpart : 0, dbg};
  if (mom && (nz != 1 || accumulate)) return cudaErrorInvalidValue;
  const int grid = (int)std::min<long>(tiles * nz, max_ctas > 0 ? max_ctas : num_sms());   // fewer CTAs leave SMs to a concurrent stream
  note_launch();
  oz_gemm_kernel<<<grid, kOzThreads, kOzSmem, st>>>(mapA, mapB, a);
  return cudaGetLastError();
}

}  // namespace sgp
