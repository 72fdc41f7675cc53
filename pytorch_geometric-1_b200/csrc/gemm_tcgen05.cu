// K5 — the dense node-feature transform  C[M,N] = A[M,K] · W[K,N] (+ bias)  on the
// 5th-generation tensor cores (tcgen05.mma, accumulator in TMEM, operands staged by
// TMA into 128-byte-swizzled shared memory).  Replaces `torch.matmul(x, self.weight)`
// of GCNConv.forward / GATConv.forward (cuBLAS SIMT sgemm in the
// reference, because its inputs are fp32).
//
// fp32 fidelity.  kind::tf32 keeps 10 mantissa bits of each operand, which alone
// would miss the fp32 parity bar.  The kernel therefore runs the 3xTF32 split
//     A = Ah + Al,  W = Wh + Wl   (h = round-to-tf32, l = tf32 of the remainder)
//     C = Ah·Wh + Al·Wh + Ah·Wl   (fp32 accumulation in TMEM)
// whose error is ~2^-21 relative per product — below fp32 accumulation noise at
// these K.  Wh/Wl are prepared once per call (tiny); Ah/Al are produced on the fly
// from the TMA-landed tile by four "splitter" warps, in place, before the MMA warp
// consumes the stage.  The layer is HBM-bound (M = 2.4M rows, K,N <= 256), so the two
// extra MMAs hide under the memory time.
//
// Kernels in this file:
//   gemm_ts_kernel<CG, PROF>  the default 3-pass kernel: A_hi / A_lo live in TENSOR memory
//                             (tcgen05.st by the splitter warps, `[a_tmem], b_desc` MMAs), CTA
//                             pairs (cta_group::2, CG = 2) whenever M > 256, TMA-store epilogue
//                             for plain outputs; see the comment block above it
//   gemm_tf32_kernel<...>     operands in shared memory: 1-pass TF32, bf16 (kind::f16), the
//                             PYGB_GEMM_TILE shapes, and PYGB_GEMM_TS=0
//   gemm_xtg_kernel<...>      K5t, dW = X^T G
// All of them issue tcgen05.mma from ONE lane picked by elect.sync inside a warp that walks the
// loop in uniform control flow (see elect_one()).  What follows describes gemm_tf32_kernel.
//
// Persistent CTAs (one per SM) of 10 warps: warp 0 TMA producer, warp 1 TMEM
// allocator + MMA issuer (one elected lane), warps 2-5 splitter, warps 6-9 epilogue
// (TMEM -> registers -> global).  Tile 128 x BN (BN = N rounded to 16, <= 256) x 32
// (one 128-byte swizzle atom of fp32), 2-6 stages.  Two other shapes are compiled and
// bit-identical (tests/test_gemm_gpu.py): 128 x BN x 16 and 256 x BN x 16 (64-byte
// swizzle atoms; the 256-row tile shares every W stage between two 128-row halves).
// On the cfg5 shapes they are not faster (profiles/r01_k5_tiles.md): the kernel is paced
// by the sum of small costs around the MMAs, not by the shared-memory fill.  Barriers per stage: full (TMA
// bytes landed) -> ready (split done, 128 arrivals) -> empty (tcgen05.commit after
// the MMAs read the stage).  Two accumulator buffers in TMEM (acc_full / acc_empty)
// let the epilogue of tile t run under the main loop of tile t+1.
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>

#include <mutex>

#include "common.cuh"

namespace pygb {
namespace {

constexpr int kBM = 128;
constexpr int kBK = 32;  // fp32 elements = 128 bytes = one swizzle atom
constexpr int kGemmThreads = 320;
constexpr int kMaxStages = 6;
constexpr int kATileBytes = kBM * kBK * 4;  // 16 KB
constexpr int kEpiRowFloats = 36;           // 32 columns + 4 pad: conflict-free 128-bit transposes
constexpr int kEpiWarpBytes = 32 * kEpiRowFloats * 4;
constexpr int kEpiBytes = 4 * kEpiWarpBytes;  // staging for the four epilogue warps

// ---- PTX wrappers ---------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
// One lane of the (converged) warp; the compiler keeps code predicated on it on the uniform
// datapath, which `if (lane == 0)` does not (every UTCHMMA then sits in an ELECT / BRA.U.ANY
// loop with R2UR moves: ~60 clocks of issue per MMA on the single issuing thread).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tcgen05_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   bar)
               : "memory");
}
__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                         uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                        uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32"
      " {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32"
      " {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
        "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
        "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ float round_tf32(float x) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
  return __uint_as_float(u);
}

// K-major, 128-byte swizzle, rows of 128 bytes packed: 8-row groups 1024 bytes apart.
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) |
         (2ull << 61);
}

constexpr int kMaxDup = 4;  // destinations per column block of the routed epilogue

// K-major, 64-byte swizzle, rows of 64 bytes packed: 8-row groups 512 bytes apart.
__device__ __forceinline__ uint64_t smem_desc_sw64(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | (32ull << 32) | (1ull << 46) |
         (4ull << 61);
}

struct GemmParams {
  const float* bias;
  int relu;  // max(x, 0) after the bias (NaN stays NaN, like torch.relu)
  void* C;        // fp32, or bf16 for the bf16 kernel
  int64_t ldc;    // elements
  int M, N, K;
  int a_kw;       // columns per plane of A (K when A is one row-major matrix)
  int a_3d;       // A is several planes: 3-D tensor map
  int BN;         // N tile (multiple of 16, <= 256)
  int stages;     // 2..kMaxStages
  int tmem_cols;  // power of two >= acc_bufs * MH * BN
  int acc_bufs;   // accumulator buffers in TMEM (2: the epilogue overlaps the next main loop)
  int split_acc;  // 3xTF32: the correction terms Al.Wh + Ah.Wl accumulate in their own TMEM
                  // columns (added in the epilogue).  The tensor core's fp32 accumulator
                  // truncates (~2^-25.5 relative per accumulation, a bias towards zero that grows
                  // with K); keeping the 2^-11-sized corrections out of the main chain cuts the
                  // number of truncating updates of the large partial sum by 3 (K = 602:
                  // 8.6e-6 -> ~3e-6 of the largest entry against fp64).  Used when TMEM has room
                  // (2 buffers x 2 x MH*BN <= 512 columns).
  int m_tiles, n_tiles;
  int tma_store;             // gemm_ts_kernel: the epilogue stores C with TMA (plain, 16-byte aligned C)
  int dbg;                   // probe builds only (PYGB_GEMM_TS_DBG): wrong numbers, for timing
  int ts_ra, ts_sb, ts_ta;   // gemm_ts_kernel: raw A tiles / W stages in shared memory, A stages in TMEM
  // routed output (row layout -> column layout fused into the epilogue): column block
  // q = col / c_fc of row r goes to c_tab[q * c_dup + d] + (c_row0 + r) * c_ldd + col % c_fc
  // for every d < c_dup (local or peer-GPU bases).  c_tab == nullptr: plain C.
  float* const* c_tab;
  int c_fc, c_dup;
  int64_t c_row0, c_ldd;
};

// Epilogue of one 32-lane quarter of a 128-row accumulator (lane = row): every thread reads 32
// columns of its row from TMEM, the warp transposes the 32x32 block through its `epi` staging
// area and writes full 128-byte lines (4 rows per instruction); bias / ReLU / bf16 conversion /
// routed destinations are applied on the way out.  `tbase` = TMEM address of column 0 of the
// accumulator (lane quarter included), `row0` = global row of lane 0 of this warp.
template <int KIND>
__device__ __forceinline__ void gemm_epilogue_rows(const GemmParams& P, uint32_t tbase, int BN,
                                                   int acc_cols, bool split_acc, uint32_t epi,
                                                   int row0, int n0, int lane, bool vec_ok) {
  float* const Cf = static_cast<float*>(P.C);
  __nv_bfloat16* const Cb = static_cast<__nv_bfloat16*>(P.C);
  const int rr = lane >> 3, jj = lane & 7;
  for (int c0 = 0; c0 < BN; c0 += 32) {
    uint32_t r[32];
    if (c0 + 32 <= BN) {
      tmem_ld32(tbase + (uint32_t)c0, r);
    } else {  // BN is a multiple of 16: a last half chunk
      tmem_ld16(tbase + (uint32_t)c0, r);
#pragma unroll
      for (int j = 16; j < 32; ++j) r[j] = 0u;
    }
    tmem_ld_wait();
    if (split_acc) {  // main + corrections, one round-to-nearest fp32 add
      uint32_t r2[32];
      if (c0 + 32 <= BN) {
        tmem_ld32(tbase + (uint32_t)(acc_cols + c0), r2);
      } else {
        tmem_ld16(tbase + (uint32_t)(acc_cols + c0), r2);
#pragma unroll
        for (int j = 16; j < 32; ++j) r2[j] = 0u;
      }
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j)
        r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(r2[j]));
    }
#pragma unroll
    for (int j = 0; j < 32; j += 4)
      asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(
                       epi + (uint32_t)((lane * kEpiRowFloats + j) * 4)),
                   "r"(r[j]), "r"(r[j + 1]), "r"(r[j + 2]), "r"(r[j + 3])
                   : "memory");
    __syncwarp();
    const int col = n0 + c0 + jj * 4;
    float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (P.bias != nullptr) {
      if (col < P.N) bv.x = __ldg(P.bias + col);
      if (col + 1 < P.N) bv.y = __ldg(P.bias + col + 1);
      if (col + 2 < P.N) bv.z = __ldg(P.bias + col + 2);
      if (col + 3 < P.N) bv.w = __ldg(P.bias + col + 3);
    }
    const bool col_ok = col < P.N;
    if (P.c_tab == nullptr) {       // plain C: the hot path stays branch-free inside the row loop
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int rl = i * 4 + rr;
        float4 v;
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                     : "r"(epi + (uint32_t)((rl * kEpiRowFloats + jj * 4) * 4)));
        v.x += bv.x; v.y += bv.y; v.z += bv.z; v.w += bv.w;
        if (P.relu) {
          v.x = v.x < 0.f ? 0.f : v.x; v.y = v.y < 0.f ? 0.f : v.y;
          v.z = v.z < 0.f ? 0.f : v.z; v.w = v.w < 0.f ? 0.f : v.w;
        }
        const int row = row0 + rl;
        if (row < P.M && col_ok && !(P.dbg & 4)) {
          if (KIND == 1) {
            __nv_bfloat16* dst = Cb + (int64_t)row * P.ldc + col;
            if (vec_ok && col + 4 <= P.N) {     // 4 bf16 = 8 bytes, 8-byte aligned
              *reinterpret_cast<uint2*>(dst) = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
            } else {
              dst[0] = __float2bfloat16_rn(v.x);
              if (col + 1 < P.N) dst[1] = __float2bfloat16_rn(v.y);
              if (col + 2 < P.N) dst[2] = __float2bfloat16_rn(v.z);
              if (col + 3 < P.N) dst[3] = __float2bfloat16_rn(v.w);
            }
          } else {
            float* dst = Cf + (int64_t)row * P.ldc + col;
            if (vec_ok && col + 4 <= P.N) {
              *reinterpret_cast<float4*>(dst) = v;
            } else {
              dst[0] = v.x;
              if (col + 1 < P.N) dst[1] = v.y;
              if (col + 2 < P.N) dst[2] = v.z;
              if (col + 3 < P.N) dst[3] = v.w;
            }
          }
        }
      }
    } else {                        // routed: N and c_fc are multiples of 4 -> whole vectors
      const int qb = col_ok ? col / P.c_fc : 0;
      const int64_t off = P.c_row0 * P.c_ldd + (col - qb * P.c_fc);
      for (int d = 0; d < P.c_dup; ++d) {
        float* rb = col_ok ? P.c_tab[qb * P.c_dup + d] + off : nullptr;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int rl = i * 4 + rr;
          float4 v;
          asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                       : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                       : "r"(epi + (uint32_t)((rl * kEpiRowFloats + jj * 4) * 4)));
          v.x += bv.x; v.y += bv.y; v.z += bv.z; v.w += bv.w;
          if (P.relu) {
            v.x = v.x < 0.f ? 0.f : v.x; v.y = v.y < 0.f ? 0.f : v.y;
            v.z = v.z < 0.f ? 0.f : v.z; v.w = v.w < 0.f ? 0.f : v.w;
          }
          const int row = row0 + rl;
          if (row < P.M && col_ok)
            *reinterpret_cast<float4*>(rb + (int64_t)row * P.c_ldd) = v;
        }
      }
    }
    __syncwarp();
  }
}

// PASSES = 3: 3xTF32 split (fp32-faithful);  PASSES = 1: plain TF32.
// Persistent: CTA b works on tiles b, b+grid, ...  (n fastest, so the A tile of a
// row block is re-read from L2 when N spans several tiles).
// MH = 1: tile 128 x BN x 32 (128-byte swizzle atoms).  MH = 2: tile 256 x BN x 16 (64-byte
// swizzle atoms): the two 128-row halves share every W tile, so a k-step of 32 moves
// 2*16 + 2*BN*64*2/1024 KB for 256 rows instead of 16 + 2*BN*128/1024 KB for 128 — 40 % fewer
// bytes from L2 per flop at BN = 256, where the shared-memory fill rate bounds the main loop.
// KIND = 0: fp32 operands, kind::tf32 (1 or 3 passes).  KIND = 1: bf16 operands and bf16
// output, kind::f16 with fp32 accumulation in TMEM — one pass, no splitter work; BKT counts
// ELEMENTS per k-step (64 bf16 = one 128-byte swizzle row, 32 bf16 = one 64-byte row).
template <int PASSES, int MH, int BKT, int KIND = 0>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tf32_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapBh,
                 const __grid_constant__ CUtensorMap mapBl, const GemmParams P) {
  constexpr int kBK = BKT;                        // shadows the namespace constants
  constexpr int kEsz = KIND == 1 ? 2 : 4;         // operand element size
  constexpr int kTileM = MH * kBM;
  constexpr int kATileBytes = kTileM * kBK * kEsz;
  constexpr bool SW64 = BKT * kEsz == 64;         // 64-byte rows -> 64-byte swizzle atoms
  static_assert(BKT * kEsz == 64 || BKT * kEsz == 128, "k-step of one 64- or 128-byte swizzle row");
  static_assert(KIND == 0 || PASSES == 1, "the bf16 kernel is single-pass");
  extern __shared__ __align__(1024) unsigned char gsm[];
  __shared__ __align__(8) uint64_t bars[3 * kMaxStages + 4];
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int BN = P.BN, S = P.stages;
  const uint32_t b_bytes = (uint32_t)BN * kBK * kEsz;
  const uint32_t stage_bytes = (PASSES == 3 ? 2 : 1) * (kATileBytes + b_bytes);
  const uint32_t base = (smem_u32(gsm) + 1023u) & ~1023u;  // swizzle atoms: 1024-byte aligned
  auto a_hi = [&](int s) { return base + (uint32_t)s * stage_bytes; };
  auto a_lo = [&](int s) { return a_hi(s) + kATileBytes; };
  auto b_hi = [&](int s) { return a_hi(s) + (PASSES == 3 ? 2 : 1) * kATileBytes; };
  auto b_lo = [&](int s) { return b_hi(s) + b_bytes; };
  const uint32_t bar0 = smem_u32(bars);
  auto full = [&](int s) { return bar0 + 8u * s; };
  auto ready = [&](int s) { return bar0 + 8u * (kMaxStages + s); };
  auto empty = [&](int s) { return bar0 + 8u * (2 * kMaxStages + s); };
  auto acc_full = [&](int b) { return bar0 + 8u * (3 * kMaxStages + b); };
  auto acc_empty = [&](int b) { return bar0 + 8u * (3 * kMaxStages + 2 + b); };

  const int nk = (P.K + kBK - 1) / kBK;
  const int n_tiles_total = P.m_tiles * P.n_tiles;
  const int acc_cols = MH * BN;                   // TMEM columns of one accumulator
  const int acc_stride = acc_cols * (P.split_acc ? 2 : 1);   // columns of one buffer
  const bool split_acc = PASSES == 3 && P.split_acc != 0;
  const int nbuf = P.acc_bufs;                    // 2, or 1 when 2 * acc_stride > 512

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(full(s), 1);
      mbar_init(ready(s), 128);
      mbar_init(empty(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(acc_full(b), 1);
      mbar_init(acc_empty(b), 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // TMEM allocation (whole warp), address lands in shared memory
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(&tmem_base_s)),
                 "r"((uint32_t)P.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_d = tmem_base_s;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mapBh) : "memory");
      if (PASSES == 3) asm volatile("prefetch.tensormap [%0];" ::"l"(&mapBl) : "memory");
      int it = 0;
      for (int tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x) {
        const int m0 = (tile / P.n_tiles) * kTileM, n0 = (tile % P.n_tiles) * BN;
        for (int kt = 0; kt < nk; ++kt, ++it) {
          const int s = it % S;
          const uint32_t ph = (uint32_t)(it / S) & 1u;
          mbar_wait(empty(s), ph ^ 1u);
          mbar_expect_tx(full(s), kATileBytes + (PASSES == 3 ? 2 : 1) * b_bytes);
          {  // A is a_slices planes of [M, a_kw] columns (one plane: ordinary row-major)
            const int kcol = kt * kBK;
            if (P.a_3d) {
              const int sl = kcol / P.a_kw;
              tma_load_3d(a_hi(s), &mapA, full(s), kcol - sl * P.a_kw, m0, sl);
            } else {
              tma_load_2d(a_hi(s), &mapA, full(s), kcol, m0);
            }
          }
          tma_load_2d(b_hi(s), &mapBh, full(s), kt * kBK, n0);
          if (PASSES == 3) tma_load_2d(b_lo(s), &mapBl, full(s), kt * kBK, n0);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the whole warp walks the loop, one elected lane issues =====
    {
      // D fp32; A/B format 2 = tf32, 1 = bf16; both K-major
      constexpr uint32_t kFmt = KIND == 1 ? 1u : 2u;
      const uint32_t idesc = (1u << 4) | (kFmt << 7) | (kFmt << 10) | ((uint32_t)(BN >> 3) << 17) |
                             ((uint32_t)(kBM >> 4) << 24);
      int it = 0, tl = 0;
      for (int tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x, ++tl) {
        const int buf = tl % nbuf;
        mbar_wait(acc_empty(buf), ((uint32_t)(tl / nbuf) & 1u) ^ 1u);  // epilogue drained this buffer
        tcgen05_fence_after();
        const uint32_t dcol = tmem_d + (uint32_t)(buf * acc_stride);
        for (int kt = 0; kt < nk; ++kt, ++it) {
          const int s = it % S;
          const uint32_t ph = (uint32_t)(it / S) & 1u;
          mbar_wait(PASSES == 3 ? ready(s) : full(s), ph);
          tcgen05_fence_after();
          const uint64_t dbh = SW64 ? smem_desc_sw64(b_hi(s)) : smem_desc_sw128(b_hi(s));
          const uint64_t dbl = SW64 ? smem_desc_sw64(b_lo(s)) : smem_desc_sw128(b_lo(s));
          if (elect_one()) {
#pragma unroll
          for (int h = 0; h < MH; ++h) {       // 128-row halves of the tile share the W stage
            const uint32_t hoff = (uint32_t)(h * kBM * kBK * kEsz);
            const uint64_t dah = SW64 ? smem_desc_sw64(a_hi(s) + hoff) : smem_desc_sw128(a_hi(s) + hoff);
            const uint64_t dal = SW64 ? smem_desc_sw64(a_lo(s) + hoff) : smem_desc_sw128(a_lo(s) + hoff);
            const uint32_t dc = dcol + (uint32_t)(h * BN);
#pragma unroll
            for (int k = 0; k < kBK * kEsz / 32; ++k) {  // UMMA_K = 8 tf32 / 16 bf16 = 32 bytes -> +2 in the address field
              const uint32_t acc = (kt > 0 || k > 0) ? 1u : 0u;
              if (KIND == 1) mma_f16(dc, dah + 2u * k, dbh + 2u * k, idesc, acc);
              else mma_tf32(dc, dah + 2u * k, dbh + 2u * k, idesc, acc);
              if (PASSES == 3) {
                const uint32_t dcc = split_acc ? dc + (uint32_t)acc_cols : dc;
                mma_tf32(dcc, dal + 2u * k, dbh + 2u * k, idesc, split_acc ? acc : 1u);
                mma_tf32(dcc, dah + 2u * k, dbl + 2u * k, idesc, 1u);
              }
            }
          }
          tcgen05_commit(empty(s));  // frees the stage once the MMAs have read it
          }
          __syncwarp();
        }
        if (elect_one()) tcgen05_commit(acc_full(buf));  // accumulator of this tile complete
        __syncwarp();
      }
    }
  } else if (warp < 6) {
    // ===== splitter: warps 2..5 (only for the 3xTF32 split) =====
    if (PASSES == 3) {
      const int t = threadIdx.x - 64;  // 0..127
      int it = 0;
      for (int tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x) {
        for (int kt = 0; kt < nk; ++kt, ++it) {
          const int s = it % S;
          const uint32_t ph = (uint32_t)(it / S) & 1u;
          mbar_wait(full(s), ph);
          const uint32_t ah = a_hi(s), al = a_lo(s);
#pragma unroll
          for (int i = 0; i < kATileBytes / 16 / 128; ++i) {
            const uint32_t off = (uint32_t)(t + i * 128) * 16u;
            float4 v;
            asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                         : "r"(ah + off));
            float4 h, l;
            h.x = round_tf32(v.x); h.y = round_tf32(v.y);
            h.z = round_tf32(v.z); h.w = round_tf32(v.w);
            l.x = round_tf32(v.x - h.x); l.y = round_tf32(v.y - h.y);
            l.z = round_tf32(v.z - h.z); l.w = round_tf32(v.w - h.w);
            asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(ah + off), "f"(h.x),
                         "f"(h.y), "f"(h.z), "f"(h.w)
                         : "memory");
            asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(al + off), "f"(l.x),
                         "f"(l.y), "f"(l.z), "f"(l.w)
                         : "memory");
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // visible to the tensor core
          mbar_arrive(ready(s));
        }
      }
    }
  } else {
    // ===== epilogue: warps 6..9.  TMEM lane = accumulator row: each thread reads 32 columns of
    // its row, the warp transposes the 32x32 block through shared memory and writes full
    // 128-byte lines (4 rows per instruction) =====
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const uint32_t epi = base + (uint32_t)S * stage_bytes + (uint32_t)(q * kEpiWarpBytes);
    const bool vec_ok = (P.ldc % 4 == 0) && ((reinterpret_cast<uintptr_t>(P.C) & 15) == 0);
    int tl = 0;
    for (int tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x, ++tl) {
      const int m0 = (tile / P.n_tiles) * kTileM, n0 = (tile % P.n_tiles) * BN;
      const int buf = tl % nbuf;
      mbar_wait(acc_full(buf), (uint32_t)(tl / nbuf) & 1u);
      tcgen05_fence_after();
      for (int h = 0; h < MH; ++h) {
      const int row0 = m0 + h * kBM + q * 32;
      const uint32_t tbase = tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * acc_stride + h * BN);
        gemm_epilogue_rows<KIND>(P, tbase, BN, acc_cols, split_acc, epi, row0, n0, lane, vec_ok);
      }
      tcgen05_fence_before();
      mbar_arrive(acc_empty(buf));  // 128 arrivals: the MMA warp may overwrite this buffer
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d),
                 "r"((uint32_t)P.tmem_cols)
                 : "memory");
  }
}

// ---- 3xTF32 with the A operand in TMEM ("TS" form of tcgen05.mma) -------------------------
// The shared-memory port of an SM moves 128 bytes per clock, and in the kernel above every
// k-step of 32 sends through it: the TMA fill (A + Wh + Wl), the splitter's read of A and its
// writes of Ah and Al, and the operand reads of 12 MMAs (Ah twice, Al once, Wh twice, Wl once)
// — 272 KB at BN = 256, i.e. ~2,100 clocks for 1,536 clocks of tensor work: the port, not HBM
// and not the tensor pipe, paces that kernel (profiles/r02_k5_ts.md).  Here the splitter warps
// read the raw fp32 tile once and write Ah / Al straight into TENSOR memory (tcgen05.st, lane =
// row, 32 + 32 columns per stage), and the MMAs take A from there: per k-step only the fill and
// the W reads touch shared memory (112 KB at BN = 128).  TMEM holds two accumulator buffers plus
// the A ring (64 columns per stage), so BN <= 128: wider N runs as several n-tiles of one
// 128-row block back to back on the same CTA (A comes from HBM once, then from L2).
// Same products, same accumulation order per output element as gemm_tf32_kernel<3, 1, 32>.
__device__ __forceinline__ void mma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc,
                                            uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0],"
      " {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
      "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
      "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]),
      "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]),
      "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() {
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

constexpr int kTsACols = 2 * kBK;   // TMEM columns of one A stage: Ah (32) + Al (32)
constexpr int kTsThreads = 480;     // + warp 10: the W producer, warps 11..14: second splitter group
constexpr int kTsMaxRA = 12;        // raw A tiles in flight (16 KB each)
constexpr int kTsMaxSB = 6;         // W stages (Wh + Wl tiles of BN x 32)

// Probe (PYGB_GEMM_TS_DBG & 16): clocks CTA 0 spends in each wait, per role.
static __device__ unsigned long long g_ts_prof[16];
static __device__ unsigned long long g_ts_cta[2 * 160];   // per CTA: clocks, ns
__device__ __forceinline__ void mbar_wait_prof(uint32_t bar, uint32_t parity, bool prof,
                                               unsigned long long& acc) {
  if (!prof) {
    mbar_wait(bar, parity);
    return;
  }
  const long long t0 = clock64();
  mbar_wait(bar, parity);
  acc += (unsigned long long)(clock64() - t0);
}

// Three rings, because the three streams have different latencies: raw A tiles come from HBM
// (~3 us under load: an SM needs >= 100 KB of loads in flight to stream at its share of the HBM
// rate, and the raw tile is all that has to wait that long), W tiles come from L2, and the split
// A lives in TMEM for the short splitter -> MMA hop.
//   rawA[RA]  : warp 0 (TMA) -> a_full  -> splitter (reads into registers) -> a_empty
//   W[SB]     : warp 10 (TMA) -> b_full -> MMA ........ tcgen05.commit -> b_empty
//   tmemA[SB] : splitter (tcgen05.st) -> ready -> MMA .. (the same commit: both rings have depth
//               SB and move in lock step, so b_empty has two waiters: warp 10 and the splitter)
// ---- epilogue with TMA stores --------------------------------------------------------------
// The transposing epilogue above spends ~1,200 clocks per 32x32 block and warp on its dependent
// ld.shared -> bias -> st.global chain (measured without the stores: gemm_ts_kernel was waiting for
// its epilogue 98 % of the time at N = 256).  Here a thread keeps its row: bias / ReLU on the 32
// registers, eight 16-byte st.shared into a 128-byte-swizzled 32x32 block, and one lane hands the
// block to the TMA engine (cp.async.bulk.tensor store; rows >= M and columns >= N are clipped by
// the tensor map).  Two staging blocks per warp: the store of one block overlaps the next.
constexpr int kEpiTmaWarpBytes = 2 * 4096;
constexpr int kEpiTmaBytes = 4 * kEpiTmaWarpBytes;

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map),
               "r"(src), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {   // <= N groups still reading shared memory
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tma_store_wait_all() {
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// `stage` = this warp's two 4 KB staging blocks (1024-byte aligned); `blk` counts the blocks this
// warp has stored so far (selects the staging block; carried across tiles).
struct EpiProf {   // probe: clocks in tmem wait + math, staging wait, st.shared + fence, store issue
  bool on;
  unsigned long long a, b, c, d;
};

// one 32-row x 32-column block held in r[] (thread = row): bias, ReLU, staging, TMA store
__device__ __forceinline__ void epi_store_block(const GemmParams& P, const CUtensorMap* mapC,
                                                uint32_t (&r)[32], uint32_t stage, int row0, int col0,
                                                int lane, int& blk, EpiProf& pf, long long t0) {
  if (P.bias != nullptr) {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float b = col0 + j < P.N ? __ldg(P.bias + col0 + j) : 0.f;   // same address in every lane
      r[j] = __float_as_uint(__uint_as_float(r[j]) + b);
    }
  }
  if (P.relu) {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float v = __uint_as_float(r[j]);
      r[j] = __float_as_uint(v < 0.f ? 0.f : v);
    }
  }
  const uint32_t dst = stage + (uint32_t)((blk & 1) * 4096);
  const uint32_t sw = (uint32_t)(lane & 7);
  const long long t1 = pf.on ? clock64() : 0;
  // the store that read this staging block two blocks ago has finished reading it (elect.sync
  // picks the same lane of a full warp every time: the lane that issued that store)
  if (elect_one()) tma_store_wait_read<1>();
  __syncwarp();
  const long long t2 = pf.on ? clock64() : 0;
#pragma unroll
  for (int j = 0; j < 8; ++j)
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(
                     dst + (uint32_t)lane * 128u + (((uint32_t)j ^ sw) << 4)),
                 "r"(r[4 * j]), "r"(r[4 * j + 1]), "r"(r[4 * j + 2]), "r"(r[4 * j + 3])
                 : "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncwarp();
  const long long t3 = pf.on ? clock64() : 0;
  if (elect_one()) {
    if (!(P.dbg & 4)) {
      tma_store_2d(mapC, dst, col0, row0);
      tma_store_commit();
    }
  }
  ++blk;
  if (pf.on) {
    pf.a += (unsigned long long)(t1 - t0);
    pf.b += (unsigned long long)(t2 - t1);
    pf.c += (unsigned long long)(t3 - t2);
    pf.d += (unsigned long long)(clock64() - t3);
  }
}

__device__ __forceinline__ void epi_tmem_load(uint32_t taddr, bool full, uint32_t (&r)[32]) {
  if (full) tmem_ld32(taddr, r);
  else tmem_ld16(taddr, r);   // BN is a multiple of 16: a last half block
}

__device__ __forceinline__ void gemm_epilogue_rows_tma(const GemmParams& P, const CUtensorMap* mapC,
                                                       uint32_t tbase, int BN, int acc_cols,
                                                       bool split_acc, uint32_t stage, int row0,
                                                       int n0, int lane, int& blk, EpiProf& pf) {
  const int c_end = BN < P.N - n0 ? BN : P.N - n0;   // columns of this tile that exist
  // (the TMEM load is not software-pipelined across blocks: registers named by an in-flight
  // tcgen05.ld must not be touched before tcgen05.wait::ld, and under register pressure the
  // compiler spilled them right after the load)
  for (int c0 = 0; c0 < c_end; c0 += 32) {
    const long long t0 = pf.on ? clock64() : 0;
    const bool full = c0 + 32 <= BN;
    uint32_t r[32];
    epi_tmem_load(tbase + (uint32_t)c0, full, r);
    tmem_ld_wait();
    if (split_acc) {   // main + corrections, one round-to-nearest fp32 add
      uint32_t r2[32];
      epi_tmem_load(tbase + (uint32_t)(acc_cols + c0), full, r2);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j)
        r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(r2[j]));
    }
    if (!full) {
#pragma unroll
      for (int j = 16; j < 32; ++j) r[j] = 0u;
    }
    epi_store_block(P, mapC, r, stage, row0, n0 + c0, lane, blk, pf, t0);
  }
}

// ---- CTA pairs (cta_group::2) ------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::
                   : "memory");
}
// shared::cluster address of the same shared-memory location in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load whose bytes are counted on a barrier of the pair's leader CTA (`bar` = its
// shared::cluster address); the tile lands in this CTA's shared memory
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                                 int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
// completion of the pair's MMAs arrives on the barrier at this offset in BOTH CTAs
__device__ __forceinline__ void tcgen05_commit_pair(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64"
      " [%0], %1;" ::"r"(bar),
      "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void mma_tf32_ts_pair(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc,
                                                 uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// CG = 2: two CTAs of a cluster (the two SMs of a TPC) work on 256 rows as ONE tensor-core
// operation (cta_group::2, M = 256): each CTA loads, splits and keeps in TMEM its own 128 rows
// of A and drains its own 128 accumulator rows, but holds only HALF of every W tile (BN/2 rows
// of Wh and of Wl) — the tensor cores of both SMs read both halves.  That halves the W fill and
// the W operand reads per CTA, which is what the shared-memory port was short of
// (profiles/r02_k5_ts.md).  The leader (cluster rank 0) issues every MMA; its b_full barrier
// counts the TMA bytes of both CTAs, its `ready` / `acc_empty` barriers take the arrivals of both
// CTAs' splitter / epilogue warps, and its commits arrive on b_empty / acc_full of both CTAs.
template <int CG, bool PROF>   // PROF: the PYGB_GEMM_TS_DBG & 16 probe build (extra live registers)
__global__ void __launch_bounds__(kTsThreads, 1)
gemm_ts_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapBh,
               const __grid_constant__ CUtensorMap mapBl, const __grid_constant__ CUtensorMap mapC,
               const GemmParams P) {
  extern __shared__ __align__(1024) unsigned char gsm[];
  __shared__ __align__(8) uint64_t bars[2 * kTsMaxRA + 3 * kTsMaxSB + 4];
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int BN = P.BN, RA = P.ts_ra, SB = P.ts_sb, TA = P.ts_ta;
  const uint32_t b_bytes = (uint32_t)(BN / CG) * kBK * 4;   // this CTA's part of a W tile
  const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;
  const bool leader = rank == 0;
  const int unit0 = (int)blockIdx.x / CG, ustride = (int)gridDim.x / CG;   // units of CG*128 rows
  const uint32_t base = (smem_u32(gsm) + 1023u) & ~1023u;
  const uint32_t b_base = base + (uint32_t)RA * kATileBytes;
  auto a_raw = [&](int s) { return base + (uint32_t)s * kATileBytes; };
  auto b_hi = [&](int s) { return b_base + (uint32_t)s * 2u * b_bytes; };
  auto b_lo = [&](int s) { return b_hi(s) + b_bytes; };
  const uint32_t bar0 = smem_u32(bars);
  auto a_full = [&](int s) { return bar0 + 8u * s; };
  auto a_empty = [&](int s) { return bar0 + 8u * (kTsMaxRA + s); };
  auto b_full = [&](int s) { return bar0 + 8u * (2 * kTsMaxRA + s); };
  auto b_empty = [&](int s) { return bar0 + 8u * (2 * kTsMaxRA + kTsMaxSB + s); };
  auto ready = [&](int s) { return bar0 + 8u * (2 * kTsMaxRA + 2 * kTsMaxSB + s); };
  auto acc_full = [&](int b) { return bar0 + 8u * (2 * kTsMaxRA + 3 * kTsMaxSB + b); };
  auto acc_empty = [&](int b) {
    return bar0 + 8u * (2 * kTsMaxRA + 3 * kTsMaxSB + 2 + b);
  };

  // ring position + phase, advanced without divisions (these loops run on single threads)
  struct Ring {
    int s;
    uint32_t ph;
    __device__ __forceinline__ void next(int depth) {
      if (++s == depth) {
        s = 0;
        ph ^= 1u;
      }
    }
  };
  const int nk = (P.K + kBK - 1) / kBK;
  const bool split_acc = P.split_acc != 0;
  const int acc_stride = BN * (split_acc ? 2 : 1);
  const uint32_t a_col0 = (uint32_t)(2 * acc_stride);   // the A ring follows the two accumulators

  if (threadIdx.x == 0) {
    for (int s = 0; s < RA; ++s) {
      mbar_init(a_full(s), 1);
      mbar_init(a_empty(s), 4);    // one arrival per splitter warp
    }
    for (int s = 0; s < SB; ++s) {
      mbar_init(b_full(s), 1);
      mbar_init(b_empty(s), 1);
    }
    for (int s = 0; s < TA; ++s) {
      mbar_init(ready(s), 4 * CG);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(acc_full(b), 1);
      mbar_init(acc_empty(b), 4 * CG);  // one per epilogue warp (of both CTAs)
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {   // the same warp of both CTAs of a pair
    if (CG == 2) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                       smem_u32(&tmem_base_s)),
                   "r"((uint32_t)P.tmem_cols)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                       smem_u32(&tmem_base_s)),
                   "r"((uint32_t)P.tmem_cols)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  tcgen05_fence_before();
  if (CG == 2) cluster_sync_all();   // the peer's barriers are initialised before anything remote
  else __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_d = tmem_base_s;
  // every role walks the same sequence of k-steps: 128-row blocks b, b + grid, ...; all n-tiles
  // of a block back to back (the block's A tiles come from HBM once, then from L2)
  const int my_blocks = P.m_tiles > unit0 ? (P.m_tiles - unit0 + ustride - 1) / ustride : 0;
  const int n_tiles_cta = my_blocks * P.n_tiles;
  const bool prof = PROF && (P.dbg & 16) && blockIdx.x == 0;
  unsigned long long w0 = 0, w1 = 0, w2 = 0;
  const long long t_begin = clock64();
  unsigned long long t_begin_ns = 0;
  if (PROF && (P.dbg & 16)) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_begin_ns));

  if (warp == 0) {
    // ===== raw A producer =====
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
      Ring ra{0, 0u};
      for (int mb = unit0; mb < P.m_tiles; mb += ustride) {
        const int m0 = (mb * CG + (int)rank) * kBM;
        for (int kk = 0, kt = 0; kk < P.n_tiles * nk; ++kk, kt = kt + 1 == nk ? 0 : kt + 1, ra.next(RA)) {
          const int s = ra.s;
          mbar_wait_prof(a_empty(s), ra.ph ^ 1u, prof, w0);
          mbar_expect_tx(a_full(s), kATileBytes);
          const int kcol = kt * kBK;
          if (P.a_3d) {
            const int sl = kcol / P.a_kw;
            tma_load_3d(a_raw(s), &mapA, a_full(s), kcol - sl * P.a_kw, m0, sl);
          } else {
            tma_load_2d(a_raw(s), &mapA, a_full(s), kcol, m0);
          }
        }
      }
      if (prof) { g_ts_prof[0] = w0; g_ts_prof[1] = (unsigned long long)(clock64() - t_begin); }
    }
  } else if (warp == 10) {
    // ===== W producer =====
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mapBh) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mapBl) : "memory");
      Ring rb{0, 0u};
      for (int tl = 0, nt = 0; tl < n_tiles_cta; ++tl, nt = nt + 1 == P.n_tiles ? 0 : nt + 1) {
        const int n0 = nt * BN + (int)rank * (BN / CG);   // this CTA's rows of the W tile
        for (int kt = 0; kt < nk; ++kt, rb.next(SB)) {
          const int s = rb.s;
          mbar_wait_prof(b_empty(s), rb.ph ^ 1u, prof, w0);
          if (CG == 2) {
            const uint32_t lbar = mapa_u32(b_full(s), 0u);   // the leader's barrier counts both halves
            if (leader) mbar_expect_tx(b_full(s), 2 * CG * b_bytes);
            tma_load_2d_pair(b_hi(s), &mapBh, lbar, kt * kBK, n0);
            tma_load_2d_pair(b_lo(s), &mapBl, lbar, kt * kBK, n0);
          } else {
            mbar_expect_tx(b_full(s), 2 * b_bytes);
            tma_load_2d(b_hi(s), &mapBh, b_full(s), kt * kBK, n0);
            tma_load_2d(b_lo(s), &mapBl, b_full(s), kt * kBK, n0);
          }
        }
      }
      if (prof) { g_ts_prof[2] = w0; g_ts_prof[3] = (unsigned long long)(clock64() - t_begin); }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: A from TMEM, W from shared memory.  The whole warp walks the loop
    // (uniform control flow); one elected lane issues.  CTA pairs: the leader only =====
    if (leader) {
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) |
                             ((uint32_t)((CG * kBM) >> 4) << 24);
      Ring rb{0, 0u};
      for (int tl = 0; tl < n_tiles_cta; ++tl) {
        const int buf = tl & 1;
        mbar_wait_prof(acc_empty(buf), ((uint32_t)(tl >> 1) & 1u) ^ 1u, prof, w0);
        tcgen05_fence_after();
        const uint32_t dc = tmem_d + (uint32_t)(buf * acc_stride);
        const uint32_t dcc = split_acc ? dc + (uint32_t)BN : dc;
        for (int kt = 0; kt < nk; ++kt, rb.next(SB)) {
          const int sb = rb.s, ta = rb.s;
          mbar_wait_prof(b_full(sb), rb.ph, prof, w1);
          mbar_wait_prof(ready(ta), rb.ph, prof, w2);
          tcgen05_fence_after();
          const uint64_t dbh = smem_desc_sw128(b_hi(sb));
          const uint64_t dbl = smem_desc_sw128(b_lo(sb));
          const uint32_t ah = tmem_d + a_col0 + (uint32_t)(ta * kTsACols);
          const uint32_t al = ah + (uint32_t)kBK;
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < kBK / 8; ++k) {   // UMMA_K = 8 tf32: 8 TMEM columns / 32 bytes of W
              const uint32_t acc = (kt > 0 || k > 0) ? 1u : 0u;
              if (CG == 2) {
                mma_tf32_ts_pair(dc, ah + 8u * k, dbh + 2u * k, idesc, acc);
                mma_tf32_ts_pair(dcc, al + 8u * k, dbh + 2u * k, idesc, split_acc ? acc : 1u);
                mma_tf32_ts_pair(dcc, ah + 8u * k, dbl + 2u * k, idesc, 1u);
              } else {
                mma_tf32_ts(dc, ah + 8u * k, dbh + 2u * k, idesc, acc);
                if (P.dbg & 1) continue;
                mma_tf32_ts(dcc, al + 8u * k, dbh + 2u * k, idesc, split_acc ? acc : 1u);
                mma_tf32_ts(dcc, ah + 8u * k, dbl + 2u * k, idesc, 1u);
              }
            }
            // W stage and TMEM A stage (same ring position) are free — in both CTAs of a pair
            if (CG == 2) tcgen05_commit_pair(b_empty(sb));
            else tcgen05_commit(b_empty(sb));
          }
          __syncwarp();
        }
        if (elect_one()) {
          if (CG == 2) tcgen05_commit_pair(acc_full(buf));
          else tcgen05_commit(acc_full(buf));
        }
        __syncwarp();
      }
      if (prof && lane == 0) {
        g_ts_prof[4] = w0; g_ts_prof[5] = w1; g_ts_prof[6] = w2;
        g_ts_prof[7] = (unsigned long long)(clock64() - t_begin);
      }
    }
  } else if (warp < 6 || warp >= 11) {
    // ===== splitter: two groups of four warps (2..5 and 11..14) take alternate k-steps — one
    // group's chain wait -> ld.shared -> cvt -> tcgen05.st -> wait::st -> arrive is ~1,000 clocks
    // per step, twice what the MMAs of a CTA pair need.  Thread = tile row = TMEM lane
    // (warp % 4 = lane quarter) =====
    const int grp = warp < 6 ? 0 : 1;
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t ta0 = tmem_d + ((uint32_t)(q * 32) << 16) + a_col0;
    const uint32_t sw = (uint32_t)(row & 7);
    const int n_steps = n_tiles_cta * nk;
    Ring ra{0, 0u}, rt{0, 0u};
    if (grp) {
      ra.next(RA);
      rt.next(TA);
    }
    for (int it = grp; it < n_steps; it += 2, ra.next(RA), ra.next(RA), rt.next(TA), rt.next(TA)) {
      const int sa = ra.s, ta = rt.s;
      mbar_wait_prof(a_full(sa), ra.ph, prof, w0);
      const uint32_t src = a_raw(sa) + (uint32_t)row * 128u;
      uint32_t hi[32], lo[32];
      if (P.dbg & 2) {
#pragma unroll
        for (int c = 0; c < 32; ++c) hi[c] = lo[c] = 0u;
      } else
#pragma unroll
      for (int c = 0; c < 8; ++c) {   // 16-byte chunk c of the row sits at chunk c ^ (row % 8)
        float4 v;
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                     : "r"(src + (((uint32_t)c ^ sw) << 4)));
        const float hx = round_tf32(v.x), hy = round_tf32(v.y);
        const float hz = round_tf32(v.z), hw = round_tf32(v.w);
        hi[4 * c + 0] = __float_as_uint(hx); hi[4 * c + 1] = __float_as_uint(hy);
        hi[4 * c + 2] = __float_as_uint(hz); hi[4 * c + 3] = __float_as_uint(hw);
        lo[4 * c + 0] = __float_as_uint(round_tf32(v.x - hx));
        lo[4 * c + 1] = __float_as_uint(round_tf32(v.y - hy));
        lo[4 * c + 2] = __float_as_uint(round_tf32(v.z - hz));
        lo[4 * c + 3] = __float_as_uint(round_tf32(v.w - hw));
      }
      mbar_wait_prof(b_empty(ta), rt.ph ^ 1u, prof, w1);   // MMAs that read this stage are done
      tcgen05_fence_after();
      if (!(P.dbg & 8)) {
        tmem_st32(ta0 + (uint32_t)(ta * kTsACols), hi);
        tmem_st32(ta0 + (uint32_t)(ta * kTsACols + kBK), lo);
      }
      // the stores consumed every register the loads filled: the raw tile may be refilled.
      // One arrival per warp (128 arrivals on one mbarrier serialise: ~1,000 clocks per k-step)
      __syncwarp();
      if (lane == 0) mbar_arrive(a_empty(sa));
      tmem_st_wait();
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CG == 2) mbar_arrive_cluster(mapa_u32(ready(ta), 0u));   // the leader issues the MMAs
        else mbar_arrive(ready(ta));
      }
    }
    if (prof && threadIdx.x == 64) {
      g_ts_prof[8] = w0; g_ts_prof[9] = w1;
      g_ts_prof[10] = (unsigned long long)(clock64() - t_begin);
    }
  } else {
    // ===== epilogue: warps 6..9 =====
    const int q = warp & 3;
    const uint32_t epi0 = b_base + (uint32_t)SB * 2u * b_bytes;   // 1024-byte aligned
    const uint32_t epi = epi0 + (uint32_t)(q * kEpiWarpBytes);
    const uint32_t stage = epi0 + (uint32_t)(q * kEpiTmaWarpBytes);
    const bool vec_ok = (P.ldc % 4 == 0) && ((reinterpret_cast<uintptr_t>(P.C) & 15) == 0);
    int blk = 0;
    EpiProf pf{prof && warp == 6, 0, 0, 0, 0};
    if (P.tma_store && lane == 0) asm volatile("prefetch.tensormap [%0];" ::"l"(&mapC) : "memory");
    for (int tl = 0, nt = 0, mb = unit0; tl < n_tiles_cta; ++tl) {
      const int buf = tl & 1;
      mbar_wait_prof(acc_full(buf), (uint32_t)(tl >> 1) & 1u, prof, w0);
      tcgen05_fence_after();
      const uint32_t tbase = tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * acc_stride);
      if (P.tma_store)
        gemm_epilogue_rows_tma(P, &mapC, tbase, BN, BN, split_acc, stage,
                               (mb * CG + (int)rank) * kBM + q * 32, nt * BN, lane, blk,
                               pf);
      else
        gemm_epilogue_rows<0>(P, tbase, BN, BN, split_acc, epi, (mb * CG + (int)rank) * kBM + q * 32,
                              nt * BN, lane, vec_ok);
      if (++nt == P.n_tiles) {
        nt = 0;
        mb += ustride;
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CG == 2) mbar_arrive_cluster(mapa_u32(acc_empty(buf), 0u));
        else mbar_arrive(acc_empty(buf));
      }
    }
    if (P.tma_store && elect_one()) tma_store_wait_all();   // shared memory is read, C is written
    if (PROF && (P.dbg & 16) && threadIdx.x == 192 && blockIdx.x < 160) {
      unsigned long long ns;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(ns));
      g_ts_cta[2 * blockIdx.x] = (unsigned long long)(clock64() - t_begin);
      g_ts_cta[2 * blockIdx.x + 1] = ns - t_begin_ns;
    }
    if (prof && threadIdx.x == 192) {
      g_ts_prof[11] = w0; g_ts_prof[12] = (unsigned long long)(clock64() - t_begin);
      g_ts_prof[13] = (unsigned long long)n_tiles_cta * (unsigned long long)nk;
      g_ts_prof[14] = pf.a + (pf.b << 32);   // packed: the probe prints them split
      g_ts_prof[15] = pf.c + (pf.d << 32);
    }
  }
  tcgen05_fence_before();
  if (CG == 2) cluster_sync_all();   // the peer's shared memory and TMEM are in use until here
  else __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    if (CG == 2)
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_d),
                   "r"((uint32_t)P.tmem_cols)
                   : "memory");
    else
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d),
                   "r"((uint32_t)P.tmem_cols)
                   : "memory");
  }
}

// W [K,N] (row stride ldw) -> Wh, Wl as [Np][Kp] (K-major, zero padded)
__global__ void __launch_bounds__(256)
gemm_prep_weight_kernel(const float* __restrict__ W, int64_t ldw, int K, int N, int Kp, int Np,
                        float* __restrict__ Wh, float* __restrict__ Wl) {
  const int64_t total = (int64_t)Np * Kp;
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * 256) {
    const int n = (int)(t / Kp), k = (int)(t - (int64_t)n * Kp);
    float v = (n < N && k < K) ? W[(int64_t)k * ldw + n] : 0.f;
    const float h = round_tf32(v);
    Wh[t] = h;
    Wl[t] = round_tf32(v - h);
  }
}

// bf16 W [K,N] (row stride ldw) -> Wt [Np][Kp] (K-major, zero padded)
__global__ void __launch_bounds__(256)
gemm_prep_weight_bf16_kernel(const __nv_bfloat16* __restrict__ W, int64_t ldw, int K, int N, int Kp,
                             int Np, __nv_bfloat16* __restrict__ Wt) {
  const int64_t total = (int64_t)Np * Kp;
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * 256) {
    const int n = (int)(t / Kp), k = (int)(t - (int64_t)n * Kp);
    Wt[t] = (n < N && k < K) ? W[(int64_t)k * ldw + n] : __float2bfloat16_rn(0.f);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, []() {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// 2-D fp32 row-major [rows, cols] with row stride ld (elements); box = [box_rows, 32 cols]
int make_map(CUtensorMap* map, const void* ptr, int64_t rows, int64_t cols, int64_t ld,
             int box_rows, int bk, bool atom32 = false, int esz = 4) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return PYGB_E_UNSUPPORTED;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * esz};
  cuuint32_t box[2] = {(cuuint32_t)bk, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, esz == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                  2, const_cast<void*>(ptr), dims, strides,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                         : (bk * esz == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B),
                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) for [%lld x %lld] ld %lld", (int)r,
              (long long)rows, (long long)cols, (long long)ld);
    return PYGB_E_CUDA;
  }
  return PYGB_OK;
}

// A as `slices` planes of [rows, cols] (row stride ld, plane stride `plane` elements).
int make_map_a(CUtensorMap* map, const float* ptr, int64_t rows, int64_t cols, int64_t ld,
               int64_t slices, int64_t plane, int box_rows, int bk) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return PYGB_E_UNSUPPORTED;
  }
  cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)slices};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 4, (cuuint64_t)(slices > 1 ? plane : ld * rows) * 4};
  cuuint32_t box[3] = {(cuuint32_t)bk, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(ptr), dims, strides,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  bk == 16 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) for A [%lld x %lld x %lld] ld %lld", (int)r,
              (long long)slices, (long long)rows, (long long)cols, (long long)ld);
    return PYGB_E_CUDA;
  }
  return PYGB_OK;
}

inline int round_up(int x, int m) { return (x + m - 1) / m * m; }

// =====================================================================================
// K5t — dW = X^T · G : the weight gradient of X·W (autograd of A.9 / A.10, row a15).
//
// The reduction runs over the ROWS of X [R, Kx] and G [R, Ng] (R = 2.4 M nodes, Kx, Ng <=
// 256), so both operands are "MN-major" for the tensor core: a TMA box of 16 rows x 32
// floats (TMA SWIZZLE_128B_ATOM_32B) IS four canonical MN-major SW128_32B atoms (4 rows
// of 128 bytes each; cute/atom/mma_traits_sm100.hpp, make_umma_desc<Major::MN>), so no
// transposed copy of either matrix is made.  Split-K over persistent CTAs: CTA b owns a
// contiguous range of 16-row steps and accumulates its whole [Kx, Ng] partial in TMEM
// (MH = ceil(Kx/128) halves of 128 lanes x BN columns <= 512 columns); the partials are
// added in CTA order by a second kernel — deterministic, no atomics.  The tensor core's
// fp32 accumulator TRUNCATES, which biases a long accumulation chain (1.2e-4 relative over
// the 16.5 K rows a CTA owns at cfg5): the accumulator is therefore drained every `chunk`
// steps (1,024 rows) and the chunks are added with round-to-nearest fp32 in the CTA's
// partial (L2-resident read-modify-write by the epilogue warps; a second TMEM buffer
// hides the drain when 2·MH·BN <= 512 columns).  3xTF32 like the
// forward: both tiles are split in place (X = Xh + Xl, G = Gh + Gl) by the splitter warps
// and D += Xh·Gh + Xl·Gh + Xh·Gl.  HBM-bound: X and G are read exactly once.
// =====================================================================================
constexpr int kTRows = 16;                  // rows per stage = two MMA k-steps of 8
constexpr int kBoxBytes = kTRows * 128;     // one box: 16 rows x 32 floats

// MN-major tf32 operands have ONE legal shared-memory layout (cutlass sm100_common.inl:
// "for mn-major tf32 operands, SW128_32B is the only available smem layout"): 128-byte rows
// whose 32-byte chunks are XOR-swizzled with (row % 4) — TMA's SWIZZLE_128B_ATOM_32B, UMMA
// layout type 1 — in atoms of 4 rows (512 bytes).  32-element chunks along M/N are `lbo`
// bytes apart, 4-row atoms along K `sbo` bytes apart; one MMA (K = 8) reads two atoms.
__device__ __forceinline__ uint64_t smem_desc_mn_sw128_32b(uint32_t saddr, uint32_t lbo_bytes) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) |
         (32ull << 32) /* sbo = 512 bytes */ | (1ull << 46) | (1ull << 61);
}

struct XtgParams {
  int R, Kx, Ng;
  int MH;         // 128-row halves of the output (Kx <= 128: 1, else 2)
  int BN;         // Ng rounded to 16
  int nbx, nbg;   // boxes per stage: MH*4 of X, ceil(BN/32) of G
  int stages;
  int tmem_cols;
  int nk;         // 16-row steps in total
  int chunk;      // steps per accumulation chunk (the TMEM accumulator is drained after each)
  int nbuf;       // accumulator buffers in TMEM: 2 when 2*MH*BN <= 512 columns
  float* part;    // [grid][MH*128][BN]
};

template <int PASSES>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_xtg_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapG,
                const XtgParams P) {
  extern __shared__ __align__(1024) unsigned char gsm[];
  __shared__ __align__(8) uint64_t bars[3 * kMaxStages + 4];
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = P.stages, BN = P.BN, MH = P.MH, nbuf = P.nbuf;
  const uint32_t x_bytes = (uint32_t)P.nbx * kBoxBytes, g_bytes = (uint32_t)P.nbg * kBoxBytes;
  const uint32_t stage_bytes = (PASSES == 3 ? 2 : 1) * (x_bytes + g_bytes);
  const uint32_t base = (smem_u32(gsm) + 1023u) & ~1023u;
  auto x_hi = [&](int s) { return base + (uint32_t)s * stage_bytes; };
  auto x_lo = [&](int s) { return x_hi(s) + x_bytes; };
  auto g_hi = [&](int s) { return x_hi(s) + (PASSES == 3 ? 2 : 1) * x_bytes; };
  auto g_lo = [&](int s) { return g_hi(s) + g_bytes; };
  const uint32_t bar0 = smem_u32(bars);
  auto full = [&](int s) { return bar0 + 8u * s; };
  auto ready = [&](int s) { return bar0 + 8u * (kMaxStages + s); };
  auto empty = [&](int s) { return bar0 + 8u * (2 * kMaxStages + s); };
  auto acc_full = [&](int b) { return bar0 + 8u * (3 * kMaxStages + b); };
  auto acc_empty = [&](int b) { return bar0 + 8u * (3 * kMaxStages + 2 + b); };

  // this CTA's contiguous range of 16-row steps (grid <= nk: at least one step each)
  const int k_begin = (int)((int64_t)P.nk * blockIdx.x / gridDim.x);
  const int k_end = (int)((int64_t)P.nk * (blockIdx.x + 1) / gridDim.x);
  const int nsteps = k_end - k_begin;
  const int nchunks = (nsteps + P.chunk - 1) / P.chunk;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(full(s), 1);
      mbar_init(ready(s), 128);
      mbar_init(empty(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(acc_full(b), 1);
      mbar_init(acc_empty(b), 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(&tmem_base_s)),
                 "r"((uint32_t)P.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_d = tmem_base_s;

  if (warp == 0) {
    // ===== TMA producer: nbx + nbg boxes of 16 rows x 32 columns per stage =====
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mapX) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mapG) : "memory");
      for (int it = 0; it < nsteps; ++it) {
        const int s = it % S;
        const uint32_t ph = (uint32_t)(it / S) & 1u;
        mbar_wait(empty(s), ph ^ 1u);
        mbar_expect_tx(full(s), x_bytes + g_bytes);
        const int r0 = (k_begin + it) * kTRows;
        for (int c = 0; c < P.nbx; ++c)
          tma_load_2d(x_hi(s) + (uint32_t)c * kBoxBytes, &mapX, full(s), c * 32, r0);
        for (int c = 0; c < P.nbg; ++c)
          tma_load_2d(g_hi(s) + (uint32_t)c * kBoxBytes, &mapG, full(s), c * 32, r0);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: A = X^T and B = G, both MN-major (whole warp walks, one lane issues) =====
    {
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) |
                             ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(kBM >> 4) << 24);
      int it = 0;
      for (int ch = 0; ch < nchunks; ++ch) {
        const int buf = ch % nbuf;
        mbar_wait(acc_empty(buf), ((uint32_t)(ch / nbuf) & 1u) ^ 1u);  // drained by the epilogue
        tcgen05_fence_after();
        const uint32_t dcol = tmem_d + (uint32_t)(buf * MH * BN);
        const int it_end = min(nsteps, (ch + 1) * P.chunk);
        for (int i0 = it; it < it_end; ++it) {
          const int s = it % S;
          const uint32_t ph = (uint32_t)(it / S) & 1u;
          mbar_wait(PASSES == 3 ? ready(s) : full(s), ph);
          tcgen05_fence_after();
          if (elect_one()) {
#pragma unroll
          for (int j = 0; j < kTRows / 8; ++j) {   // MMA k-steps of 8 rows = two 4-row atoms, 1024 bytes apart
            const uint64_t dgh = smem_desc_mn_sw128_32b(g_hi(s) + j * 1024u, kBoxBytes);
            const uint64_t dgl = smem_desc_mn_sw128_32b(g_lo(s) + j * 1024u, kBoxBytes);
            for (int h = 0; h < MH; ++h) {
              const uint32_t hoff = (uint32_t)(h * 4) * kBoxBytes + j * 1024u;
              const uint64_t dxh = smem_desc_mn_sw128_32b(x_hi(s) + hoff, kBoxBytes);
              const uint64_t dxl = smem_desc_mn_sw128_32b(x_lo(s) + hoff, kBoxBytes);
              const uint32_t dc = dcol + (uint32_t)(h * BN);
              mma_tf32(dc, dxh, dgh, idesc, (it > i0 || j > 0) ? 1u : 0u);
              if (PASSES == 3) {
                mma_tf32(dc, dxl, dgh, idesc, 1u);
                mma_tf32(dc, dxh, dgl, idesc, 1u);
              }
            }
          }
          tcgen05_commit(empty(s));
          }
          __syncwarp();
        }
        if (elect_one()) tcgen05_commit(acc_full(buf));
        __syncwarp();
      }
    }
  } else if (warp < 6) {
    // ===== splitter: both tiles, in place (layout-agnostic: element-wise on the raw bytes) =====
    if (PASSES == 3) {
      const int t = threadIdx.x - 64;
      const int nvec = (int)((x_bytes + g_bytes) / 16 / 128);   // float4 per thread: boxes * 1
      for (int it = 0; it < nsteps; ++it) {
        const int s = it % S;
        const uint32_t ph = (uint32_t)(it / S) & 1u;
        mbar_wait(full(s), ph);
        const uint32_t xh = x_hi(s);
        const uint32_t lo_x = x_bytes, lo_g = g_bytes;           // hi -> lo distance per region
        const uint32_t g_start = (uint32_t)P.nbx * 128u;         // first float4 of the G region
#pragma unroll 4
        for (int i = 0; i < nvec; ++i) {
          const uint32_t v4 = (uint32_t)(t + i * 128);
          const bool is_g = v4 >= g_start;
          // hi regions: X at [0, x_bytes), G at [2*x_bytes, 2*x_bytes + g_bytes)
          const uint32_t src = xh + (is_g ? 2u * x_bytes + (v4 - g_start) * 16u : v4 * 16u);
          const uint32_t dst = src + (is_g ? lo_g : lo_x);
          float4 v;
          asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                       : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                       : "r"(src));
          float4 h, l;
          h.x = round_tf32(v.x); h.y = round_tf32(v.y);
          h.z = round_tf32(v.z); h.w = round_tf32(v.w);
          l.x = round_tf32(v.x - h.x); l.y = round_tf32(v.y - h.y);
          l.z = round_tf32(v.z - h.z); l.w = round_tf32(v.w - h.w);
          asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(src), "f"(h.x), "f"(h.y),
                       "f"(h.z), "f"(h.w)
                       : "memory");
          asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "f"(l.x), "f"(l.y),
                       "f"(l.z), "f"(l.w)
                       : "memory");
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_arrive(ready(s));
      }
    }
  } else {
    // ===== epilogue: drain every chunk into this CTA's partial [MH*128, BN] (fp32 adds with
    // round-to-nearest; every thread owns its addresses, so plain loads and stores) =====
    const int q = warp & 3;
    for (int ch = 0; ch < nchunks; ++ch) {
      const int buf = ch % nbuf;
      mbar_wait(acc_full(buf), (uint32_t)(ch / nbuf) & 1u);
      tcgen05_fence_after();
      for (int h = 0; h < MH; ++h) {
        float* prow = P.part + ((int64_t)blockIdx.x * (MH * kBM) + h * kBM + q * 32 + lane) * BN;
        const uint32_t tbase =
            tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * MH * BN + h * BN);
        for (int c0 = 0; c0 < BN; c0 += 32) {
          uint32_t r[32];
          const int nc = c0 + 32 <= BN ? 32 : 16;
          if (nc == 32) tmem_ld32(tbase + (uint32_t)c0, r);
          else tmem_ld16(tbase + (uint32_t)c0, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            if (j < nc) {
              float4 v = make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]),
                                     __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
              float4* dst = reinterpret_cast<float4*>(prow + c0 + j);
              if (ch > 0) {
                const float4 o = *dst;
                v.x += o.x, v.y += o.y, v.z += o.z, v.w += o.w;
              }
              *dst = v;
            }
          }
        }
      }
      tcgen05_fence_before();
      mbar_arrive(acc_empty(buf));  // 128 arrivals: the MMA warp may reuse this buffer
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d),
                 "r"((uint32_t)P.tmem_cols)
                 : "memory");
  }
}

// dW[r, c] = sum over the CTAs' partials, in CTA order
__global__ void __launch_bounds__(256)
gemm_xtg_reduce_kernel(const float* __restrict__ part, int n_part, int rows_p, int BN, int Kx,
                       int Ng, float* __restrict__ dW, int64_t ldw) {
  const int total = Kx * Ng;
  for (int t = blockIdx.x * 256 + threadIdx.x; t < total; t += gridDim.x * 256) {
    const int r = t / Ng, c = t - r * Ng;
    const float* p = part + (int64_t)r * BN + c;
    float acc = 0.f;
    for (int b = 0; b < n_part; ++b) acc += p[(int64_t)b * rows_p * BN];
    dW[(int64_t)r * ldw + c] = acc;
  }
}

// 2-D fp32 [rows, cols] (row stride ld), box = 16 rows x 32 columns, SWIZZLE_128B_ATOM_32B
int make_map_rows16(CUtensorMap* map, const float* ptr, int64_t rows, int64_t cols, int64_t ld) {
  return make_map(map, ptr, rows, cols, ld, kTRows, 32, true);
}

struct XtgPlan {
  int MH, BN, nbx, nbg, stages, tmem_cols, grid, nk, nbuf;
  size_t smem, part_floats;
};

int xtg_plan(int64_t R, int64_t Kx, int64_t Ng, int passes, XtgPlan* pl) {
  int sms = 148, dev = 0;
  PYGB_CUDA(cudaGetDevice(&dev));
  PYGB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  pl->MH = Kx <= kBM ? 1 : 2;
  pl->BN = round_up((int)Ng, 16);
  pl->nbx = pl->MH * 4;
  pl->nbg = (pl->BN + 31) / 32;
  const size_t stage_bytes = (size_t)(passes == 3 ? 2 : 1) * (pl->nbx + pl->nbg) * kBoxBytes;
  int stages = (int)((226 * 1024 - 1024) / stage_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  pl->stages = stages;
  pl->smem = stages * stage_bytes + 1024;
  pl->nbuf = 2 * pl->MH * pl->BN <= 512 ? 2 : 1;
  int tc = 32;
  while (tc < pl->nbuf * pl->MH * pl->BN) tc <<= 1;
  pl->tmem_cols = tc;
  pl->nk = (int)ceil_div(R, kTRows);
  pl->grid = pl->nk < sms ? pl->nk : sms;
  if (pl->grid < 1) pl->grid = 1;
  pl->part_floats = (size_t)pl->grid * pl->MH * kBM * pl->BN;
  return PYGB_OK;
}

}  // namespace
}  // namespace pygb

using namespace pygb;

extern "C" {

size_t pygb_gemm_weight_workspace(int64_t K, int64_t N) {
  const size_t Kp = (size_t)round_up((int)K, kBK), Np = (size_t)round_up((int)N, 16);
  return 2 * align_up(Kp * Np * sizeof(float), 1024) + 1024;
}

struct GemmRoute {
  float* const* tab;
  int64_t fc, dup, row0, ldd;
  int64_t max_ctas;   // > 0: cap of the persistent grid (NVLink-bound epilogue: a few SMs are enough)
};

static int gemm_xw_impl(const float* A, int64_t lda, int64_t a_slices, int64_t a_plane, int64_t M,
                        int64_t K, const float* W,
                        int64_t ldw, int64_t N, const float* bias, int relu, float* C,
                        int64_t ldc, const GemmRoute* route, int passes, void* ws,
                        size_t ws_bytes, pygb_stream_t s) {
  PYGB_REQUIRE(M >= 0 && K > 0 && N > 0 && M < (1ll << 31) - 256 && K < (1 << 20) && N < (1 << 20),
               "pygb_gemm_xw: bad sizes");
  PYGB_REQUIRE(passes == 1 || passes == 3, "pygb_gemm_xw: passes must be 1 (tf32) or 3 (3xtf32)");
  PYGB_REQUIRE(a_slices >= 1 && K % a_slices == 0, "pygb_gemm_xw: K must be a_slices planes");
  const int64_t a_kw = K / a_slices;
  PYGB_REQUIRE(lda >= a_kw && ldw >= N && ldc >= N, "pygb_gemm_xw: bad leading dimensions");
  PYGB_REQUIRE(a_slices == 1 || (a_kw % kBK == 0 && a_plane >= lda * M && a_plane % 4 == 0),
               "pygb_gemm_xw: sliced A needs %d-column multiples per plane and a 16-byte "
               "aligned plane stride", kBK);
  if (M == 0) return PYGB_OK;
  PYGB_REQUIRE(A && W && (C || route) && ws, "pygb_gemm_xw: null pointer");
  if (!aligned16(A) || (lda % 4) != 0) {
    set_error("pygb_gemm_xw: A must be 16-byte aligned with a row stride that is a multiple of 4");
    return PYGB_E_UNSUPPORTED;
  }
  if (ws_bytes < pygb_gemm_weight_workspace(K, N)) {
    set_error("pygb_gemm_xw: workspace too small");
    return PYGB_E_WORKSPACE;
  }
  cudaStream_t st = as_stream(s);
  const int Kp = round_up((int)K, kBK), Np = round_up((int)N, 16);
  char* wsb = static_cast<char*>(ws);
  wsb = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(wsb), 1024));
  float* Wh = reinterpret_cast<float*>(wsb);
  float* Wl = reinterpret_cast<float*>(wsb + align_up((size_t)Kp * Np * sizeof(float), 1024));
  {
    int64_t blocks = ceil_div((int64_t)Kp * Np, 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    gemm_prep_weight_kernel<<<(unsigned)blocks, 256, 0, st>>>(W, ldw, (int)K, (int)N, Kp, Np, Wh,
                                                             Wl);
    PYGB_LAUNCH_CHECK();
  }
  GemmParams P;
  P.dbg = 0;
  P.tma_store = 0;
  P.ts_ra = P.ts_sb = P.ts_ta = 0;
  P.bias = bias;
  P.relu = relu != 0;
  P.c_tab = route ? route->tab : nullptr;
  P.c_fc = route ? (int)route->fc : 1;
  P.c_dup = route ? (int)route->dup : 0;
  P.c_row0 = route ? route->row0 : 0;
  P.c_ldd = route ? route->ldd : 0;
  P.C = C;
  P.ldc = ldc;
  P.M = (int)M;
  P.N = (int)N;
  P.K = (int)K;
  P.a_kw = (int)a_kw;
  P.a_3d = a_slices > 1;
  P.BN = Np < 256 ? Np : 256;
  P.n_tiles = (int)ceil_div(Np, P.BN);
  int sms = 148;
  {
    int dev = 0;
    PYGB_CUDA(cudaGetDevice(&dev));
    PYGB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  }
  // Tile shape.  256 x BN x 16 shares every W stage between two 128-row halves: worth it when
  // W traffic dominates the stage (3 passes, wide BN) and there are enough tiles for every SM.
  int mh = 1, bk = kBK;
  // a k-step of 16 wastes less of the last step when K is just above a multiple of 32
  // (K = 100: 112 instead of 128 columns of MMAs and W traffic; profiles/r01_k5_tiles.md)
  if (round_up((int)K, 16) < round_up((int)K, 32) && K > 32) bk = 16;
  if (const char* e = getenv("PYGB_GEMM_TILE")) {     // tuning / tests: "128x32", "128x16", "256x16"
    int tm = 0, tk = 0;
    if (sscanf(e, "%dx%d", &tm, &tk) == 2 && (tm == 128 || tm == 256) && (tk == 16 || tk == 32) &&
        !(tm == 256 && tk == 32)) {
      mh = tm / 128;
      bk = tk;
    }
  }
  // 3xTF32 default: the kernel that keeps Ah / Al in tensor memory (gemm_ts_kernel).
  // PYGB_GEMM_TS=0 or an explicit PYGB_GEMM_TILE selects the shared-memory-operand kernels.
  bool use_ts = passes == 3 && getenv("PYGB_GEMM_TILE") == nullptr;
  if (const char* e = getenv("PYGB_GEMM_TS")) use_ts = use_ts && atoi(e) != 0;
  if (use_ts) {
    // BN <= 128 (TMEM: two accumulators + the A ring).  The correction terms get their own
    // accumulator (split_acc) when 4*BN + the ring fit: BN <= 64, also chosen for N <= 128 with a
    // long K, where the truncating accumulator matters (see GemmParams::split_acc).
    int nt = (int)ceil_div(Np, 128);
    if (Np > 64 && Np <= 128 && K > 256) nt = 2;
    P.n_tiles = nt;
    P.BN = round_up((int)ceil_div(Np, nt), 16);
    P.split_acc = P.BN <= 64 ? 1 : 0;
    if (const char* e = getenv("PYGB_GEMM_SPLIT_ACC")) P.split_acc = P.split_acc && atoi(e) != 0;
    const int acc_stride = P.BN * (P.split_acc ? 2 : 1);
    P.acc_bufs = 2;
    // CTA pairs (cta_group::2) whenever there are at least two 256-row units; PYGB_GEMM_CG=1: single CTAs
    int cg = M > 2 * kBM ? 2 : 1;
    if (const char* e = getenv("PYGB_GEMM_CG")) cg = atoi(e) == 1 ? 1 : cg;
    P.m_tiles = (int)ceil_div(M, cg * kBM);   // units of cg * 128 rows
    // ring depths: TMEM stages from what the accumulators leave; W stages (L2 latency) 3, or 4
    // when the tiles are small; every remaining 16 KB of shared memory is a raw A tile in flight
    const int nk = (int)ceil_div(K, kBK);
    const size_t b_stage = 2 * (size_t)(P.BN / cg) * kBK * 4;   // per CTA: its part of Wh and Wl
    // C through TMA stores when it is an ordinary 16-byte aligned matrix (not routed)
    // (N % 4: the store engine clips columns >= N in 16-byte units — measured: with N = 47 it
    // wrote column 47 of a wider output matrix)
    P.tma_store = (!route && C && aligned16(C) && ldc % 4 == 0 && N % 4 == 0) ? 1 : 0;
    if (const char* e = getenv("PYGB_GEMM_TMA_STORE")) P.tma_store = P.tma_store && atoi(e) != 0;
    const size_t epi_bytes = P.tma_store ? (size_t)kEpiTmaBytes : (size_t)kEpiBytes;
    const size_t budget = 226 * 1024 - 1024 - epi_bytes;
    int ta = (512 - 2 * acc_stride) / kTsACols;
    int sb = 4;
    if (const char* e = getenv("PYGB_GEMM_TS_SB")) sb = atoi(e);
    if (sb > ta) sb = ta;
    if (sb > kTsMaxSB) sb = kTsMaxSB;
    while (sb > 2 && (budget - sb * b_stage) / kATileBytes < 4) --sb;   // keep >= 4 raw A tiles
    if (sb < 2) sb = 2;
    ta = sb;
    int ra = (int)((budget - sb * b_stage) / kATileBytes);
    if (const char* e = getenv("PYGB_GEMM_TS_RA")) ra = ra < atoi(e) ? ra : atoi(e);
    if (ra > kTsMaxRA) ra = kTsMaxRA;
    PYGB_REQUIRE(ta >= 2 && ra >= 2, "pygb_gemm_xw: tile does not fit shared / tensor memory");
    P.dbg = 0;
    if (const char* e = getenv("PYGB_GEMM_TS_DBG")) P.dbg = atoi(e);
    P.ts_ra = ra;
    P.ts_sb = sb;
    P.ts_ta = ta;
    P.stages = sb;
    int tc = 32;
    while (tc < 2 * acc_stride + ta * kTsACols) tc <<= 1;
    P.tmem_cols = tc;
    const size_t smem = ra * (size_t)kATileBytes + sb * b_stage + epi_bytes + 1024;
    (void)nk;
    CUtensorMap mapA, mapBh, mapBl, mapC;
    int rc = a_slices > 1 ? make_map_a(&mapA, A, M, a_kw, lda, a_slices, a_plane, kBM, kBK)
                          : make_map(&mapA, A, M, K, lda, kBM, kBK);
    if (rc) return rc;
    rc = make_map(&mapBh, Wh, Np, Kp, Kp, P.BN / cg, kBK);
    if (rc) return rc;
    rc = make_map(&mapBl, Wl, Np, Kp, Kp, P.BN / cg, kBK);
    if (rc) return rc;
    if (P.tma_store) {
      rc = make_map(&mapC, C, M, N, ldc, 32, 32);   // 32 x 32 fp32 blocks, 128-byte swizzle
      if (rc) return rc;
    } else {
      mapC = mapA;   // unused
    }
    int64_t units = P.m_tiles < sms / cg ? P.m_tiles : sms / cg;
    if (route && route->max_ctas > 0 && units * cg > route->max_ctas)
      units = route->max_ctas / cg > 0 ? route->max_ctas / cg : 1;
    const int64_t ctas = units * cg;
    if (cg == 2) {
      auto kern = (P.dbg & 16) ? gemm_ts_kernel<2, true> : gemm_ts_kernel<2, false>;
      PYGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)(226 * 1024)));
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3((unsigned)ctas);
      cfg.blockDim = dim3(kTsThreads);
      cfg.dynamicSmemBytes = smem;
      cfg.stream = st;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = 2;
      attr[0].val.clusterDim.y = 1;
      attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      PYGB_CUDA(cudaLaunchKernelEx(&cfg, kern, mapA, mapBh, mapBl, mapC, P));
    } else {
      auto kern = (P.dbg & 16) ? gemm_ts_kernel<1, true> : gemm_ts_kernel<1, false>;
      PYGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)(226 * 1024)));
      kern<<<(unsigned)ctas, kTsThreads, smem, st>>>(mapA, mapBh, mapBl, mapC, P);
    }
    PYGB_LAUNCH_CHECK();
    if (P.dbg & 16) {   // probe: synchronises and prints CTA 0's wait clocks
      unsigned long long h[16], hc[2 * 160];
      PYGB_CUDA(cudaStreamSynchronize(st));
      PYGB_CUDA(cudaMemcpyFromSymbol(h, g_ts_prof, sizeof(h)));
      PYGB_CUDA(cudaMemcpyFromSymbol(hc, g_ts_cta, sizeof(hc)));
      {
        unsigned long long cmin = ~0ull, cmax = 0, nmin = ~0ull, nmax = 0;
        for (int i = 0; i < (int)ctas && i < 160; ++i) {
          if (hc[2 * i] < cmin) cmin = hc[2 * i];
          if (hc[2 * i] > cmax) cmax = hc[2 * i];
          if (hc[2 * i + 1] < nmin) nmin = hc[2 * i + 1];
          if (hc[2 * i + 1] > nmax) nmax = hc[2 * i + 1];
        }
        fprintf(stderr, "[ts prof] per-CTA clocks min %llu max %llu, ns min %llu max %llu (CTA0 %llu clk %llu ns)\n",
                cmin, cmax, nmin, nmax, hc[0], hc[1]);
      }
      fprintf(stderr,
              "[ts prof] cg=%d M=%d K=%d N=%d BN=%d ra=%d sb=%d ta=%d steps=%llu | A-prod wait a_empty %llu of %llu"
              " | W-prod wait b_empty %llu of %llu | MMA wait acc_empty %llu b_full %llu ready %llu of"
              " %llu (epi: tmem ld %llu, stage wait %llu, st.shared+fence %llu, issue %llu) | split wait a_full %llu b_empty %llu of %llu | epi wait acc_full %llu of %llu\n",
              cg, P.M, P.K, P.N, P.BN, ra, sb, ta, h[13], h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7],
              h[14] & 0xffffffffull, h[14] >> 32, h[15] & 0xffffffffull, h[15] >> 32, h[8], h[9],
              h[10], h[11], h[12]);
    }
    return PYGB_OK;
  }
  const int acc_cols = mh * P.BN;
  P.split_acc = (passes == 3 && 4 * acc_cols <= 512) ? 1 : 0;
  if (const char* e = getenv("PYGB_GEMM_SPLIT_ACC")) P.split_acc = P.split_acc && atoi(e) != 0;
  const int acc_stride = acc_cols * (P.split_acc ? 2 : 1);
  P.acc_bufs = 2 * acc_stride <= 512 ? 2 : 1;
  int tc = 32;
  while (tc < P.acc_bufs * acc_stride) tc <<= 1;
  P.tmem_cols = tc;
  P.m_tiles = (int)ceil_div(M, mh * kBM);
  const size_t b_bytes = (size_t)P.BN * bk * 4;
  const size_t a_bytes = (size_t)mh * kBM * bk * 4;
  const size_t stage_bytes = (passes == 3 ? 2 : 1) * (a_bytes + b_bytes);
  const size_t budget = 226 * 1024;
  int stages = (int)((budget - 1024 - kEpiBytes) / stage_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  const int nk = (int)ceil_div(K, bk);
  if (stages > nk) stages = nk < 2 ? 2 : nk;
  PYGB_REQUIRE(stages >= 2, "pygb_gemm_xw: tile does not fit shared memory");
  P.stages = stages;
  const size_t smem = stages * stage_bytes + kEpiBytes + 1024;

  CUtensorMap mapA, mapBh, mapBl;
  int rc = a_slices > 1 ? make_map_a(&mapA, A, M, a_kw, lda, a_slices, a_plane, mh * kBM, bk)
                        : make_map(&mapA, A, M, K, lda, mh * kBM, bk);
  if (rc) return rc;
  rc = make_map(&mapBh, Wh, Np, Kp, Kp, P.BN, bk);
  if (rc) return rc;
  rc = make_map(&mapBl, Wl, Np, Kp, Kp, P.BN, bk);
  if (rc) return rc;

  const int64_t tiles = (int64_t)P.m_tiles * P.n_tiles;
  int64_t ctas = tiles < sms ? tiles : sms;
  if (route && route->max_ctas > 0 && ctas > route->max_ctas) ctas = route->max_ctas;
  dim3 grid((unsigned)ctas);
#define PYGB_GEMM_LAUNCH(PS, MHV, BKV)                                                         \
  do {                                                                                         \
    PYGB_CUDA(cudaFuncSetAttribute(gemm_tf32_kernel<PS, MHV, BKV>,                             \
                                   cudaFuncAttributeMaxDynamicSharedMemorySize,                \
                                   (int)(226 * 1024)));                                        \
    gemm_tf32_kernel<PS, MHV, BKV><<<grid, kGemmThreads, smem, st>>>(mapA, mapBh, mapBl, P);   \
  } while (0)
  if (passes == 3) {
    if (mh == 2) PYGB_GEMM_LAUNCH(3, 2, 16);
    else if (bk == 16) PYGB_GEMM_LAUNCH(3, 1, 16);
    else PYGB_GEMM_LAUNCH(3, 1, 32);
  } else {
    if (mh == 2) PYGB_GEMM_LAUNCH(1, 2, 16);
    else if (bk == 16) PYGB_GEMM_LAUNCH(1, 1, 16);
    else PYGB_GEMM_LAUNCH(1, 1, 32);
  }
#undef PYGB_GEMM_LAUNCH
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

size_t pygb_gemm_xtg_workspace(int64_t R, int64_t Kx, int64_t Ng) {
  if (R <= 0 || Kx <= 0 || Ng <= 0 || Kx > 256 || Ng > 256) return 256;
  XtgPlan pl;
  if (xtg_plan(R, Kx, Ng, 3, &pl) != PYGB_OK) return 0;
  return align_up(pl.part_floats * sizeof(float), 256) + 256;
}

int pygb_gemm_xtg(const float* X, int64_t ldx, const float* G, int64_t ldg, int64_t R, int64_t Kx,
                  int64_t Ng, float* dW, int64_t ldw, int passes, void* ws, size_t ws_bytes,
                  pygb_stream_t s) {
  PYGB_REQUIRE(R >= 0 && Kx > 0 && Ng > 0 && R < (1ll << 31) - 256 && ldx >= Kx && ldg >= Ng &&
                   ldw >= Ng,
               "pygb_gemm_xtg: bad sizes");
  PYGB_REQUIRE(passes == 1 || passes == 3, "pygb_gemm_xtg: passes must be 1 (tf32) or 3 (3xtf32)");
  PYGB_REQUIRE(dW, "pygb_gemm_xtg: null pointer");
  cudaStream_t st = as_stream(s);
  if (R == 0) {
    PYGB_CUDA(cudaMemset2DAsync(dW, (size_t)ldw * 4, 0, (size_t)Ng * 4, (size_t)Kx, st));
    return PYGB_OK;
  }
  PYGB_REQUIRE(X && G && ws, "pygb_gemm_xtg: null pointer");
  if (Kx > 256 || Ng > 256 || !aligned16(X) || !aligned16(G) || ldx % 4 != 0 || ldg % 4 != 0) {
    set_error("pygb_gemm_xtg: needs Kx, Ng <= 256 and 16-byte aligned rows of X and G");
    return PYGB_E_UNSUPPORTED;
  }
  XtgPlan pl;
  int rc = xtg_plan(R, Kx, Ng, passes, &pl);
  if (rc) return rc;
  PYGB_REQUIRE(pl.stages >= 2, "pygb_gemm_xtg: tile does not fit shared memory");
  if (ws_bytes < pygb_gemm_xtg_workspace(R, Kx, Ng)) {
    set_error("pygb_gemm_xtg: workspace too small");
    return PYGB_E_WORKSPACE;
  }
  XtgParams P;
  P.R = (int)R;
  P.Kx = (int)Kx;
  P.Ng = (int)Ng;
  P.MH = pl.MH;
  P.BN = pl.BN;
  P.nbx = pl.nbx;
  P.nbg = pl.nbg;
  P.stages = pl.stages;
  P.tmem_cols = pl.tmem_cols;
  P.nk = pl.nk;
  P.chunk = 1024 / kTRows;
  P.nbuf = pl.nbuf;
  P.part = reinterpret_cast<float*>(align_up(reinterpret_cast<size_t>(ws), 256));
  CUtensorMap mapX, mapG;
  rc = make_map_rows16(&mapX, X, R, Kx, ldx);
  if (rc) return rc;
  rc = make_map_rows16(&mapG, G, R, Ng, ldg);
  if (rc) return rc;
  if (passes == 3) {
    PYGB_CUDA(cudaFuncSetAttribute(gemm_xtg_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)(226 * 1024)));
    gemm_xtg_kernel<3><<<pl.grid, kGemmThreads, pl.smem, st>>>(mapX, mapG, P);
  } else {
    PYGB_CUDA(cudaFuncSetAttribute(gemm_xtg_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)(226 * 1024)));
    gemm_xtg_kernel<1><<<pl.grid, kGemmThreads, pl.smem, st>>>(mapX, mapG, P);
  }
  PYGB_LAUNCH_CHECK();
  int64_t blocks = ceil_div(Kx * Ng, 256);
  gemm_xtg_reduce_kernel<<<(unsigned)blocks, 256, 0, st>>>(P.part, pl.grid, pl.MH * kBM, pl.BN,
                                                          (int)Kx, (int)Ng, dW, ldw);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

int pygb_gemm_xw_bf16(const void* A, int64_t lda, int64_t M, int64_t K, const void* W, int64_t ldw,
                      int64_t N, const float* bias, int relu, void* C, int64_t ldc, void* ws,
                      size_t ws_bytes, pygb_stream_t s) {
  PYGB_REQUIRE(M >= 0 && K > 0 && N > 0 && M < (1ll << 31) - 256 && K < (1 << 20) && N < (1 << 20),
               "pygb_gemm_xw_bf16: bad sizes");
  PYGB_REQUIRE(lda >= K && ldw >= N && ldc >= N, "pygb_gemm_xw_bf16: bad leading dimensions");
  if (M == 0) return PYGB_OK;
  PYGB_REQUIRE(A && W && C && ws, "pygb_gemm_xw_bf16: null pointer");
  if (!aligned16(A) || (lda % 8) != 0) {
    set_error("pygb_gemm_xw_bf16: A must be 16-byte aligned with a row stride that is a multiple of 8");
    return PYGB_E_UNSUPPORTED;
  }
  if (ws_bytes < pygb_gemm_weight_workspace(K, N)) {
    set_error("pygb_gemm_xw_bf16: workspace too small");
    return PYGB_E_WORKSPACE;
  }
  cudaStream_t st = as_stream(s);
  // k-step: 64 bf16 (128-byte rows) unless 32 wastes less of the last step
  int bk = 64;
  if (round_up((int)K, 32) < round_up((int)K, 64) && K > 64) bk = 32;
  const int Kp = round_up((int)K, 64), Np = round_up((int)N, 16);
  __nv_bfloat16* Wt = reinterpret_cast<__nv_bfloat16*>(align_up(reinterpret_cast<size_t>(ws), 1024));
  {
    int64_t blocks = ceil_div((int64_t)Kp * Np, 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    gemm_prep_weight_bf16_kernel<<<(unsigned)blocks, 256, 0, st>>>(
        static_cast<const __nv_bfloat16*>(W), ldw, (int)K, (int)N, Kp, Np, Wt);
    PYGB_LAUNCH_CHECK();
  }
  GemmParams P;
  P.dbg = 0;
  P.tma_store = 0;
  P.ts_ra = P.ts_sb = P.ts_ta = 0;
  P.bias = bias;
  P.relu = relu != 0;
  P.c_tab = nullptr;
  P.c_fc = 1;
  P.c_dup = 0;
  P.c_row0 = 0;
  P.c_ldd = 0;
  P.C = C;
  P.ldc = ldc;
  P.M = (int)M;
  P.N = (int)N;
  P.K = (int)K;
  P.a_kw = (int)K;
  P.a_3d = 0;
  P.BN = Np < 256 ? Np : 256;
  P.n_tiles = (int)ceil_div(Np, P.BN);
  int sms = 148, dev = 0;
  PYGB_CUDA(cudaGetDevice(&dev));
  PYGB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  P.split_acc = 0;
  P.acc_bufs = 2 * P.BN <= 512 ? 2 : 1;
  int tc = 32;
  while (tc < P.acc_bufs * P.BN) tc <<= 1;
  P.tmem_cols = tc;
  P.m_tiles = (int)ceil_div(M, kBM);
  const size_t stage_bytes = (size_t)(kBM + P.BN) * bk * 2;
  int stages = (int)((226 * 1024 - 1024 - kEpiBytes) / stage_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  const int nk = (int)ceil_div(K, bk);
  if (stages > nk) stages = nk < 2 ? 2 : nk;
  P.stages = stages;
  const size_t smem = stages * stage_bytes + kEpiBytes + 1024;
  CUtensorMap mapA, mapB;
  int rc = make_map(&mapA, A, M, K, lda, kBM, bk, false, 2);
  if (rc) return rc;
  rc = make_map(&mapB, Wt, Np, Kp, Kp, P.BN, bk, false, 2);
  if (rc) return rc;
  const int64_t tiles = (int64_t)P.m_tiles * P.n_tiles;
  dim3 grid((unsigned)(tiles < sms ? tiles : sms));
  if (bk == 64) {
    PYGB_CUDA(cudaFuncSetAttribute(gemm_tf32_kernel<1, 1, 64, 1>,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(226 * 1024)));
    gemm_tf32_kernel<1, 1, 64, 1><<<grid, kGemmThreads, smem, st>>>(mapA, mapB, mapB, P);
  } else {
    PYGB_CUDA(cudaFuncSetAttribute(gemm_tf32_kernel<1, 1, 32, 1>,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(226 * 1024)));
    gemm_tf32_kernel<1, 1, 32, 1><<<grid, kGemmThreads, smem, st>>>(mapA, mapB, mapB, P);
  }
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

int pygb_gemm_xw(const float* A, int64_t lda, int64_t a_slices, int64_t a_plane, int64_t M,
                 int64_t K, const float* W, int64_t ldw, int64_t N, const float* bias, int relu,
                 float* C, int64_t ldc, int passes, void* ws, size_t ws_bytes, pygb_stream_t s) {
  PYGB_REQUIRE(C || M == 0, "pygb_gemm_xw: null pointer");
  return gemm_xw_impl(A, lda, a_slices, a_plane, M, K, W, ldw, N, bias, relu, C, ldc, nullptr,
                      passes, ws, ws_bytes, s);
}

int pygb_gemm_xw_routed(const float* A, int64_t lda, int64_t a_slices, int64_t a_plane, int64_t M,
                        int64_t K, const float* W, int64_t ldw, int64_t N, const float* bias,
                        int relu,
                        float* const* dst_tab, int64_t Fc, int32_t n_dup, int64_t dst_row0,
                        int64_t ldd, int passes, int64_t max_ctas, void* ws, size_t ws_bytes,
                        pygb_stream_t s) {
  PYGB_REQUIRE(dst_tab && Fc > 0 && Fc % 4 == 0 && N % Fc == 0 && n_dup >= 1 && n_dup <= kMaxDup &&
                   dst_row0 >= 0 && ldd >= Fc && ldd % 4 == 0,
               "pygb_gemm_xw_routed: N must be a whole number of column blocks of Fc (multiple "
               "of 4) columns, 1 <= n_dup <= %d", kMaxDup);
  GemmRoute r{dst_tab, Fc, n_dup, dst_row0, ldd, max_ctas};
  return gemm_xw_impl(A, lda, a_slices, a_plane, M, K, W, ldw, N, bias, relu, nullptr, N, &r,
                      passes, ws, ws_bytes, s);
}

}  // extern "C"
