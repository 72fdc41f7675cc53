// K2b v2 — streaming fused GATConv backward (autograd of SURVEY A.10 / A.7, row a15):
// two passes on the skeleton of gat_stream.cu, neither of which writes or reads an
// [E', H] tensor.  With  s_ij = a_dst[i,h] + a_src[j,h],  l_ij = leaky_relu(s_ij),
// alpha_ij = exp(l_ij - max_i) / (sum_i + 1e-16)  and  out_i = sum_j alpha_ij xw_j :
//
//   d alpha_ij = <g_i, xw_j>                       (per head)
//   d l_ij     = alpha_ij (d alpha_ij - r_i),      r_i = sum_k alpha_ik d alpha_ik = <g_i, out_i>
//   d s_ij     = d l_ij * (s_ij > 0 ? 1 : slope)
//   d a_dst[i] = sum_j d s_ij       d a_src[j] = sum_i d s_ij       d xw_j = sum_i alpha_ij g_i
//
// r_i needs no pass over the edges (out is saved by the forward), so alpha and ds can be
// RECOMPUTED wherever they are needed from per-node records:
//   rec[i,h] = (a_dst, row_max * log2(e), 1 / (row_sum + 1e-16), r)      (pygb_gat_backward_prep)
//
//   mode 0 (destination-sorted walk): row-resident g_i and rec[i]; gathers xw_j and
//           a_src[j]; accumulates d a_dst[i].
//   mode 1 (source-sorted walk): row-resident xw_j and a_src[j]; gathers g_i and rec[i];
//           accumulates d xw_j and d a_src[j].
//
// Sums run in slot order; rows cut by a sub-item boundary leave partial sums that the
// fix-up kernel adds in slot order — no float atomics, bit-identical from run to run.
#include <float.h>
#include <math.h>

#include <mutex>
#include <type_traits>

#include "stream_common.cuh"

namespace pygb {
namespace {

constexpr int kModeDst = 0, kModeSrc = 1;
constexpr float kLog2eB = 1.4426950408889634f;

struct GatBwdParams {
  const float* row_feat;  // mode 0: grad_out [n_rows, F]; mode 1: xw [n_rows, F]
  int64_t ld_row;
  const float* gat_feat;  // mode 0: xw; mode 1: grad_out
  int64_t ld_gat;
  const float* a_src;     // [n_src, H]
  const float4* rec;      // [n_dst, H]
  int F, H, C, LH;        // LH = C / 4 lanes per head (power of two)
  float slope;
  const int* rowptr;
  const int* idxf;
  int64_t n_rows;
  int E;
  const int* sub_info;
  const int* nz_rows;
  int Ts;
  int n_sub;
  float* d_feat;  // mode 1: [n_rows, ldd]
  int64_t ldd;
  float* d_a;     // [n_rows, H]
  float* partial; // [2*n_sub][pstride]
  int* tail_row;
  int Fp;         // F in mode 1, 0 in mode 0
};

template <int MODE, int LPF, int VPL>
struct BwdCfg {
  static constexpr int EPW = 32 / LPF;
  static constexpr int NSTAGE = 4;
  static constexpr int G = VPL >= 4 ? 1 : (VPL >= 2 ? 2 : 4);
  static constexpr int SIDE_SZ = MODE == kModeSrc ? 16 : 4;   // rec (float4) or a_src (float)
  static constexpr int DATA_B = VPL * 512;
  static constexpr int SIDE_B = VPL * 32 * SIDE_SZ;
  static constexpr int META_B = (EPW * G * 4 + 15) / 16 * 16;
  static constexpr int STAGE_B = G * (DATA_B + SIDE_B) + META_B;
  static constexpr int WARP_B = NSTAGE * STAGE_B;
  static constexpr int CTA_B = WARP_B * kStreamWarps;
};

__device__ __forceinline__ float ex2f_b(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float lds_f32_b(uint32_t a) {
  float r;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(r) : "r"(a));
  return r;
}

// LHT = lanes per head known at compile time (1, 2, 4) or 0: read P.LH.
template <int MODE, int LPF, int VPL, bool EXACT, int LHT>
__global__ void __launch_bounds__(kStreamWarps * 32)
gat_bwd_stream_kernel(const GatBwdParams P) {
  using Cf = BwdCfg<MODE, LPF, VPL>;
  constexpr int EPW = Cf::EPW, KV = 4, G = Cf::G, NSTAGE = Cf::NSTAGE;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int fl = lane & (LPF - 1), es = lane / LPF;
  const int F = P.F, H = P.H, LH = LHT ? LHT : P.LH;
  // mode 1 stages ONE 16-byte record per (slot, head), copied by the first lane of the
  // head and read by all its lanes; mode 0 stages a_src[j, head] per lane (4 bytes)
  const int side_b = MODE == kModeSrc ? EPW * H * 16 : Cf::SIDE_B;
  const int stage_b = G * (Cf::DATA_B + side_b) + Cf::META_B;
  const uint32_t wbase =
      (uint32_t)__cvta_generic_to_shared(smem_raw) + (uint32_t)(warp * NSTAGE * stage_b);
  const uint32_t data0 = wbase + (uint32_t)(lane * 16);
  const uint32_t meta0 = wbase + (uint32_t)(G * (Cf::DATA_B + side_b) + es * G * 4);
  const int units = F / KV;
  bool act[VPL], lead_h[VPL];
  int head[VPL];
  uint32_t side0[VPL];
  const char* sideg[VPL];  // global base of the staged side operand of unit v
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    act[v] = EXACT || (fl + v * LPF < units);
    head[v] = act[v] ? ((fl + v * LPF) * KV) / P.C : 0;
    lead_h[v] = act[v] && ((fl + v * LPF) & (LH - 1)) == 0;
    if (MODE == kModeSrc) {
      side0[v] = wbase + (uint32_t)(G * Cf::DATA_B + (es * H + head[v]) * 16);
      sideg[v] = reinterpret_cast<const char*>(P.rec) + head[v] * 16;
    } else {
      side0[v] = wbase + (uint32_t)(G * Cf::DATA_B + lane * 4 + v * 128);
      sideg[v] = reinterpret_cast<const char*>(P.a_src) + head[v] * 4;
    }
  }
  const uint32_t ldside = (uint32_t)(MODE == kModeSrc ? H * 16 : H * 4);
  const int W = gridDim.x * kStreamWarps;
  const char* gb = reinterpret_cast<const char*>(P.gat_feat) + fl * 16;
  asm volatile("" : "+l"(gb));
  const uint32_t ldgb = (uint32_t)(P.ld_gat * sizeof(float));
  const int* __restrict__ idxf = P.idxf;
  const float* __restrict__ a_src = P.a_src;
  const float4* __restrict__ rec = P.rec;
  const int* __restrict__ nz_rows = P.nz_rows;
  const int E = P.E, Ts = P.Ts;
  const int n_items = (P.n_sub + EPW - 1) / EPW;
  const int grp_lane0 = es * LPF;
  const bool lead = fl == 0;
  const float slope = P.slope;
  const int pstride = (P.Fp + H + 3) & ~3;

  // accumulators of the current row
  float acc[VPL][KV], da[VPL];
  // row-resident operands (current row, next row prefetched)
  float rf[VPL][KV], rs[VPL][4], rf_n[VPL][KV], rs_n[VPL][4];
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    da[v] = 0.f;
#pragma unroll
    for (int i = 0; i < KV; ++i) acc[v][i] = 0.f;
  }

  auto load_row = [&](int row, float (&f)[VPL][KV], float (&s)[VPL][4]) {
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
      float4 t = make_float4(0.f, 0.f, 0.f, 0.f), r = make_float4(0.f, 0.f, 0.f, 0.f);
      if (act[v]) {
        t = __ldg(reinterpret_cast<const float4*>(P.row_feat + (int64_t)row * P.ld_row +
                                                  (fl + v * LPF) * KV));
        if (MODE == kModeDst) r = __ldg(rec + (int64_t)row * H + head[v]);
        else r.x = __ldg(a_src + (int64_t)row * H + head[v]);
      }
      f[v][0] = t.x, f[v][1] = t.y, f[v][2] = t.z, f[v][3] = t.w;
      s[v][0] = r.x, s[v][1] = r.y, s[v][2] = r.z, s[v][3] = r.w;
    }
  };

  for (int item = blockIdx.x + gridDim.x * warp; item < n_items; item += W) {
    const int s = item * EPW + es;
    const bool live = s < P.n_sub;
    const int pg0 = live ? s * Ts : 0;
    const int n_g = live ? min(Ts, E - pg0) : 0;
    const int info = live ? __ldg(P.sub_info + s) : 0;
    int rank = info & 0x7fffffff;
    bool left_open = info < 0;
    int row_cur = live ? __ldg(nz_rows + rank) : 0;
    int row_nxt = live ? __ldg(nz_rows + rank + 1) : 0;
    load_row(row_cur, rf, rs);
    load_row(row_nxt, rf_n, rs_n);
    int kdone = 0;
    const int* __restrict__ my_idx = idxf + pg0 + fl;
    int idx_cur = (fl < n_g) ? __ldg(my_idx) : 0;
    int idx_nxt = (LPF + fl < n_g) ? __ldg(my_idx + LPF) : 0;

    auto write_piece = [&](float* pr) {
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        if (!act[v]) continue;
        const int f0 = (fl + v * LPF) * KV;
        if (MODE == kModeSrc) store_partial<KV>(acc[v], pr + f0);
        if (f0 % P.C == 0) pr[P.Fp + head[v]] = da[v];
      }
    };
    auto reset = [&]() {
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        da[v] = 0.f;
#pragma unroll
        for (int i = 0; i < KV; ++i) acc[v][i] = 0.f;
      }
    };
    auto flush = [&](int k) {
      if (!left_open) {
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
          if (!act[v]) continue;
          const int f0 = (fl + v * LPF) * KV;
          if (MODE == kModeSrc)
            *reinterpret_cast<float4*>(P.d_feat + (int64_t)row_cur * P.ldd + f0) =
                make_float4(acc[v][0], acc[v][1], acc[v][2], acc[v][3]);
          if (f0 % P.C == 0) P.d_a[(int64_t)row_cur * H + head[v]] = da[v];
        }
      } else {
        write_piece(P.partial + (int64_t)(2 * s) * pstride);
      }
      reset();
      left_open = false;
      kdone = k + 1;
      ++rank;
      row_cur = row_nxt;
      row_nxt = __ldg(nz_rows + rank + 1);
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
#pragma unroll
        for (int i = 0; i < KV; ++i) rf[v][i] = rf_n[v][i], rs[v][i] = rs_n[v][i];
      }
      load_row(row_nxt, rf_n, rs_n);
    };

    const bool all_full = (int64_t)(item * EPW + EPW) * Ts <= (int64_t)E;

    auto run = [&](auto full_tag) {
      constexpr bool FULL = decltype(full_tag)::value;
      const int nsteps = FULL ? Ts : __shfl_sync(kFull, n_g, 0);
      const int nstages = (nsteps + G - 1) / G;

      auto issue_stage = [&](int st) {
        const uint32_t so = (uint32_t)((st & (NSTAGE - 1)) * stage_b);
        const int k0 = st * G;
        if (st > 0 && (k0 & (LPF - 1)) == 0) {
          idx_cur = idx_nxt;
          idx_nxt = (FULL ? k0 + LPF < Ts : k0 + LPF + fl < n_g) ? __ldg(my_idx + k0 + LPF) : 0;
        }
        int craw[G];
        uint32_t c[G];
        const int lane0 = grp_lane0 + (k0 & (LPF - 1));
#pragma unroll
        for (int g = 0; g < G; ++g) {
          craw[g] = __shfl_sync(kFull, idx_cur, lane0 + g);
          c[g] = (uint32_t)(craw[g] & 0x7fffffff);
        }
#pragma unroll
        for (int g = 0; g < G; ++g) {
          if (FULL || k0 + g < n_g) {
#pragma unroll
            for (int v = 0; v < VPL; ++v) {
              if (act[v])
                cp_async16(data0 + so + (uint32_t)(g * Cf::DATA_B + v * 512),
                           gb + (uint64_t)c[g] * ldgb + v * LPF * 16);
              if (MODE == kModeSrc) {
                if (lead_h[v])
                  cp_async16(side0[v] + so + (uint32_t)(g * side_b),
                             sideg[v] + (uint64_t)c[g] * ldside);
              } else if (act[v]) {
                cp_async_small<4>(side0[v] + so + (uint32_t)(g * side_b),
                                  sideg[v] + (uint64_t)c[g] * ldside);
              }
            }
          }
        }
        if (lead) sts_words<G>(meta0 + so, craw);
        cp_async_commit();
      };

      auto consume_stage = [&](int st) {
        const uint32_t so = (uint32_t)((st & (NSTAGE - 1)) * stage_b);
        const int k0 = st * G;
        uint32_t flags[G];
        lds_words<G>(meta0 + so, flags);
#pragma unroll
        for (int g = 0; g < G; ++g) {
          const int k = k0 + g;
          const bool valid = FULL || k < n_g;
#pragma unroll
          for (int v = 0; v < VPL; ++v) {
            const bool on = valid && act[v];
            float f[KV] = {0.f, 0.f, 0.f, 0.f};
            float alpha = 0.f, dot = 0.f, dlr = 0.f, rdot = 0.f;
            if (on) {
              const uint4 qv = lds128(data0 + so + (uint32_t)(g * Cf::DATA_B + v * 512));
              f[0] = __uint_as_float(qv.x), f[1] = __uint_as_float(qv.y);
              f[2] = __uint_as_float(qv.z), f[3] = __uint_as_float(qv.w);
              float sv, m2, inv;
              if (MODE == kModeSrc) {
                const uint4 rq = lds128(side0[v] + so + (uint32_t)(g * side_b));
                sv = __uint_as_float(rq.x) + rs[v][0];
                m2 = __uint_as_float(rq.y), inv = __uint_as_float(rq.z), rdot = __uint_as_float(rq.w);
              } else {
                sv = rs[v][0] + lds_f32_b(side0[v] + so + (uint32_t)(g * side_b));
                m2 = rs[v][1], inv = rs[v][2], rdot = rs[v][3];
              }
              const bool pos = sv > 0.f;
              const float l = pos ? sv : sv * slope;
              dlr = pos ? 1.f : slope;
              alpha = ex2f_b(fmaf(l, kLog2eB, -m2)) * inv;
              dot = fmaf(f[3], rf[v][3], fmaf(f[2], rf[v][2], fmaf(f[1], rf[v][1], f[0] * rf[v][0])));
            }
            // <g_i, xw_j> over the C channels of the head: LH adjacent lanes
            if (LHT) {
#pragma unroll
              for (int o = 1; o < LHT; o <<= 1) dot += __shfl_xor_sync(kFull, dot, o);
            } else {
              for (int o = 1; o < LH; o <<= 1) dot += __shfl_xor_sync(kFull, dot, o);
            }
            if (on) {
              da[v] = fmaf(alpha * dlr, dot - rdot, da[v]);
              if (MODE == kModeSrc) {
#pragma unroll
                for (int i = 0; i < KV; ++i) acc[v][i] = fmaf(alpha, f[i], acc[v][i]);
              }
            }
          }
          if (valid && (int)flags[g] < 0) flush(k);
        }
      };

#pragma unroll 1
      for (int st = 0; st < NSTAGE - 1; ++st) {
        if (st < nstages) issue_stage(st);
        else cp_async_commit();
      }
#pragma unroll 1
      for (int st = 0; st < nstages; ++st) {
        if (st + NSTAGE - 1 < nstages) issue_stage(st + NSTAGE - 1);
        else cp_async_commit();
        cp_async_wait<NSTAGE - 1>();
        __syncwarp();
        consume_stage(st);
        __syncwarp();
      }
    };
    if (all_full) run(std::true_type{});
    else run(std::false_type{});

    if (kdone < n_g) {  // the row continues in the next sub-item
      write_piece(P.partial + (int64_t)(2 * s + (left_open ? 0 : 1)) * pstride);
      if (lead && !left_open) P.tail_row[s] = row_cur;
      reset();
    }
  }
  cp_async_wait<0>();
}

// Adds the pieces of rows cut by sub-item boundaries (slot order) and zeroes the rows
// that have no slot.
template <int MODE>
__global__ void __launch_bounds__(256)
gat_bwd_fixup_kernel(const GatBwdParams P) {
  const int lane = threadIdx.x & 31;
  const int gw = (blockIdx.x * 256 + threadIdx.x) >> 5;
  const int nw = (gridDim.x * 256) >> 5;
  const int F = P.F, H = P.H;
  const int pstride = (P.Fp + H + 3) & ~3;
  for (int c = gw; c < P.n_sub; c += nw) {
    const int row = P.tail_row[c];
    if (row < 0) continue;
    const int end = P.rowptr[row + 1];
    const int c_last = (end - 1) / P.Ts;
    if (MODE == kModeSrc) {
      for (int f = lane; f < F; f += 32) {
        float A = P.partial[(int64_t)(2 * c + 1) * pstride + f];
        for (int cc = c + 1; cc <= c_last; ++cc) A += P.partial[(int64_t)(2 * cc) * pstride + f];
        P.d_feat[(int64_t)row * P.ldd + f] = A;
      }
    }
    for (int h = lane; h < H; h += 32) {
      float D = P.partial[(int64_t)(2 * c + 1) * pstride + P.Fp + h];
      for (int cc = c + 1; cc <= c_last; ++cc) D += P.partial[(int64_t)(2 * cc) * pstride + P.Fp + h];
      P.d_a[(int64_t)row * H + h] = D;
    }
  }
  for (int64_t base = (int64_t)gw * 32; base < P.n_rows; base += (int64_t)nw * 32) {
    const int64_t row = base + lane;
    const bool empty = row < P.n_rows && P.rowptr[row] == P.rowptr[row + 1];
    unsigned mask = __ballot_sync(kFull, empty);
    while (mask) {
      const int l = __ffs(mask) - 1;
      mask &= mask - 1;
      if (MODE == kModeSrc)
        for (int f = lane; f < F; f += 32) P.d_feat[(base + l) * P.ldd + f] = 0.f;
      for (int h = lane; h < H; h += 32) P.d_a[(base + l) * H + h] = 0.f;
    }
  }
}

// rec[i,h] = (a_dst, row_max * log2 e, 1 / (row_sum + 1e-16), <g_i, out_i - bias>_h)
__global__ void __launch_bounds__(256)
gat_bwd_prep_kernel(const float* __restrict__ a_dst, const float* __restrict__ row_max,
                    const float* __restrict__ row_sum, const float* __restrict__ out, int64_t ldo,
                    const float* __restrict__ bias, const float* __restrict__ g, int64_t ldg,
                    int64_t n, int H, int C, float4* __restrict__ rec) {
  const int64_t total = n * H;
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
    const int64_t node = t / H;
    const int h = (int)(t - node * H);
    const float* gr = g + node * ldg + (int64_t)h * C;
    const float* orow = out + node * ldo + (int64_t)h * C;
    float r = 0.f;
    for (int c = 0; c < C; ++c) {
      float o = __ldg(orow + c);
      if (bias) o -= __ldg(bias + h * C + c);
      r = fmaf(__ldg(gr + c), o, r);
    }
    rec[t] = make_float4(a_dst ? __ldg(a_dst + t) : 0.f, __ldg(row_max + t) * kLog2eB,
                         1.f / (__ldg(row_sum + t) + 1e-16f), r);
  }
}

template <int MODE, int LPF, int VPL, bool EXACT, int LHT>
int launch_bwd_k(const GatBwdParams& P, cudaStream_t st) {
  using Cf = BwdCfg<MODE, LPF, VPL>;
  auto kern = gat_bwd_stream_kernel<MODE, LPF, VPL, EXACT, LHT>;
  // same geometry as the kernel: the side operand of mode 1 is H records per slot
  const int side_b = MODE == kModeSrc ? Cf::EPW * P.H * 16 : Cf::SIDE_B;
  const int cta_b = Cf::NSTAGE * (Cf::G * (Cf::DATA_B + side_b) + Cf::META_B) * kStreamWarps;
  static std::mutex mu;
  static int ctas_cache[64] = {0}, smem_cache[64] = {0};
  int dev = 0;
  PYGB_CUDA(cudaGetDevice(&dev));
  int ctas_per_sm = 0;
  cudaError_t attr_err = cudaSuccess;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (dev >= 0 && dev < 64 && smem_cache[dev] == cta_b) ctas_per_sm = ctas_cache[dev];
    if (ctas_per_sm == 0) {
      attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cf::CTA_B);
      if (attr_err == cudaSuccess)
        attr_err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, kern,
                                                                 kStreamWarps * 32, cta_b);
      if (ctas_per_sm < 1) ctas_per_sm = 1;
      if (attr_err == cudaSuccess && dev >= 0 && dev < 64) {
        ctas_cache[dev] = ctas_per_sm;
        smem_cache[dev] = cta_b;
      }
    }
  }
  PYGB_CUDA(attr_err);
  const int64_t n_items = ceil_div(P.n_sub, Cf::EPW);
  int64_t grid = (int64_t)sm_count() * ctas_per_sm;
  const int64_t need = ceil_div(n_items, kStreamWarps);
  if (grid > need) grid = need;
  if (grid < 1) grid = 1;
  kern<<<(unsigned)grid, kStreamWarps * 32, cta_b, st>>>(P);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

template <int MODE, int LPF, int VPL, bool EXACT>
int launch_bwd_lh(const GatBwdParams& P, cudaStream_t st) {
  switch (P.LH) {  // C = 4, 8, 16 get the unrolled head reduction
    case 1: return launch_bwd_k<MODE, LPF, VPL, EXACT, 1>(P, st);
    case 2: return launch_bwd_k<MODE, LPF, VPL, EXACT, 2>(P, st);
    case 4: return launch_bwd_k<MODE, LPF, VPL, EXACT, 4>(P, st);
    default: return launch_bwd_k<MODE, LPF, VPL, EXACT, 0>(P, st);
  }
}

template <int MODE, int LPF, int VPL>
int launch_bwd_one(const GatBwdParams& P, cudaStream_t st) {
  if (P.F / 4 == LPF * VPL) return launch_bwd_lh<MODE, LPF, VPL, true>(P, st);
  return launch_bwd_lh<MODE, LPF, VPL, false>(P, st);
}

template <int MODE>
int launch_bwd(const GatBwdParams& P, cudaStream_t st) {
  const int units = P.F / 4;
  int rc;
  if (units <= 4) rc = launch_bwd_one<MODE, 4, 1>(P, st);
  else if (units <= 8) rc = launch_bwd_one<MODE, 8, 1>(P, st);
  else if (units <= 16) rc = launch_bwd_one<MODE, 16, 1>(P, st);
  else if (units <= 32) rc = launch_bwd_one<MODE, 32, 1>(P, st);
  else if (units <= 64) rc = launch_bwd_one<MODE, 32, 2>(P, st);
  else rc = launch_bwd_one<MODE, 32, 4>(P, st);
  if (rc) return rc;
  int64_t work = (int64_t)P.n_sub * 32;
  if (work < P.n_rows) work = P.n_rows;
  gat_bwd_fixup_kernel<MODE><<<grid1d(work), 256, 0, st>>>(P);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

inline size_t bwd_pstride(int mode, int H, int C) {
  const int64_t Fp = mode == kModeSrc ? (int64_t)H * C : 0;
  return (size_t)((Fp + H + 3) & ~3);
}

}  // namespace
}  // namespace pygb

using namespace pygb;

extern "C" {

int pygb_gat_backward_prep(const float* a_dst, const float* row_max, const float* row_sum,
                           const float* out, int64_t ldo, const float* bias,
                           const float* grad_out, int64_t ldg, int64_t n_dst, int H, int C,
                           float* rec, pygb_stream_t s) {
  PYGB_REQUIRE(n_dst >= 0 && H > 0 && C > 0 && ldo >= (int64_t)H * C && ldg >= (int64_t)H * C,
               "pygb_gat_backward_prep: bad sizes");
  if (n_dst == 0) return PYGB_OK;
  PYGB_REQUIRE(row_max && row_sum && out && grad_out && rec && aligned16(rec),
               "pygb_gat_backward_prep: null or misaligned pointer");
  gat_bwd_prep_kernel<<<grid1d(n_dst * H), 256, 0, as_stream(s)>>>(
      a_dst, row_max, row_sum, out, ldo, bias, grad_out, ldg, n_dst, H, C,
      reinterpret_cast<float4*>(rec));
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

size_t pygb_gat_backward_stream_workspace(int mode, int64_t n_sub, int H, int C) {
  return align_up((size_t)(2 * n_sub) * bwd_pstride(mode, H, C) * sizeof(float), 256) +
         align_up((size_t)n_sub * sizeof(int), 256) + 256;
}

int pygb_gat_backward_stream(int mode, const float* row_feat, int64_t ld_row,
                             const float* gat_feat, int64_t ld_gat, const float* a_src,
                             const float* rec, int H, int C, float negative_slope,
                             const int32_t* rowptr, const int32_t* idx_flagged, int64_t n_rows,
                             int64_t E, const int32_t* nz_rows, const int32_t* sub_info,
                             int32_t Ts, int64_t n_sub, float* d_feat, int64_t ldd, float* d_a,
                             void* ws, size_t ws_bytes, pygb_stream_t s) {
  PYGB_REQUIRE(mode == kModeDst || mode == kModeSrc, "pygb_gat_backward_stream: bad mode");
  PYGB_REQUIRE(H > 0 && C > 0 && n_rows >= 0 && E >= 0 && E < (1ll << 31) - 1024 &&
                   n_rows < (1ll << 31) - 64,
               "pygb_gat_backward_stream: bad sizes");
  const int64_t F = (int64_t)H * C;
  const int LH = C / 4;
  if (C % 4 != 0 || (LH & (LH - 1)) != 0 || LH > 32 || F / 4 > 128) {
    set_error("pygb_gat_backward_stream: needs C in {4,8,16,32,64,128} and H*C <= 512");
    return PYGB_E_UNSUPPORTED;
  }
  PYGB_REQUIRE(Ts >= 32 && (Ts & (Ts - 1)) == 0 && n_sub == ceil_div(E, Ts),
               "pygb_gat_backward_stream: bad plan");
  if (n_rows == 0) return PYGB_OK;
  PYGB_REQUIRE(rowptr && d_a && a_src && rec && row_feat && (mode == kModeDst || d_feat) &&
                   (E == 0 || (gat_feat && idx_flagged && nz_rows && sub_info)),
               "pygb_gat_backward_stream: null pointer");
  PYGB_REQUIRE(aligned16(row_feat) && aligned16(gat_feat) && aligned16(rec) && ld_row % 4 == 0 &&
                   ld_gat % 4 == 0 && ld_row >= F && ld_gat >= F && ld_gat * 4 < (1ll << 31) &&
                   (mode == kModeDst || (aligned16(d_feat) && ldd % 4 == 0 && ldd >= F)),
               "pygb_gat_backward_stream: feature rows must be 16-byte aligned");
  if (ws_bytes < pygb_gat_backward_stream_workspace(mode, n_sub, H, C)) {
    set_error("pygb_gat_backward_stream: workspace too small");
    return PYGB_E_WORKSPACE;
  }
  cudaStream_t st = as_stream(s);
  Workspace Wk(ws, ws_bytes);
  GatBwdParams P;
  P.row_feat = row_feat;
  P.ld_row = ld_row;
  P.gat_feat = gat_feat;
  P.ld_gat = ld_gat;
  P.a_src = a_src;
  P.rec = reinterpret_cast<const float4*>(rec);
  P.F = (int)F;
  P.H = H;
  P.C = C;
  P.LH = LH;
  P.slope = negative_slope;
  P.rowptr = rowptr;
  P.idxf = idx_flagged;
  P.n_rows = n_rows;
  P.E = (int)E;
  P.sub_info = sub_info;
  P.nz_rows = nz_rows;
  P.Ts = Ts;
  P.n_sub = (int)n_sub;
  P.d_feat = d_feat;
  P.ldd = ldd;
  P.d_a = d_a;
  P.Fp = mode == kModeSrc ? (int)F : 0;
  P.partial = Wk.take<float>((size_t)(2 * n_sub) * bwd_pstride(mode, H, C));
  P.tail_row = Wk.take<int>((size_t)n_sub);
  if (n_sub > 0) PYGB_CUDA(cudaMemsetAsync(P.tail_row, 0xff, (size_t)n_sub * sizeof(int), st));
  return mode == kModeDst ? launch_bwd<kModeDst>(P, st) : launch_bwd<kModeSrc>(P, st);
}

}  // extern "C"
