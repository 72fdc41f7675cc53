// K2 v2 — streaming fused GATConv propagate: leaky-relu logits,
// per-destination softmax and the alpha-weighted sum in ONE pass over the
// destination-sorted CSR, on the same skeleton as seg_stream.cu (equal-slot
// sub-items per lane group, flagged index stream, cp.async-staged ring).
//
// The logit splits as a_dst[i,h] + a_src[j,h].  Each step stages the gathered
// xw[j] row and the 4-byte a_src[j, head] of every 16-byte unit through shared
// memory; the consumer keeps an ONLINE softmax per (row, head): running maximum
// m, running denominator s and the accumulator, rescaled by exp(m - m') when the
// maximum moves.  The result equals the reference's max-subtracted softmax up to
// fp32 rounding (~1e-7), and no [E', ...] tensor is ever written.  Rows cut by a
// sub-item boundary leave (acc, m, s) pieces that the fix-up kernel merges in
// slot order — deterministic, no atomics.
//
// cfg3 (Reddit-shaped, 8 heads x 8): xw is 60 MB and stays in L2, so the kernel
// is bound by L2->SM bandwidth and issue slots, not HBM.
#include <float.h>
#include <math.h>

#include <mutex>
#include <type_traits>

#include "stream_common.cuh"

namespace pygb {
namespace {

struct GatParams {
  const void* xw;
  int64_t ldx;
  int F, H, C;
  const float* a_src;
  const float* a_dst;  // nullable
  float slope;
  const int* rowptr;
  const int* idxf;
  int64_t n_rows;
  int E;
  const float* bias;
  void* out;
  int64_t ldo;
  float* row_max;  // nullable
  float* row_sum;  // nullable
  const int* sub_info;
  const int* nz_rows;
  int Ts;
  int n_sub;
  float* partial;  // [2*n_sub][Fp + 2*H]
  int* tail_row;
  int Fp;
};

template <typename T, int LPF, int VPL>
struct GatCfg {
  static constexpr int EPW = 32 / LPF;
  static constexpr int KV = Elem<T>::kVec;
  static constexpr int NSTAGE = 4;
  static constexpr int G = VPL >= 4 ? 1 : (VPL >= 2 ? 2 : 4);
  static constexpr int DATA_B = VPL * 512;  // gathered rows, one step
  static constexpr int ASRC_B = VPL * 128;  // a_src per (lane, v), one step
  static constexpr int META_B = (EPW * G * 4 + 15) / 16 * 16;
  static constexpr int STAGE_B = G * (DATA_B + ASRC_B) + META_B;
  static constexpr int WARP_B = NSTAGE * STAGE_B;
  static constexpr int CTA_B = WARP_B * kStreamWarps;
};

__device__ __forceinline__ float leaky_relu(float x, float slope) { return x > 0.f ? x : x * slope; }
__device__ __forceinline__ float ex2f(float x) {  // 2^x, flush-to-zero, one MUFU
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
__device__ __forceinline__ float lds_f32(uint32_t a) {
  float r;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(r) : "r"(a));
  return r;
}

// The running maximum m is kept in the log2 domain (logit * log2(e)) so that every
// exponential is a single ex2.approx; row_max is converted back on output.
template <typename T, int LPF, int VPL, bool EXACT>
__global__ void __launch_bounds__(kStreamWarps * 32)
gat_stream_kernel(const GatParams P) {
  using Cf = GatCfg<T, LPF, VPL>;
  constexpr int EPW = Cf::EPW, KV = Cf::KV, G = Cf::G, NSTAGE = Cf::NSTAGE;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int fl = lane & (LPF - 1), es = lane / LPF;
  const uint32_t wbase = (uint32_t)__cvta_generic_to_shared(smem_raw) + (uint32_t)(warp * Cf::WARP_B);
  const uint32_t data0 = wbase + (uint32_t)(lane * 16);
  const uint32_t asrc0 = wbase + (uint32_t)(G * Cf::DATA_B + lane * 4);
  const uint32_t meta0 = wbase + (uint32_t)(G * (Cf::DATA_B + Cf::ASRC_B) + es * G * 4);
  const int F = P.F, H = P.H;
  const int units = F / KV;  // launcher guarantees C % KV == 0, hence F % KV == 0
  bool act[VPL];
  int head[VPL];
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    act[v] = EXACT || (fl + v * LPF < units);
    head[v] = act[v] ? ((fl + v * LPF) * KV) / P.C : 0;
  }
  const int W = gridDim.x * kStreamWarps;
  const char* xb = static_cast<const char*>(P.xw) + fl * 16;
  asm volatile("" : "+l"(xb));
  const uint32_t ldxb = (uint32_t)(P.ldx * sizeof(T));
  const int* __restrict__ idxf = P.idxf;
  const float* __restrict__ a_src = P.a_src;
  const float* __restrict__ a_dst = P.a_dst;
  const int* __restrict__ nz_rows = P.nz_rows;
  const int E = P.E, Ts = P.Ts;
  const int n_items = (P.n_sub + EPW - 1) / EPW;
  const int grp_lane0 = es * LPF;
  const bool lead = fl == 0;
  const float slope = P.slope;
  const int pstride = (P.Fp + 2 * H + 3) & ~3;  // pieces stay 16-byte aligned

  float acc[VPL][KV], m[VPL], ssum[VPL];
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    m[v] = -FLT_MAX;
    ssum[v] = 0.f;
#pragma unroll
    for (int i = 0; i < KV; ++i) acc[v][i] = 0.f;
  }

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
    float adst[VPL], adst_nxt[VPL];
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
      adst[v] = (live && a_dst) ? __ldg(a_dst + (int64_t)row_cur * H + head[v]) : 0.f;
      adst_nxt[v] = (live && a_dst) ? __ldg(a_dst + (int64_t)row_nxt * H + head[v]) : 0.f;
    }
    int kdone = 0;  // slots already closed by a flush
    const int nsteps = __shfl_sync(kFull, n_g, 0);
    const int nstages = (nsteps + G - 1) / G;
    const int* __restrict__ my_idx = idxf + pg0 + fl;
    int idx_cur = (fl < n_g) ? __ldg(my_idx) : 0;
    int idx_nxt = (LPF + fl < n_g) ? __ldg(my_idx + LPF) : 0;

    auto write_piece = [&](float* pr) {  // (acc, m, s) of an unfinished row
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        if (!act[v]) continue;
        const int f0 = (fl + v * LPF) * KV;
        store_partial<KV>(acc[v], pr + f0);
        if (f0 % P.C == 0) {
          pr[P.Fp + head[v]] = m[v];  // log2 domain
          pr[P.Fp + H + head[v]] = ssum[v];
        }
      }
    };
    auto reset = [&]() {
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        m[v] = -FLT_MAX;
        ssum[v] = 0.f;
#pragma unroll
        for (int i = 0; i < KV; ++i) acc[v][i] = 0.f;
      }
    };
    auto flush = [&](int k) {
      if (!left_open) {
        T* orow = static_cast<T*>(P.out) + (int64_t)row_cur * P.ldo;
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
          if (!act[v]) continue;
          const int f0 = (fl + v * LPF) * KV;
          const float inv = 1.f / (ssum[v] + 1e-16f);
          float r[KV];
#pragma unroll
          for (int i = 0; i < KV; ++i) r[i] = acc[v][i] * inv;
          store_row<T, KV>(r, f0, F, 1.f, 0, P.bias, orow);
          if (f0 % P.C == 0) {
            if (P.row_max) P.row_max[(int64_t)row_cur * H + head[v]] = m[v] * kLn2;
            if (P.row_sum) P.row_sum[(int64_t)row_cur * H + head[v]] = ssum[v];
          }
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
        adst[v] = adst_nxt[v];
        adst_nxt[v] = a_dst ? __ldg(a_dst + (int64_t)row_nxt * H + head[v]) : 0.f;
      }
    };

    const bool all_full = (int64_t)(item * EPW + EPW) * Ts <= (int64_t)E;

    auto run = [&](auto full_tag) {
      constexpr bool FULL = decltype(full_tag)::value;
      const int nsteps = FULL ? Ts : __shfl_sync(kFull, n_g, 0);
      const int nstages = (nsteps + G - 1) / G;

      auto issue_stage = [&](int st) {
        const uint32_t so = (uint32_t)((st & (NSTAGE - 1)) * Cf::STAGE_B);
        const int k0 = st * G;
        if (st > 0 && (k0 & (LPF - 1)) == 0) {
          idx_cur = idx_nxt;
          idx_nxt = (FULL ? k0 + LPF < Ts : k0 + LPF + fl < n_g) ? __ldg(my_idx + k0 + LPF) : 0;
        }
        int craw[G];
        const char* src[G];
        const float* as[G];
        const int lane0 = grp_lane0 + (k0 & (LPF - 1));
#pragma unroll
        for (int g = 0; g < G; ++g) {
          craw[g] = __shfl_sync(kFull, idx_cur, lane0 + g);
          const uint32_t c = (uint32_t)(craw[g] & 0x7fffffff);
          src[g] = xb + (uint64_t)c * ldxb;
          as[g] = a_src + (uint64_t)c * H;
        }
#pragma unroll
        for (int g = 0; g < G; ++g) {
          if (FULL || k0 + g < n_g) {
#pragma unroll
            for (int v = 0; v < VPL; ++v) {
              if (act[v]) {
                cp_async16(data0 + so + (uint32_t)(g * Cf::DATA_B + v * 512), src[g] + v * LPF * 16);
                cp_async_small<4>(asrc0 + so + (uint32_t)(g * Cf::ASRC_B + v * 128),
                                  as[g] + head[v]);
              }
            }
          }
        }
        if (lead) sts_words<G>(meta0 + so, craw);
        cp_async_commit();
      };

      auto consume_stage = [&](int st) {
        const uint32_t so = (uint32_t)((st & (NSTAGE - 1)) * Cf::STAGE_B);
        const int k0 = st * G;
        uint32_t flags[G];
        lds_words<G>(meta0 + so, flags);
#pragma unroll
        for (int g = 0; g < G; ++g) {
          const int k = k0 + g;
          const bool valid = FULL || k < n_g;
#pragma unroll
          for (int v = 0; v < VPL; ++v) {
            float l2 = -FLT_MAX;
            uint4 qv = make_uint4(0, 0, 0, 0);
            if (valid && act[v]) {
              qv = lds128(data0 + so + (uint32_t)(g * Cf::DATA_B + v * 512));
              const float asv = lds_f32(asrc0 + so + (uint32_t)(g * Cf::ASRC_B + v * 128));
              l2 = leaky_relu(adst[v] + asv, slope) * kLog2e;
            }
            // the maximum rarely moves after the first few edges of a row: rescale only then
            if (__any_sync(kFull, l2 > m[v])) {
              const float mn = fmaxf(m[v], l2);
              const float sc = ex2f(m[v] - mn);
              m[v] = mn;
              ssum[v] *= sc;
#pragma unroll
              for (int i = 0; i < KV; ++i) acc[v][i] *= sc;
            }
            if (valid && act[v]) {
              const float p = ex2f(l2 - m[v]);
              ssum[v] += p;
              float f[KV];
              Elem<T>::unpack(qv, f);
#pragma unroll
              for (int i = 0; i < KV; ++i) acc[v][i] = fmaf(p, f[i], acc[v][i]);
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

// Merges the (acc, m, s) pieces of rows cut by sub-item boundaries and writes the
// rows that have no incoming edge (sum over nothing = 0, + bias).
template <typename T>
__global__ void __launch_bounds__(256)
gat_stream_fixup_kernel(const GatParams P) {
  const int lane = threadIdx.x & 31;
  const int gw = (blockIdx.x * 256 + threadIdx.x) >> 5;
  const int nw = (gridDim.x * 256) >> 5;
  const int F = P.F, H = P.H, C = P.C;
  const int pstride = (P.Fp + 2 * H + 3) & ~3;  // pieces stay 16-byte aligned
  T* out = static_cast<T*>(P.out);
  for (int c = gw; c < P.n_sub; c += nw) {
    const int row = P.tail_row[c];
    if (row < 0) continue;
    const int end = P.rowptr[row + 1];
    const int c_last = (end - 1) / P.Ts;
    for (int f = lane; f < F; f += 32) {
      const int h = f / C;
      // pass 1: the row maximum over the pieces
      float M = P.partial[(int64_t)(2 * c + 1) * pstride + P.Fp + h];
      for (int cc = c + 1; cc <= c_last; ++cc)
        M = fmaxf(M, P.partial[(int64_t)(2 * cc) * pstride + P.Fp + h]);
      // pass 2: rescaled sums, in slot order
      float S = 0.f, A = 0.f;
      for (int cc = c; cc <= c_last; ++cc) {
        const float* pr = P.partial + (int64_t)(cc == c ? 2 * c + 1 : 2 * cc) * pstride;
        const float sc = ex2f(pr[P.Fp + h] - M);
        S = fmaf(pr[P.Fp + H + h], sc, S);
        A = fmaf(pr[f], sc, A);
      }
      float r = A / (S + 1e-16f);
      if (P.bias) r += __ldg(P.bias + f);
      Elem<T>::store(out + (int64_t)row * P.ldo + f, r);
      if (f % C == 0) {
        if (P.row_max) P.row_max[(int64_t)row * H + h] = M * kLn2;
        if (P.row_sum) P.row_sum[(int64_t)row * H + h] = S;
      }
    }
  }
  for (int64_t base = (int64_t)gw * 32; base < P.n_rows; base += (int64_t)nw * 32) {
    const int64_t row = base + lane;
    const bool empty = row < P.n_rows && P.rowptr[row] == P.rowptr[row + 1];
    unsigned mask = __ballot_sync(kFull, empty);
    while (mask) {
      const int l = __ffs(mask) - 1;
      mask &= mask - 1;
      T* orow = out + (base + l) * P.ldo;
      for (int f = lane; f < F; f += 32)
        Elem<T>::store(orow + f, P.bias ? __ldg(P.bias + f) : 0.f);
      for (int h = lane; h < H; h += 32) {
        if (P.row_max) P.row_max[(base + l) * H + h] = -FLT_MAX;
        if (P.row_sum) P.row_sum[(base + l) * H + h] = 0.f;
      }
    }
  }
}

template <typename T, int LPF, int VPL, bool EXACT>
int launch_gat_k(const GatParams& P, cudaStream_t st) {
  using Cf = GatCfg<T, LPF, VPL>;
  auto kern = gat_stream_kernel<T, LPF, VPL, EXACT>;
  static std::mutex mu;
  static int ctas_cache[64] = {0};
  int dev = 0;
  PYGB_CUDA(cudaGetDevice(&dev));
  int ctas_per_sm = 0;
  cudaError_t attr_err = cudaSuccess;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (dev >= 0 && dev < 64) ctas_per_sm = ctas_cache[dev];
    if (ctas_per_sm == 0) {
      attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cf::CTA_B);
      if (attr_err == cudaSuccess)
        attr_err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, kern,
                                                                 kStreamWarps * 32, Cf::CTA_B);
      if (ctas_per_sm < 1) ctas_per_sm = 1;
      if (attr_err == cudaSuccess && dev >= 0 && dev < 64) ctas_cache[dev] = ctas_per_sm;
    }
  }
  PYGB_CUDA(attr_err);
  const int64_t n_items = ceil_div(P.n_sub, Cf::EPW);
  int64_t grid = (int64_t)sm_count() * ctas_per_sm;
  const int64_t need = ceil_div(n_items, kStreamWarps);
  if (grid > need) grid = need;
  if (grid < 1) grid = 1;
  kern<<<(unsigned)grid, kStreamWarps * 32, Cf::CTA_B, st>>>(P);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

template <typename T, int LPF, int VPL>
int launch_gat_one(const GatParams& P, cudaStream_t st) {
  if (P.F / Elem<T>::kVec == LPF * VPL) return launch_gat_k<T, LPF, VPL, true>(P, st);
  return launch_gat_k<T, LPF, VPL, false>(P, st);
}

template <typename T>
int launch_gat(const GatParams& P, cudaStream_t st) {
  constexpr int KV = Elem<T>::kVec;
  const int units = P.F / KV;
  int rc;
  if (units <= 4) rc = launch_gat_one<T, 4, 1>(P, st);
  else if (units <= 8) rc = launch_gat_one<T, 8, 1>(P, st);
  else if (units <= 16) rc = launch_gat_one<T, 16, 1>(P, st);
  else if (units <= 32) rc = launch_gat_one<T, 32, 1>(P, st);
  else if (units <= 64) rc = launch_gat_one<T, 32, 2>(P, st);
  else rc = launch_gat_one<T, 32, 4>(P, st);
  if (rc) return rc;
  int64_t work = (int64_t)P.n_sub * 32;
  if (work < P.n_rows) work = P.n_rows;
  gat_stream_fixup_kernel<T><<<grid1d(work), 256, 0, st>>>(P);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

}  // namespace
}  // namespace pygb

using namespace pygb;

extern "C" {

size_t pygb_gat_stream_workspace(int64_t n_sub, int H, int C) {
  const int64_t Fp = (int64_t)H * C;
  return align_up((size_t)(2 * n_sub) * (size_t)((Fp + 2 * H + 3) & ~3) * sizeof(float), 256) +
         align_up((size_t)n_sub * sizeof(int), 256) + 256;
}

int pygb_gat_forward_stream(int dtype, const void* xw, int64_t ldx, const float* a_src,
                            const float* a_dst, int H, int C, float negative_slope,
                            const int32_t* rowptr, const int32_t* idx_flagged, int64_t n_rows,
                            int64_t E, const int32_t* nz_rows, const int32_t* sub_info, int32_t Ts,
                            int64_t n_sub, const float* bias, void* out, int64_t ldo,
                            float* row_max, float* row_sum, void* ws, size_t ws_bytes,
                            pygb_stream_t s) {
  PYGB_REQUIRE(dtype == PYGB_F32 || dtype == PYGB_BF16, "pygb_gat_forward_stream: bad dtype");
  const int kv = dtype == PYGB_F32 ? 4 : 8;
  PYGB_REQUIRE(H > 0 && C > 0 && n_rows >= 0 && E >= 0 && E < (1ll << 31) - 1024 &&
                   n_rows < (1ll << 31) - 64 && ldo >= (int64_t)H * C,
               "pygb_gat_forward_stream: bad sizes");
  const int64_t F = (int64_t)H * C;
  if (C % kv != 0 || F / kv > 128) {
    set_error("pygb_gat_forward_stream: needs C %% %d == 0 and H*C <= %d", kv, 128 * kv);
    return PYGB_E_UNSUPPORTED;
  }
  PYGB_REQUIRE(Ts >= 32 && (Ts & (Ts - 1)) == 0 && n_sub == ceil_div(E, Ts),
               "pygb_gat_forward_stream: bad plan");
  if (n_rows == 0) return PYGB_OK;
  PYGB_REQUIRE(rowptr && out && a_src && (E == 0 || (xw && idx_flagged && nz_rows && sub_info)),
               "pygb_gat_forward_stream: null pointer");
  const int esz = dtype == PYGB_F32 ? 4 : 2;
  PYGB_REQUIRE(aligned16(xw) && aligned16(out) && ldx % kv == 0 && ldo % kv == 0 && ldx >= F &&
                   ldx * esz < (1ll << 31),
               "pygb_gat_forward_stream: xw/out rows must be 16-byte aligned");
  if (ws_bytes < pygb_gat_stream_workspace(n_sub, H, C)) {
    set_error("pygb_gat_forward_stream: workspace too small");
    return PYGB_E_WORKSPACE;
  }
  cudaStream_t st = as_stream(s);
  Workspace Wk(ws, ws_bytes);
  GatParams P;
  P.xw = xw;
  P.ldx = ldx;
  P.F = (int)F;
  P.H = H;
  P.C = C;
  P.a_src = a_src;
  P.a_dst = a_dst;
  P.slope = negative_slope;
  P.rowptr = rowptr;
  P.idxf = idx_flagged;
  P.n_rows = n_rows;
  P.E = (int)E;
  P.bias = bias;
  P.out = out;
  P.ldo = ldo;
  P.row_max = row_max;
  P.row_sum = row_sum;
  P.sub_info = sub_info;
  P.nz_rows = nz_rows;
  P.Ts = Ts;
  P.n_sub = (int)n_sub;
  P.Fp = (int)F;
  P.partial = Wk.take<float>((size_t)(2 * n_sub) * (size_t)((F + 2 * H + 3) & ~3));
  P.tail_row = Wk.take<int>((size_t)n_sub);
  if (n_sub > 0) PYGB_CUDA(cudaMemsetAsync(P.tail_row, 0xff, (size_t)n_sub * sizeof(int), st));
  return dtype == PYGB_F32 ? launch_gat<float>(P, st) : launch_gat<__nv_bfloat16>(P, st);
}

}  // extern "C"
