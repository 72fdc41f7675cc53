// K1 — segmented gather-reduce (aggr = add | mean), K3 — segmented max with
// argmax, and the row gather used by generic message() functions.
//
// Layout: features are row-major [n, ld] (fp32 or bf16); the graph is a CSR
// sorted by destination (rowptr int32 [n_rows+1], idx int32 [E]).  One warp
// owns one destination row.  The 32 lanes are split as (edge slot, feature
// lane): LPF lanes cover one feature row with 16-byte vectors (VPL vectors per
// lane), and 32/LPF edge slots walk different edges of the same row, so narrow
// rows still issue full-width requests.  The index/weight stream of a row is
// read coalesced, 32 entries at a time, and broadcast by shuffle; U edges are
// in flight per slot before the first FMA.  Partial sums of the edge slots are
// combined by a fixed xor-shuffle tree: no atomics, run-to-run deterministic.
//
// HBM-bound: algorithmic bytes per edge = 4 (idx) + 4 (weight) + F*sizeof(T),
// plus one output row per destination.
#include <float.h>
#include <math.h>

#include "common.cuh"

namespace pygb {
namespace {

constexpr int kWarpsPerCta = 8;

template <int VPL>
struct Unroll {
  static constexpr int kU = VPL >= 3 ? 2 : (VPL == 2 ? 4 : 8);
};

// ---------------------------------------------------------------------------
// K1
// ---------------------------------------------------------------------------
// HEADW: the weight is per (edge, head): ew[[wperm]p * H + head(feature)], head = f / C
// (GAT's alpha-weighted aggregation and its transposed backward).
template <typename T, bool VEC, int LPF, int VPL, bool HEADW>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
seg_gather_reduce_kernel(const T* __restrict__ x, int64_t ldx, int F,
                         const int* __restrict__ rowptr, const int* __restrict__ idx,
                         const float* __restrict__ ew, const int* __restrict__ wperm,
                         const float* __restrict__ bias, int64_t n_rows, int mean,
                         T* __restrict__ out, int64_t ldo, int H, int C) {
  constexpr int EPW = 32 / LPF;
  constexpr int U = Unroll<VPL>::kU;
  constexpr int KV = VEC ? Elem<T>::kVec : 1;
  const int64_t row = (int64_t)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
  if (row >= n_rows) return;
  const int lane = threadIdx.x & 31;
  const int fl = lane & (LPF - 1);
  const int es = lane / LPF;
  const int beg = rowptr[row], end = rowptr[row + 1];
  const int units = VEC ? (F + KV - 1) / KV : F;

  for (int unit0 = 0; unit0 < units; unit0 += LPF * VPL) {
    float acc[VPL][KV];
    int head[VPL];
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
      const int unit = unit0 + fl + v * LPF;
      head[v] = (HEADW && unit < units) ? (unit * KV) / C : 0;
#pragma unroll
      for (int i = 0; i < KV; ++i) acc[v][i] = 0.f;
    }

    for (int base = beg; base < end; base += 32) {
      const int n = min(32, end - base);
      int my_idx = 0;
      float my_w = 1.f;  // HEADW: bit pattern of the weight-row index instead
      if (lane < n) {
        my_idx = __ldg(idx + base + lane);
        if (HEADW)
          my_w = __int_as_float(wperm ? __ldg(wperm + base + lane) : base + lane);
        else if (ew)
          my_w = wperm ? __ldg(ew + __ldg(wperm + base + lane)) : __ldg(ew + base + lane);
      }
      for (int j = 0; j < n; j += EPW * U) {
        float vals[U][VPL][KV];
        float ws[U][HEADW ? VPL : 1];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int k = j + u * EPW + es;
          const int c = __shfl_sync(kFull, my_idx, k & 31);
          const float ww = __shfl_sync(kFull, my_w, k & 31);
          const bool valid = k < n;
          if (!HEADW) ws[u][0] = valid ? ww : 0.f;
          const T* xr = x + (int64_t)c * ldx;
#pragma unroll
          for (int v = 0; v < VPL; ++v) {
            const int unit = unit0 + fl + v * LPF;
            if (valid && unit < units) {
              if (VEC) {
                uint4 q = ldg_nc_v4(xr + (int64_t)unit * KV);
                Elem<T>::unpack(q, vals[u][v]);
              } else {
                vals[u][v][0] = Elem<T>::load(xr + unit);
              }
              if (HEADW) ws[u][v] = __ldg(ew + (int64_t)__float_as_int(ww) * H + head[v]);
            } else {
#pragma unroll
              for (int i = 0; i < KV; ++i) vals[u][v][i] = 0.f;
              if (HEADW) ws[u][v] = 0.f;
            }
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int v = 0; v < VPL; ++v)
#pragma unroll
            for (int i = 0; i < KV; ++i)
              acc[v][i] = fmaf(ws[u][HEADW ? v : 0], vals[u][v][i], acc[v][i]);
      }
    }
    // combine edge slots (fixed tree)
#pragma unroll
    for (int off = LPF; off < 32; off <<= 1)
#pragma unroll
      for (int v = 0; v < VPL; ++v)
#pragma unroll
        for (int i = 0; i < KV; ++i) acc[v][i] += __shfl_xor_sync(kFull, acc[v][i], off);

    if (es == 0) {
      const float cnt = (float)max(end - beg, 1);
      T* orow = out + row * ldo;
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        const int unit = unit0 + fl + v * LPF;
        if (unit >= units) continue;
        const int f0 = unit * KV;
#pragma unroll
        for (int i = 0; i < KV; ++i) {
          float r = acc[v][i];
          if (mean) r = r / cnt;
          if (bias && f0 + i < F) r += __ldg(bias + f0 + i);
          acc[v][i] = r;
        }
        if (VEC && f0 + KV <= F) {
          *reinterpret_cast<uint4*>(orow + f0) = Elem<T>::pack(acc[v]);
        } else {
#pragma unroll
          for (int i = 0; i < KV; ++i)
            if (f0 + i < F) Elem<T>::store(orow + f0 + i, acc[v][i]);
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// K3 — max + argmax.  Strict '>' in ascending edge order inside a slot, ties
// between slots go to the smaller edge id => the reference CPU loop's winner.
// ---------------------------------------------------------------------------
template <typename T, bool VEC, int LPF, int VPL>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
seg_max_kernel(const T* __restrict__ src, int64_t lds, int F, const int* __restrict__ rowptr,
               const int* __restrict__ idx, const int* __restrict__ eid, int64_t n_rows,
               int e_fill, float zero_below, int mask_arg, T* __restrict__ out, int64_t ldo,
               int64_t* __restrict__ arg) {
  constexpr int EPW = 32 / LPF;
  constexpr int U = Unroll<VPL>::kU;
  constexpr int KV = VEC ? Elem<T>::kVec : 1;
  const int64_t row = (int64_t)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
  if (row >= n_rows) return;
  const int lane = threadIdx.x & 31;
  const int fl = lane & (LPF - 1);
  const int es = lane / LPF;
  const int beg = rowptr[row], end = rowptr[row + 1];
  const int units = VEC ? (F + KV - 1) / KV : F;

  for (int unit0 = 0; unit0 < units; unit0 += LPF * VPL) {
    float best[VPL][KV];
    int barg[VPL][KV];
#pragma unroll
    for (int v = 0; v < VPL; ++v)
#pragma unroll
      for (int i = 0; i < KV; ++i) {
        best[v][i] = -FLT_MAX;
        barg[v][i] = e_fill;
      }
    for (int base = beg; base < end; base += 32) {
      const int n = min(32, end - base);
      int my_idx = 0, my_eid = e_fill;
      if (lane < n) {
        my_idx = __ldg(idx + base + lane);
        my_eid = eid ? __ldg(eid + base + lane) : base + lane;
      }
      for (int j = 0; j < n; j += EPW * U) {
        float vals[U][VPL][KV];
        int ids[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int k = j + u * EPW + es;
          const int c = __shfl_sync(kFull, my_idx, k & 31);
          ids[u] = __shfl_sync(kFull, my_eid, k & 31);
          const bool valid = k < n;
          const T* xr = src + (int64_t)c * lds;
#pragma unroll
          for (int v = 0; v < VPL; ++v) {
            const int unit = unit0 + fl + v * LPF;
            if (valid && unit < units) {
              if (VEC) {
                uint4 q = ldg_nc_v4(xr + (int64_t)unit * KV);
                Elem<T>::unpack(q, vals[u][v]);
              } else {
                vals[u][v][0] = Elem<T>::load(xr + unit);
              }
            } else {
#pragma unroll
              for (int i = 0; i < KV; ++i) vals[u][v][i] = -FLT_MAX;  // never wins
            }
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int v = 0; v < VPL; ++v)
#pragma unroll
            for (int i = 0; i < KV; ++i)
              if (vals[u][v][i] > best[v][i]) {
                best[v][i] = vals[u][v][i];
                barg[v][i] = ids[u];
              }
      }
    }
#pragma unroll
    for (int off = LPF; off < 32; off <<= 1)
#pragma unroll
      for (int v = 0; v < VPL; ++v)
#pragma unroll
        for (int i = 0; i < KV; ++i) {
          const float ov = __shfl_xor_sync(kFull, best[v][i], off);
          const int oa = __shfl_xor_sync(kFull, barg[v][i], off);
          if (ov > best[v][i] || (ov == best[v][i] && oa < barg[v][i])) {
            best[v][i] = ov;
            barg[v][i] = oa;
          }
        }
    if (es == 0) {
      T* orow = out + row * ldo;
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        const int unit = unit0 + fl + v * LPF;
        if (unit >= units) continue;
        const int f0 = unit * KV;
#pragma unroll
        for (int i = 0; i < KV; ++i) {
          float r = best[v][i];
          if (barg[v][i] == e_fill) r = 0.f;  // no entry (Q5)
          if (r < zero_below) {               // scatter_'s -10000 rule (Q2)
            r = 0.f;
            if (mask_arg) barg[v][i] = e_fill;  // in-place overwrite: no gradient
          }
          best[v][i] = r;
        }
        if (VEC && f0 + KV <= F) {
          *reinterpret_cast<uint4*>(orow + f0) = Elem<T>::pack(best[v]);
        } else {
#pragma unroll
          for (int i = 0; i < KV; ++i)
            if (f0 + i < F) Elem<T>::store(orow + f0 + i, best[v][i]);
        }
        if (arg) {
#pragma unroll
          for (int i = 0; i < KV; ++i)
            if (f0 + i < F) arg[row * (int64_t)F + f0 + i] = (int64_t)barg[v][i];
        }
      }
    }
  }
}

__global__ void __launch_bounds__(256)
seg_max_backward_kernel(const float* __restrict__ g, int64_t ldg, const int64_t* __restrict__ arg,
                        int64_t total, int F, int64_t E, float* __restrict__ gs, int64_t lds) {
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * 256) {
    const int64_t i = t / F;
    const int f = (int)(t - i * F);
    const int64_t a = arg[t];
    if (a >= 0 && a < E) gs[a * lds + f] = g[i * ldg + f];
  }
}

// ---------------------------------------------------------------------------
// row gather  out[e,:] = x[index[e],:]
// ---------------------------------------------------------------------------
template <typename T, bool VEC>
__global__ void __launch_bounds__(256)
gather_rows_kernel(const T* __restrict__ x, int64_t ldx, int F, const int64_t* __restrict__ index,
                   int64_t E, T* __restrict__ out, int64_t ldo) {
  constexpr int KV = VEC ? Elem<T>::kVec : 1;
  const int units = F / KV;  // VEC path requires F % KV == 0
  const int64_t total = E * units;
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * 256) {
    const int64_t e = t / units;
    const int u = (int)(t - e * units);
    const int64_t r = __ldg(index + e);
    if (VEC) {
      uint4 q = ldg_nc_v4(x + r * ldx + (int64_t)u * KV);
      *reinterpret_cast<uint4*>(out + e * ldo + (int64_t)u * KV) = q;
    } else {
      out[e * ldo + u] = x[r * ldx + u];
    }
  }
}

// ---------------------------------------------------------------------------
// dispatch helpers
// ---------------------------------------------------------------------------
struct Shape {
  bool vec;
  int lpf, vpl;
};

template <typename T>
Shape pick_shape(const void* x, int64_t ldx, int64_t F, const void* out, int64_t ldo,
                 int64_t C = 0) {
  constexpr int KV = Elem<T>::kVec;
  const int64_t units_v = (F + KV - 1) / KV;
  bool vec = aligned16(x) && aligned16(out) && (ldx % KV == 0) && (ldo % KV == 0) &&
             (ldx >= units_v * KV) && (C == 0 || C % KV == 0);
  int64_t units = vec ? units_v : F;
  Shape s;
  s.vec = vec;
  if (units <= 32) {
    int l = 1;
    while (l < units) l <<= 1;
    s.lpf = l;
    s.vpl = 1;
  } else {
    s.lpf = 32;
    int v = (int)((units + 31) / 32);
    if (v > 4) v = 4;
    if (!vec && v == 3) v = 4;
    s.vpl = v;
  }
  return s;
}

#define PYGB_SHAPE_CASES(T, KERNEL_LAUNCH)                                        \
  if (sh.vec) {                                                                   \
    switch (sh.lpf * 8 + sh.vpl) {                                                \
      case 1 * 8 + 1: KERNEL_LAUNCH(T, true, 1, 1); break;                        \
      case 2 * 8 + 1: KERNEL_LAUNCH(T, true, 2, 1); break;                        \
      case 4 * 8 + 1: KERNEL_LAUNCH(T, true, 4, 1); break;                        \
      case 8 * 8 + 1: KERNEL_LAUNCH(T, true, 8, 1); break;                        \
      case 16 * 8 + 1: KERNEL_LAUNCH(T, true, 16, 1); break;                      \
      case 32 * 8 + 1: KERNEL_LAUNCH(T, true, 32, 1); break;                      \
      case 32 * 8 + 2: KERNEL_LAUNCH(T, true, 32, 2); break;                      \
      case 32 * 8 + 3: KERNEL_LAUNCH(T, true, 32, 3); break;                      \
      default: KERNEL_LAUNCH(T, true, 32, 4); break;                              \
    }                                                                             \
  } else {                                                                        \
    switch (sh.lpf * 8 + sh.vpl) {                                                \
      case 1 * 8 + 1: KERNEL_LAUNCH(T, false, 1, 1); break;                       \
      case 2 * 8 + 1: KERNEL_LAUNCH(T, false, 2, 1); break;                       \
      case 4 * 8 + 1: KERNEL_LAUNCH(T, false, 4, 1); break;                       \
      case 8 * 8 + 1: KERNEL_LAUNCH(T, false, 8, 1); break;                       \
      case 16 * 8 + 1: KERNEL_LAUNCH(T, false, 16, 1); break;                     \
      case 32 * 8 + 1: KERNEL_LAUNCH(T, false, 32, 1); break;                     \
      case 32 * 8 + 2: KERNEL_LAUNCH(T, false, 32, 2); break;                     \
      default: KERNEL_LAUNCH(T, false, 32, 4); break;                             \
    }                                                                             \
  }

template <typename T, bool HEADW>
int launch_gather_reduce(int reduce, const void* x, int64_t ldx, int64_t F, const int* rowptr,
                         const int* idx, const float* ew, const int* wperm, const float* bias,
                         int64_t n_rows, void* out, int64_t ldo, int H, int C, cudaStream_t st) {
  Shape sh = pick_shape<T>(x, ldx, F, out, ldo, HEADW ? C : 0);
  const unsigned grid = (unsigned)ceil_div(n_rows, kWarpsPerCta);
  const int mean = reduce == PYGB_MEAN;
#define LAUNCH_K1(TT, V, L, P)                                                             \
  seg_gather_reduce_kernel<TT, V, L, P, HEADW><<<grid, kWarpsPerCta * 32, 0, st>>>(        \
      static_cast<const TT*>(x), ldx, (int)F, rowptr, idx, ew, wperm, bias, n_rows, mean,  \
      static_cast<TT*>(out), ldo, H, C)
  PYGB_SHAPE_CASES(T, LAUNCH_K1)
#undef LAUNCH_K1
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

template <typename T>
int launch_seg_max(const void* src, int64_t lds, int64_t F, const int* rowptr, const int* idx,
                   const int* eid, int64_t n_rows, int64_t e_fill, float zero_below, int mask_arg,
                   void* out, int64_t ldo, int64_t* arg, cudaStream_t st) {
  Shape sh = pick_shape<T>(src, lds, F, out, ldo);
  const unsigned grid = (unsigned)ceil_div(n_rows, kWarpsPerCta);
#define LAUNCH_K3(TT, V, L, P)                                                           \
  seg_max_kernel<TT, V, L, P><<<grid, kWarpsPerCta * 32, 0, st>>>(                       \
      static_cast<const TT*>(src), lds, (int)F, rowptr, idx, eid, n_rows, (int)e_fill,   \
      zero_below, mask_arg, static_cast<TT*>(out), ldo, arg)
  PYGB_SHAPE_CASES(T, LAUNCH_K3)
#undef LAUNCH_K3
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

template <typename T>
int launch_gather_rows(const void* x, int64_t ldx, int64_t F, const int64_t* index, int64_t E,
                       void* out, int64_t ldo, cudaStream_t st) {
  constexpr int KV = Elem<T>::kVec;
  const bool vec = aligned16(x) && aligned16(out) && ldx % KV == 0 && ldo % KV == 0 && F % KV == 0;
  const int64_t total = E * (vec ? F / KV : F);
  int64_t blocks = ceil_div(total, 256);
  if (blocks > 148 * 32) blocks = 148 * 32;
  if (vec)
    gather_rows_kernel<T, true><<<(unsigned)blocks, 256, 0, st>>>(
        static_cast<const T*>(x), ldx, (int)F, index, E, static_cast<T*>(out), ldo);
  else
    gather_rows_kernel<T, false><<<(unsigned)blocks, 256, 0, st>>>(
        static_cast<const T*>(x), ldx, (int)F, index, E, static_cast<T*>(out), ldo);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

}  // namespace
}  // namespace pygb

using namespace pygb;

extern "C" {

int pygb_seg_gather_reduce(int reduce, int dtype, const void* x, int64_t ldx, int64_t F,
                           const int32_t* rowptr, const int32_t* idx, const float* edge_w,
                           const int32_t* wperm, const float* bias, int64_t n_rows, void* out,
                           int64_t ldo, pygb_stream_t s) {
  PYGB_REQUIRE(reduce == PYGB_SUM || reduce == PYGB_MEAN, "pygb_seg_gather_reduce: bad reduce %d",
               reduce);
  PYGB_REQUIRE(dtype == PYGB_F32 || dtype == PYGB_BF16, "pygb_seg_gather_reduce: bad dtype %d",
               dtype);
  PYGB_REQUIRE(F >= 0 && n_rows >= 0 && ldx >= 0 && ldo >= F && F < (1ll << 31),
               "pygb_seg_gather_reduce: bad sizes");
  PYGB_REQUIRE(!(wperm && !edge_w), "pygb_seg_gather_reduce: wperm without edge_w");
  if (n_rows == 0 || F == 0) return PYGB_OK;
  PYGB_REQUIRE(rowptr && out, "pygb_seg_gather_reduce: null pointer");
  cudaStream_t st = as_stream(s);
  if (dtype == PYGB_F32)
    return launch_gather_reduce<float, false>(reduce, x, ldx, F, rowptr, idx, edge_w, wperm, bias,
                                              n_rows, out, ldo, 1, 1, st);
  return launch_gather_reduce<__nv_bfloat16, false>(reduce, x, ldx, F, rowptr, idx, edge_w, wperm,
                                                    bias, n_rows, out, ldo, 1, 1, st);
}

int pygb_seg_gather_reduce_heads(int dtype, const void* x, int64_t ldx, int H, int C,
                                 const int32_t* rowptr, const int32_t* idx, const float* head_w,
                                 const int32_t* wperm, const float* bias, int64_t n_rows,
                                 void* out, int64_t ldo, pygb_stream_t s) {
  PYGB_REQUIRE(dtype == PYGB_F32 || dtype == PYGB_BF16, "pygb_seg_gather_reduce_heads: bad dtype");
  PYGB_REQUIRE(H > 0 && C > 0 && (int64_t)H * C < (1ll << 31) && n_rows >= 0 &&
                   ldo >= (int64_t)H * C,
               "pygb_seg_gather_reduce_heads: bad sizes");
  if (n_rows == 0) return PYGB_OK;
  PYGB_REQUIRE(rowptr && out && head_w, "pygb_seg_gather_reduce_heads: null pointer");
  cudaStream_t st = as_stream(s);
  const int64_t F = (int64_t)H * C;
  if (dtype == PYGB_F32)
    return launch_gather_reduce<float, true>(PYGB_SUM, x, ldx, F, rowptr, idx, head_w, wperm, bias,
                                             n_rows, out, ldo, H, C, st);
  return launch_gather_reduce<__nv_bfloat16, true>(PYGB_SUM, x, ldx, F, rowptr, idx, head_w, wperm,
                                                   bias, n_rows, out, ldo, H, C, st);
}

int pygb_gather_rows(int dtype, const void* x, int64_t ldx, int64_t F, const int64_t* index,
                     int64_t E, void* out, int64_t ldo, pygb_stream_t s) {
  PYGB_REQUIRE(dtype == PYGB_F32 || dtype == PYGB_BF16, "pygb_gather_rows: bad dtype %d", dtype);
  PYGB_REQUIRE(F >= 0 && E >= 0 && ldo >= F && ldx >= 0 && F < (1ll << 31),
               "pygb_gather_rows: bad sizes");
  if (E == 0 || F == 0) return PYGB_OK;
  PYGB_REQUIRE(x && index && out, "pygb_gather_rows: null pointer");
  cudaStream_t st = as_stream(s);
  if (dtype == PYGB_F32) return launch_gather_rows<float>(x, ldx, F, index, E, out, ldo, st);
  return launch_gather_rows<__nv_bfloat16>(x, ldx, F, index, E, out, ldo, st);
}

int pygb_seg_max(int dtype, const void* src, int64_t lds, int64_t F, const int32_t* rowptr,
                 const int32_t* idx, const int32_t* eid, int64_t n_rows, int64_t E_fill,
                 float zero_below, int mask_arg, void* out, int64_t ldo, int64_t* arg,
                 pygb_stream_t s) {
  PYGB_REQUIRE(dtype == PYGB_F32 || dtype == PYGB_BF16, "pygb_seg_max: bad dtype %d", dtype);
  PYGB_REQUIRE(F >= 0 && n_rows >= 0 && ldo >= F && E_fill >= 0 && E_fill < (1ll << 31) &&
                   F < (1ll << 31),
               "pygb_seg_max: bad sizes");
  if (n_rows == 0 || F == 0) return PYGB_OK;
  PYGB_REQUIRE(rowptr && out, "pygb_seg_max: null pointer");
  cudaStream_t st = as_stream(s);
  if (dtype == PYGB_F32)
    return launch_seg_max<float>(src, lds, F, rowptr, idx, eid, n_rows, E_fill, zero_below,
                                 mask_arg, out, ldo, arg, st);
  return launch_seg_max<__nv_bfloat16>(src, lds, F, rowptr, idx, eid, n_rows, E_fill, zero_below,
                                       mask_arg, out, ldo, arg, st);
}

int pygb_seg_max_backward(const float* grad_out, int64_t ldg, const int64_t* arg, int64_t n_rows,
                          int64_t F, int64_t E, float* grad_src, int64_t lds, pygb_stream_t s) {
  PYGB_REQUIRE(n_rows >= 0 && F >= 0 && E >= 0 && F < (1ll << 31), "pygb_seg_max_backward: bad sizes");
  const int64_t total = n_rows * F;
  if (total == 0 || E == 0) return PYGB_OK;
  int64_t blocks = ceil_div(total, 256);
  if (blocks > 148 * 32) blocks = 148 * 32;
  seg_max_backward_kernel<<<(unsigned)blocks, 256, 0, as_stream(s)>>>(grad_out, ldg, arg, total,
                                                                    (int)F, E, grad_src, lds);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

}  // extern "C"
