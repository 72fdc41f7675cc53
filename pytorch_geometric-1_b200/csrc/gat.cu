// K2 — per-destination edge softmax and the fused GATConv propagate.
//
// Reference: alpha = leaky_relu((cat[x_i,x_j]*att).sum(-1)),
// softmax over the edges of each target (scatter_max, exp, scatter_add, +1e-16),
// out = scatter_add(x_j * alpha).  That is ~12 passes over E'-sized tensors and
// four [E',H] temporaries.  Here the logit splits as a_dst[i,h] + a_src[j,h]
// (two [N,H] vectors), and one warp per destination row walks its CSR slice
// twice: pass 1 takes the exact row maximum of the logits (only the 4·H-byte
// a_src rows are touched, L2 resident), pass 2 gathers xw[j] once and
// accumulates exp(l-m)·xw[j] and the denominator.  No [E',·] tensor is written.
#include <float.h>
#include <math.h>

#include "common.cuh"

namespace pygb {
namespace {

constexpr int kWarpsPerCta = 8;

__device__ __forceinline__ float leaky(float x, float slope) { return x > 0.f ? x : x * slope; }

// a_dst[n,h] = <xw[n,h,:], att[h,:C]> ; a_src[n,h] = <xw[n,h,:], att[h,C:]>
template <typename T>
__global__ void __launch_bounds__(256)
gat_scores_kernel(const T* __restrict__ xw, int64_t ldx, const float* __restrict__ att, int64_t N,
                  int H, int C, float* __restrict__ a_src, float* __restrict__ a_dst) {
  const int64_t total = N * H;
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * 256) {
    const int64_t n = t / H;
    const int h = (int)(t - n * H);
    const T* xr = xw + n * ldx + (int64_t)h * C;
    const float* ad = att + (int64_t)h * 2 * C;
    float sd = 0.f, ss = 0.f;
    for (int c = 0; c < C; ++c) {
      const float v = Elem<T>::load(xr + c);
      sd = fmaf(v, __ldg(ad + c), sd);
      ss = fmaf(v, __ldg(ad + C + c), ss);
    }
    if (a_dst) a_dst[t] = sd;
    a_src[t] = ss;
  }
}

template <typename T, bool VEC, int LPF, int VPL>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
gat_forward_kernel(const T* __restrict__ xw, int64_t ldx, int F, int H, int C,
                   const float* __restrict__ a_src, const float* __restrict__ a_dst, float slope,
                   const int* __restrict__ rowptr, const int* __restrict__ col, int64_t n_rows,
                   const float* __restrict__ bias, T* __restrict__ out, int64_t ldo,
                   float* __restrict__ row_max, float* __restrict__ row_sum) {
  constexpr int EPW = 32 / LPF;
  constexpr int U = VPL >= 2 ? 2 : 4;
  constexpr int KV = VEC ? Elem<T>::kVec : 1;
  const int64_t row = (int64_t)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
  if (row >= n_rows) return;
  const int lane = threadIdx.x & 31;
  const int fl = lane & (LPF - 1);
  const int es = lane / LPF;
  const int beg = rowptr[row], end = rowptr[row + 1];
  const int units = VEC ? (F + KV - 1) / KV : F;

  for (int unit0 = 0; unit0 < units; unit0 += LPF * VPL) {
    int head[VPL];
    float adst[VPL], m[VPL], ssum[VPL];
    float acc[VPL][KV];
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
      const int unit = unit0 + fl + v * LPF;
      head[v] = unit < units ? (unit * KV) / C : 0;
      adst[v] = a_dst ? __ldg(a_dst + row * H + head[v]) : 0.f;
      m[v] = -FLT_MAX;
      ssum[v] = 0.f;
#pragma unroll
      for (int i = 0; i < KV; ++i) acc[v][i] = 0.f;
    }
    // ---- pass 1: exact row max of the logits --------------------------------
    for (int base = beg; base < end; base += 32) {
      const int n = min(32, end - base);
      const int my_idx = lane < n ? __ldg(col + base + lane) : 0;
      for (int j = 0; j < n; j += EPW) {
        const int k = j + es;
        const int c = __shfl_sync(kFull, my_idx, k & 31);
        if (k < n) {
#pragma unroll
          for (int v = 0; v < VPL; ++v) {
            const float l = leaky(adst[v] + __ldg(a_src + (int64_t)c * H + head[v]), slope);
            m[v] = fmaxf(m[v], l);
          }
        }
      }
    }
#pragma unroll
    for (int off = LPF; off < 32; off <<= 1)
#pragma unroll
      for (int v = 0; v < VPL; ++v) m[v] = fmaxf(m[v], __shfl_xor_sync(kFull, m[v], off));

    // ---- pass 2: exp, denominator, weighted gather ---------------------------
    for (int base = beg; base < end; base += 32) {
      const int n = min(32, end - base);
      const int my_idx = lane < n ? __ldg(col + base + lane) : 0;
      for (int j = 0; j < n; j += EPW * U) {
        float vals[U][VPL][KV];
        float ex[U][VPL];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int k = j + u * EPW + es;
          const int c = __shfl_sync(kFull, my_idx, k & 31);
          const bool valid = k < n;
          const T* xr = xw + (int64_t)c * ldx;
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
              const float l = leaky(adst[v] + __ldg(a_src + (int64_t)c * H + head[v]), slope);
              ex[u][v] = __expf(l - m[v]);
            } else {
#pragma unroll
              for (int i = 0; i < KV; ++i) vals[u][v][i] = 0.f;
              ex[u][v] = 0.f;
            }
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int v = 0; v < VPL; ++v) {
            ssum[v] += ex[u][v];
#pragma unroll
            for (int i = 0; i < KV; ++i) acc[v][i] = fmaf(ex[u][v], vals[u][v][i], acc[v][i]);
          }
      }
    }
#pragma unroll
    for (int off = LPF; off < 32; off <<= 1)
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        ssum[v] += __shfl_xor_sync(kFull, ssum[v], off);
#pragma unroll
        for (int i = 0; i < KV; ++i) acc[v][i] += __shfl_xor_sync(kFull, acc[v][i], off);
      }

    if (es == 0) {
      T* orow = out + row * ldo;
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        const int unit = unit0 + fl + v * LPF;
        if (unit >= units) continue;
        const int f0 = unit * KV;
        const float den = ssum[v] + 1e-16f;
#pragma unroll
        for (int i = 0; i < KV; ++i) {
          float r = acc[v][i] / den;
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
        if (f0 % C == 0 && f0 < F) {  // first lane of the head records the statistics
          if (row_max) row_max[row * H + head[v]] = m[v];
          if (row_sum) row_sum[row * H + head[v]] = ssum[v];
        }
      }
    }
  }
}

// Per-edge quantities of the GAT backward (dst-sorted rows), fp32:
//   alpha_p, dalpha_p = <g[i,h,:], xw[j,h,:]>, ds_p = alpha_p (dalpha_p - D[i,h]) leaky'(s_p)
// with D[i,h] = <g[i,h,:], out[i,h,:] - bias[h,:]> = sum_p alpha_p dalpha_p.
// Lanes walk the (edge, head) pairs of the row; results go to ORIGINAL edge order.
__global__ void __launch_bounds__(kWarpsPerCta * 32)
gat_edge_grad_kernel(const float* __restrict__ xw, int64_t ldx, const float* __restrict__ a_src,
                     const float* __restrict__ a_dst, int H, int C, float slope,
                     const int* __restrict__ rowptr, const int* __restrict__ col,
                     const int* __restrict__ perm, int64_t n_rows, const float* __restrict__ out,
                     int64_t ldo, const float* __restrict__ bias, const float* __restrict__ g,
                     int64_t ldg, const float* __restrict__ row_max,
                     const float* __restrict__ row_sum, float* __restrict__ alpha_out,
                     float* __restrict__ ds_out) {
  extern __shared__ float d_smem[];  // [kWarpsPerCta][H]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * kWarpsPerCta + warp;
  if (row >= n_rows) return;
  float* D = d_smem + warp * H;
  const float* gr = g + row * ldg;
  const float* orow = out + row * ldo;
  for (int h = lane; h < H; h += 32) {
    float d = 0.f;
    for (int c = 0; c < C; ++c) {
      float o = orow[h * C + c];
      if (bias) o -= __ldg(bias + h * C + c);
      d = fmaf(gr[h * C + c], o, d);
    }
    D[h] = d;
  }
  __syncwarp();
  const int beg = rowptr[row], end = rowptr[row + 1];
  const int64_t work = (int64_t)(end - beg) * H;
  for (int64_t t = lane; t < work; t += 32) {
    const int p = beg + (int)(t / H);
    const int h = (int)(t % H);
    const int j = __ldg(col + p);
    const float s = (a_dst ? __ldg(a_dst + row * H + h) : 0.f) + __ldg(a_src + (int64_t)j * H + h);
    const float l = leaky(s, slope);
    const float alpha = expf(l - row_max[row * H + h]) / (row_sum[row * H + h] + 1e-16f);
    const float* xr = xw + (int64_t)j * ldx + (int64_t)h * C;
    const float* gh = gr + (int64_t)h * C;
    float dot = 0.f;
    for (int c = 0; c < C; ++c) dot = fmaf(gh[c], __ldg(xr + c), dot);
    const float dl = alpha * (dot - D[h]);
    const int64_t e = perm ? perm[p] : p;
    alpha_out[e * H + h] = alpha;
    ds_out[e * H + h] = dl * (s > 0.f ? 1.f : slope);
  }
}

// alpha only (original edge order) — used when dropout has to be applied to alpha.
__global__ void __launch_bounds__(kWarpsPerCta * 32)
gat_alpha_kernel(const float* __restrict__ a_src, const float* __restrict__ a_dst, int H,
                 float slope, const int* __restrict__ rowptr, const int* __restrict__ col,
                 const int* __restrict__ perm, int64_t n_rows, const float* __restrict__ row_max,
                 const float* __restrict__ row_sum, float* __restrict__ alpha_out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * kWarpsPerCta + warp;
  if (row >= n_rows) return;
  const int beg = rowptr[row], end = rowptr[row + 1];
  const int64_t work = (int64_t)(end - beg) * H;
  for (int64_t t = lane; t < work; t += 32) {
    const int p = beg + (int)(t / H);
    const int h = (int)(t % H);
    const int j = __ldg(col + p);
    const float s = (a_dst ? __ldg(a_dst + row * H + h) : 0.f) + __ldg(a_src + (int64_t)j * H + h);
    const float alpha =
        expf(leaky(s, slope) - row_max[row * H + h]) / (row_sum[row * H + h] + 1e-16f);
    const int64_t e = perm ? perm[p] : p;
    alpha_out[e * H + h] = alpha;
  }
}

// ---------------------------------------------------------------------------
// utils.softmax over [E,H] logits in original edge order (A.7)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kWarpsPerCta * 32)
edge_softmax_kernel(const float* __restrict__ src, int H, const int* __restrict__ rowptr,
                    const int* __restrict__ perm, int64_t n_rows, float* __restrict__ out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * kWarpsPerCta + warp;
  if (row >= n_rows) return;
  const int beg = rowptr[row], end = rowptr[row + 1];
  const int deg = end - beg;
  if (deg == 0) return;
  // lanes are (edge slot, head lane): LPH = smallest power of two >= min(H,32)
  int lph = 1;
  while (lph < H && lph < 32) lph <<= 1;
  const int hl = lane & (lph - 1), es = lane / lph, epw = 32 / lph;
  for (int h0 = 0; h0 < H; h0 += lph) {
    const int h = h0 + hl;
    const bool hv = h < H;
    float m = -FLT_MAX;
    for (int k = es; k < deg; k += epw)
      if (hv) m = fmaxf(m, src[(int64_t)perm[beg + k] * H + h]);
    for (int off = lph; off < 32; off <<= 1) m = fmaxf(m, __shfl_xor_sync(kFull, m, off));
    float s = 0.f;
    for (int k = es; k < deg; k += epw)
      if (hv) s += expf(src[(int64_t)perm[beg + k] * H + h] - m);
    for (int off = lph; off < 32; off <<= 1) s += __shfl_xor_sync(kFull, s, off);
    const float den = s + 1e-16f;
    for (int k = es; k < deg; k += epw)
      if (hv) {
        const int64_t o = (int64_t)perm[beg + k] * H + h;
        out[o] = expf(src[o] - m) / den;
      }
  }
}

// grad_src = y * (g - sum_row(y*g))
__global__ void __launch_bounds__(kWarpsPerCta * 32)
edge_softmax_bwd_kernel(const float* __restrict__ y, const float* __restrict__ g, int H,
                        const int* __restrict__ rowptr, const int* __restrict__ perm,
                        int64_t n_rows, float* __restrict__ gs) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * kWarpsPerCta + warp;
  if (row >= n_rows) return;
  const int beg = rowptr[row], end = rowptr[row + 1];
  const int deg = end - beg;
  if (deg == 0) return;
  int lph = 1;
  while (lph < H && lph < 32) lph <<= 1;
  const int hl = lane & (lph - 1), es = lane / lph, epw = 32 / lph;
  for (int h0 = 0; h0 < H; h0 += lph) {
    const int h = h0 + hl;
    const bool hv = h < H;
    float s = 0.f;
    for (int k = es; k < deg; k += epw)
      if (hv) {
        const int64_t o = (int64_t)perm[beg + k] * H + h;
        s = fmaf(y[o], g[o], s);
      }
    for (int off = lph; off < 32; off <<= 1) s += __shfl_xor_sync(kFull, s, off);
    for (int k = es; k < deg; k += epw)
      if (hv) {
        const int64_t o = (int64_t)perm[beg + k] * H + h;
        gs[o] = y[o] * (g[o] - s);
      }
  }
}

struct Shape {
  bool vec;
  int lpf, vpl;
};

template <typename T>
Shape pick_shape(const void* x, int64_t ldx, int64_t F, int C, const void* out, int64_t ldo) {
  constexpr int KV = Elem<T>::kVec;
  bool vec = aligned16(x) && aligned16(out) && (ldx % KV == 0) && (ldo % KV == 0) &&
             (C % KV == 0);
  int64_t units = vec ? F / KV : F;
  Shape s;
  s.vec = vec;
  if (units <= 32) {
    int l = 1;
    while (l < units) l <<= 1;
    s.lpf = l;
    s.vpl = 1;
  } else {
    s.lpf = 32;
    s.vpl = units > 64 ? 4 : 2;
  }
  return s;
}

template <typename T>
int launch_gat_forward(const void* xw, int64_t ldx, int H, int C, const float* a_src,
                       const float* a_dst, float slope, const int* rowptr, const int* col,
                       int64_t n_rows, const float* bias, void* out, int64_t ldo, float* row_max,
                       float* row_sum, cudaStream_t st) {
  const int F = H * C;
  Shape sh = pick_shape<T>(xw, ldx, F, C, out, ldo);
  const unsigned grid = (unsigned)ceil_div(n_rows, kWarpsPerCta);
#define LAUNCH_GAT(V, L, P)                                                                   \
  gat_forward_kernel<T, V, L, P><<<grid, kWarpsPerCta * 32, 0, st>>>(                          \
      static_cast<const T*>(xw), ldx, F, H, C, a_src, a_dst, slope, rowptr, col, n_rows, bias, \
      static_cast<T*>(out), ldo, row_max, row_sum)
  if (sh.vec) {
    switch (sh.lpf * 8 + sh.vpl) {
      case 1 * 8 + 1: LAUNCH_GAT(true, 1, 1); break;
      case 2 * 8 + 1: LAUNCH_GAT(true, 2, 1); break;
      case 4 * 8 + 1: LAUNCH_GAT(true, 4, 1); break;
      case 8 * 8 + 1: LAUNCH_GAT(true, 8, 1); break;
      case 16 * 8 + 1: LAUNCH_GAT(true, 16, 1); break;
      case 32 * 8 + 1: LAUNCH_GAT(true, 32, 1); break;
      case 32 * 8 + 2: LAUNCH_GAT(true, 32, 2); break;
      default: LAUNCH_GAT(true, 32, 4); break;
    }
  } else {
    switch (sh.lpf * 8 + sh.vpl) {
      case 1 * 8 + 1: LAUNCH_GAT(false, 1, 1); break;
      case 2 * 8 + 1: LAUNCH_GAT(false, 2, 1); break;
      case 4 * 8 + 1: LAUNCH_GAT(false, 4, 1); break;
      case 8 * 8 + 1: LAUNCH_GAT(false, 8, 1); break;
      case 16 * 8 + 1: LAUNCH_GAT(false, 16, 1); break;
      case 32 * 8 + 1: LAUNCH_GAT(false, 32, 1); break;
      case 32 * 8 + 2: LAUNCH_GAT(false, 32, 2); break;
      default: LAUNCH_GAT(false, 32, 4); break;
    }
  }
#undef LAUNCH_GAT
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

}  // namespace
}  // namespace pygb

using namespace pygb;

extern "C" {

int pygb_gat_scores(int dtype, const void* xw, int64_t ldx, const float* att, int64_t N, int H,
                    int C, float* a_src, float* a_dst, pygb_stream_t s) {
  PYGB_REQUIRE(dtype == PYGB_F32 || dtype == PYGB_BF16, "pygb_gat_scores: bad dtype");
  PYGB_REQUIRE(N >= 0 && H > 0 && C > 0 && a_src && att, "pygb_gat_scores: bad arguments");
  if (N == 0) return PYGB_OK;
  int64_t blocks = ceil_div(N * H, 256);
  if (blocks > 148 * 32) blocks = 148 * 32;
  if (dtype == PYGB_F32)
    gat_scores_kernel<float><<<(unsigned)blocks, 256, 0, as_stream(s)>>>(
        static_cast<const float*>(xw), ldx, att, N, H, C, a_src, a_dst);
  else
    gat_scores_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, as_stream(s)>>>(
        static_cast<const __nv_bfloat16*>(xw), ldx, att, N, H, C, a_src, a_dst);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

int pygb_gat_forward(int dtype, const void* xw, int64_t ldx, const float* a_src, const float* a_dst,
                     int H, int C, float negative_slope, const int32_t* rowptr, const int32_t* col,
                     int64_t n_rows, const float* bias, void* out, int64_t ldo, float* row_max,
                     float* row_sum, pygb_stream_t s) {
  PYGB_REQUIRE(dtype == PYGB_F32 || dtype == PYGB_BF16, "pygb_gat_forward: bad dtype");
  PYGB_REQUIRE(H > 0 && C > 0 && n_rows >= 0 && (int64_t)H * C < (1ll << 31) &&
                   ldo >= (int64_t)H * C,
               "pygb_gat_forward: bad sizes");
  if (n_rows == 0) return PYGB_OK;
  PYGB_REQUIRE(xw && a_src && rowptr && out, "pygb_gat_forward: null pointer");
  cudaStream_t st = as_stream(s);
  if (dtype == PYGB_F32)
    return launch_gat_forward<float>(xw, ldx, H, C, a_src, a_dst, negative_slope, rowptr, col,
                                     n_rows, bias, out, ldo, row_max, row_sum, st);
  return launch_gat_forward<__nv_bfloat16>(xw, ldx, H, C, a_src, a_dst, negative_slope, rowptr,
                                           col, n_rows, bias, out, ldo, row_max, row_sum, st);
}

int pygb_gat_edge_grad(const float* xw, int64_t ldx, const float* a_src, const float* a_dst, int H,
                       int C, float negative_slope, const int32_t* rowptr, const int32_t* col,
                       const int32_t* perm, int64_t n_rows, const float* out, int64_t ldo,
                       const float* bias, const float* grad_out, int64_t ldg,
                       const float* row_max, const float* row_sum, float* alpha, float* ds,
                       pygb_stream_t s) {
  PYGB_REQUIRE(H > 0 && C > 0 && n_rows >= 0, "pygb_gat_edge_grad: bad sizes");
  PYGB_REQUIRE((size_t)kWarpsPerCta * H * sizeof(float) <= 48 * 1024,
               "pygb_gat_edge_grad: too many heads (%d)", H);
  if (n_rows == 0) return PYGB_OK;
  PYGB_REQUIRE(xw && a_src && rowptr && out && grad_out && row_max && row_sum && alpha && ds,
               "pygb_gat_edge_grad: null pointer");
  const unsigned grid = (unsigned)ceil_div(n_rows, kWarpsPerCta);
  gat_edge_grad_kernel<<<grid, kWarpsPerCta * 32, kWarpsPerCta * H * sizeof(float), as_stream(s)>>>(
      xw, ldx, a_src, a_dst, H, C, negative_slope, rowptr, col, perm, n_rows, out, ldo, bias,
      grad_out, ldg, row_max, row_sum, alpha, ds);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

int pygb_gat_alpha(const float* a_src, const float* a_dst, int H, float negative_slope,
                   const int32_t* rowptr, const int32_t* col, const int32_t* perm, int64_t n_rows,
                   const float* row_max, const float* row_sum, float* alpha, pygb_stream_t s) {
  PYGB_REQUIRE(H > 0 && n_rows >= 0, "pygb_gat_alpha: bad sizes");
  if (n_rows == 0) return PYGB_OK;
  PYGB_REQUIRE(a_src && rowptr && row_max && row_sum && alpha, "pygb_gat_alpha: null pointer");
  const unsigned grid = (unsigned)ceil_div(n_rows, kWarpsPerCta);
  gat_alpha_kernel<<<grid, kWarpsPerCta * 32, 0, as_stream(s)>>>(
      a_src, a_dst, H, negative_slope, rowptr, col, perm, n_rows, row_max, row_sum, alpha);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

int pygb_edge_softmax(const float* src, int64_t H, const int32_t* rowptr, const int32_t* perm,
                      int64_t n_rows, float* out, pygb_stream_t s) {
  PYGB_REQUIRE(H > 0 && H < (1 << 30) && n_rows >= 0, "pygb_edge_softmax: bad sizes");
  if (n_rows == 0) return PYGB_OK;
  PYGB_REQUIRE(src && rowptr && perm && out, "pygb_edge_softmax: null pointer");
  const unsigned grid = (unsigned)ceil_div(n_rows, kWarpsPerCta);
  edge_softmax_kernel<<<grid, kWarpsPerCta * 32, 0, as_stream(s)>>>(src, (int)H, rowptr, perm,
                                                                   n_rows, out);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

int pygb_edge_softmax_backward(const float* out, const float* grad_out, int64_t H,
                               const int32_t* rowptr, const int32_t* perm, int64_t n_rows,
                               float* grad_src, pygb_stream_t s) {
  PYGB_REQUIRE(H > 0 && H < (1 << 30) && n_rows >= 0, "pygb_edge_softmax_backward: bad sizes");
  if (n_rows == 0) return PYGB_OK;
  PYGB_REQUIRE(out && grad_out && rowptr && perm && grad_src,
               "pygb_edge_softmax_backward: null pointer");
  const unsigned grid = (unsigned)ceil_div(n_rows, kWarpsPerCta);
  edge_softmax_bwd_kernel<<<grid, kWarpsPerCta * 32, 0, as_stream(s)>>>(out, grad_out, (int)H,
                                                                       rowptr, perm, n_rows,
                                                                       grad_src);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

}  // extern "C"
