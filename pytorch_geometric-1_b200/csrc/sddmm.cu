// Per-edge dot product over the destination-sorted graph (an SDDMM):
//     out[perm[p]] = < g[row(p), :], x[col[p], :] >        for every slot p
// This is the gradient of the fused weighted aggregation with respect to the edge weight,
// d/dw_e sum_i <g_i, sum_p w_p x_col(p)> = <g_dst(e), x_src(e)> (autograd of the
// `norm.view(-1,1) * x_j` message of GCNConv).  The reference gets it
// from autograd over two materialised [E,F] gathers; here nothing of size E*F exists.
//
// Work decomposition: the E slots are cut into chunks of kChunk consecutive slots; a warp
// takes chunks round-robin, finds the row of the chunk's first slot by binary search in
// rowptr (once per chunk) and walks.  A warp is 32/LPF lane groups, each handling one slot
// at a time: LPF lanes x 16-byte vectors cover a feature row (looping when F is wider), the
// dot product finishes with log2(LPF) shuffles.  g[row] is re-read per slot but consecutive
// slots share the row, so it stays in L1/L2; the x gather is the HBM stream (E*F*s bytes).
// HBM-bound integer/fp32 work: no tensor cores, no atomics (every edge is written once).
#include "common.cuh"

namespace pygb {
namespace {

constexpr int kChunk = 256;   // slots per work item
constexpr int kWarps = 8;     // per CTA

template <typename T, bool VEC>
__global__ void __launch_bounds__(kWarps * 32)
edge_dot_kernel(const T* __restrict__ g, int64_t ldg, const T* __restrict__ x, int64_t ldx, int F,
                const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                const int32_t* __restrict__ perm, int n_rows, int64_t E, int lpf,
                float* __restrict__ out) {
  constexpr int kV = Elem<T>::kVec;
  const int lane = threadIdx.x & 31;
  const int gl = lane % lpf;             // lane inside the group
  const int grp = lane / lpf;            // group inside the warp
  const int ngrp = 32 / lpf;
  const int64_t warp0 = (int64_t)blockIdx.x * kWarps + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * kWarps;
  const int64_t nchunks = (E + kChunk - 1) / kChunk;
  for (int64_t c = warp0; c < nchunks; c += nwarps) {
    const int64_t s0 = c * kChunk;
    const int64_t s1 = s0 + kChunk < E ? s0 + kChunk : E;
    // row of slot s0: largest r with rowptr[r] <= s0
    int lo = 0, hi = n_rows;             // invariant: rowptr[lo] <= s0 < rowptr[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if ((int64_t)__ldg(rowptr + mid) <= s0) lo = mid; else hi = mid;
    }
    int row = lo;
    int64_t row_end = __ldg(rowptr + row + 1);
    for (int64_t pb = s0; pb < s1; pb += ngrp) {   // uniform trip count: full-mask shuffles below
      const int64_t p = pb + grp;
      const bool live = p < s1;
      while (live && p >= row_end) {     // skips empty rows too
        ++row;
        row_end = __ldg(rowptr + row + 1);
      }
      const int j = live ? __ldg(col + p) : 0;
      const T* gr = g + (int64_t)row * ldg;
      const T* xr = x + (int64_t)j * ldx;
      float acc = 0.f;
      if (!live) {
      } else if (VEC) {
        for (int v = gl * kV; v < F; v += lpf * kV) {
          const uint4 a = *reinterpret_cast<const uint4*>(gr + v);   // reused by the next slots
          const uint4 b = ldg_nc_v4(xr + v);
          float fa[kV], fb[kV];
          Elem<T>::unpack(a, fa);
          Elem<T>::unpack(b, fb);
#pragma unroll
          for (int k = 0; k < kV; ++k) acc = fmaf(fa[k], fb[k], acc);
        }
      } else {
        for (int v = gl; v < F; v += lpf) acc = fmaf(Elem<T>::load(gr + v), Elem<T>::load(xr + v), acc);
      }
      for (int o = lpf >> 1; o > 0; o >>= 1) acc += __shfl_xor_sync(kFull, acc, o);
      if (live && gl == 0) out[__ldg(perm + p)] = acc;
    }
  }
}

template <typename T>
int launch_edge_dot(const void* g, int64_t ldg, const void* x, int64_t ldx, int64_t F,
                    const int32_t* rowptr, const int32_t* col, const int32_t* perm, int64_t n_rows,
                    int64_t E, float* out, cudaStream_t st) {
  constexpr int kV = Elem<T>::kVec;
  const bool vec = F % kV == 0 && ldg % kV == 0 && ldx % kV == 0 && aligned16(g) && aligned16(x);
  const int64_t units = vec ? F / kV : F;      // lanes that can be busy on one row
  int lpf = 1;
  while (lpf < 32 && lpf < units) lpf <<= 1;
  int sms = 148, dev = 0;
  PYGB_CUDA(cudaGetDevice(&dev));
  PYGB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  int64_t blocks = ceil_div(ceil_div(E, kChunk), kWarps);
  if (blocks > (int64_t)sms * 8) blocks = (int64_t)sms * 8;   // whole waves of resident CTAs
  if (blocks < 1) blocks = 1;
  if (vec)
    edge_dot_kernel<T, true><<<(unsigned)blocks, kWarps * 32, 0, st>>>(
        static_cast<const T*>(g), ldg, static_cast<const T*>(x), ldx, (int)F, rowptr, col, perm,
        (int)n_rows, E, lpf, out);
  else
    edge_dot_kernel<T, false><<<(unsigned)blocks, kWarps * 32, 0, st>>>(
        static_cast<const T*>(g), ldg, static_cast<const T*>(x), ldx, (int)F, rowptr, col, perm,
        (int)n_rows, E, lpf, out);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

}  // namespace
}  // namespace pygb

using namespace pygb;

extern "C" int pygb_edge_dot(int dtype, const void* g, int64_t ldg, const void* x, int64_t ldx,
                             int64_t F, const int32_t* rowptr, const int32_t* col,
                             const int32_t* perm, int64_t n_rows, int64_t E, float* out,
                             pygb_stream_t s) {
  PYGB_REQUIRE(dtype == PYGB_F32 || dtype == PYGB_BF16, "pygb_edge_dot: bad dtype %d", dtype);
  PYGB_REQUIRE(F >= 0 && n_rows >= 0 && E >= 0 && ldg >= F && ldx >= F && F < (1ll << 31) &&
                   n_rows < (1ll << 31) && E < (1ll << 31),
               "pygb_edge_dot: bad sizes");
  if (E == 0) return PYGB_OK;
  PYGB_REQUIRE(rowptr && col && perm && out && n_rows > 0, "pygb_edge_dot: null pointer");
  cudaStream_t st = as_stream(s);
  if (F == 0) {
    PYGB_CUDA(cudaMemsetAsync(out, 0, (size_t)E * sizeof(float), st));
    return PYGB_OK;
  }
  PYGB_REQUIRE(g && x, "pygb_edge_dot: null pointer");
  if (dtype == PYGB_F32)
    return launch_edge_dot<float>(g, ldg, x, ldx, F, rowptr, col, perm, n_rows, E, out, st);
  return launch_edge_dot<__nv_bfloat16>(g, ldg, x, ldx, F, rowptr, col, perm, n_rows, E, out, st);
}
