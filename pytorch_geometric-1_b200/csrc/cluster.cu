// Farthest point sampling and fixed-radius neighbour search — the point-cloud graph
// construction next to k-NN (`torch_cluster.fps` / `radius` /
// `radius_graph` in the reference: examples/pointnet++.py:9,39-41,46-48, README.md:150).
// Batched brute force over small examples (1,024-point clouds): latency- and issue-bound
// integer/fp32 work, no tensor cores.
//
//  * fps: one CTA per example.  The running minimum squared distance of every point to the
//    selected set lives in a global scratch row (L2-resident); each iteration folds in the
//    last selected point and takes a block-wide argmax — ties to the LOWER index, so the
//    result is deterministic and reproducible on the CPU.
//  * radius: one warp per query.  Lanes scan the candidates of the query's example 32 at a
//    time, a ballot ranks the hits, and the first `max_nn` hits in ascending candidate index
//    are written — the order torch_cluster's per-thread loop produces.
#include "common.cuh"

namespace pygb {
namespace {

constexpr int kFpsThreads = 256;

__global__ void __launch_bounds__(kFpsThreads)
fps_kernel(const float* __restrict__ x, int64_t ldx, int D, const int64_t* __restrict__ ptr,
           const int64_t* __restrict__ out_ptr, const int64_t* __restrict__ start,
           float* __restrict__ dmin, int64_t* __restrict__ out) {
  extern __shared__ float cur_pt[];                 // [D] coordinates of the last selected point
  __shared__ float s_val[kFpsThreads / 32];
  __shared__ int s_idx[kFpsThreads / 32];
  __shared__ int s_cur;
  const int b = blockIdx.x, t = threadIdx.x;
  const int64_t p0 = ptr[b];
  const int nb = (int)(ptr[b + 1] - p0);
  const int64_t o0 = out_ptr[b];
  const int m = (int)(out_ptr[b + 1] - o0);
  if (nb <= 0 || m <= 0) return;
  int cur = (int)start[b];
  if (cur < 0 || cur >= nb) cur = 0;
  for (int it = 0; it < m; ++it) {
    if (t == 0) out[o0 + it] = p0 + cur;
    for (int d = t; d < D; d += kFpsThreads) cur_pt[d] = x[(p0 + cur) * ldx + d];
    __syncthreads();
    float best = -1.f;
    int bi = 0x7fffffff;
    for (int i = t; i < nb; i += kFpsThreads) {
      const float* xi = x + (p0 + i) * ldx;
      float dist = 0.f;
      for (int d = 0; d < D; ++d) {
        const float df = xi[d] - cur_pt[d];
        dist = fmaf(df, df, dist);
      }
      const float dm = it == 0 ? dist : fminf(dmin[p0 + i], dist);
      dmin[p0 + i] = dm;
      if (dm > best) {            // ascending i per thread: the first maximum is kept
        best = dm;
        bi = i;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(kFull, best, o);
      const int oi = __shfl_xor_sync(kFull, bi, o);
      if (ov > best || (ov == best && oi < bi)) {
        best = ov;
        bi = oi;
      }
    }
    if ((t & 31) == 0) {
      s_val[t >> 5] = best;
      s_idx[t >> 5] = bi;
    }
    __syncthreads();
    if (t == 0) {
      float bv = s_val[0];
      int bx = s_idx[0];
      for (int w = 1; w < kFpsThreads / 32; ++w)
        if (s_val[w] > bv || (s_val[w] == bv && s_idx[w] < bx)) {
          bv = s_val[w];
          bx = s_idx[w];
        }
      s_cur = bx == 0x7fffffff ? 0 : bx;
    }
    __syncthreads();
    cur = s_cur;
  }
}

constexpr int kRadWarps = 8;

__global__ void __launch_bounds__(kRadWarps * 32)
radius_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ y, int64_t ldy,
              int D, const int64_t* __restrict__ ptr_x, const int64_t* __restrict__ batch_y,
              int64_t n_y, float r2, int max_nn, int skip_same, int* __restrict__ nbr,
              int* __restrict__ cnt) {
  const int lane = threadIdx.x & 31;
  const int64_t q = (int64_t)blockIdx.x * kRadWarps + (threadIdx.x >> 5);
  if (q >= n_y) return;
  const int64_t b = batch_y[q];
  const int64_t lo = ptr_x[b], hi = ptr_x[b + 1];
  const float* yq = y + q * ldy;
  int found = 0;
  for (int64_t c0 = lo; c0 < hi && found < max_nn; c0 += 32) {
    const int64_t j = c0 + lane;
    bool hit = false;
    if (j < hi && !(skip_same && j == q)) {
      const float* xj = x + j * ldx;
      float dist = 0.f;
      for (int d = 0; d < D; ++d) {
        const float df = xj[d] - yq[d];
        dist = fmaf(df, df, dist);
      }
      hit = dist <= r2;
    }
    const unsigned m = __ballot_sync(kFull, hit);
    const int rank = found + __popc(m & ((1u << lane) - 1u));
    if (hit && rank < max_nn) nbr[q * max_nn + rank] = (int)j;
    found += __popc(m);
  }
  if (found > max_nn) found = max_nn;
  for (int rfill = found + lane; rfill < max_nn; rfill += 32) nbr[q * max_nn + rfill] = -1;
  if (lane == 0) cnt[q] = found;
}

}  // namespace
}  // namespace pygb

using namespace pygb;

extern "C" {

int pygb_fps(const float* x, int64_t ldx, int64_t n, int D, const int64_t* ptr,
             const int64_t* out_ptr, const int64_t* start, int64_t n_examples, float* dmin,
             int64_t* out, pygb_stream_t s) {
  PYGB_REQUIRE(n >= 0 && D > 0 && ldx >= D && n_examples >= 0 && D <= 4096,
               "pygb_fps: bad sizes");
  if (n == 0 || n_examples == 0) return PYGB_OK;
  PYGB_REQUIRE(x && ptr && out_ptr && start && dmin && out, "pygb_fps: null pointer");
  fps_kernel<<<(unsigned)n_examples, kFpsThreads, (size_t)D * sizeof(float), as_stream(s)>>>(
      x, ldx, D, ptr, out_ptr, start, dmin, out);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

int pygb_radius(const float* x, int64_t ldx, const float* y, int64_t ldy, int D,
                const int64_t* ptr_x, const int64_t* batch_y, int64_t n_y, float r,
                int max_num_neighbors, int skip_same_index, int32_t* nbr, int32_t* count,
                pygb_stream_t s) {
  PYGB_REQUIRE(n_y >= 0 && D > 0 && ldx >= D && ldy >= D && max_num_neighbors > 0 && r >= 0.f,
               "pygb_radius: bad arguments");
  if (n_y == 0) return PYGB_OK;
  PYGB_REQUIRE(x && y && ptr_x && batch_y && nbr && count, "pygb_radius: null pointer");
  const unsigned grid = (unsigned)ceil_div(n_y, kRadWarps);
  radius_kernel<<<grid, kRadWarps * 32, 0, as_stream(s)>>>(x, ldx, y, ldy, D, ptr_x, batch_y, n_y,
                                                          r * r, max_num_neighbors,
                                                          skip_same_index, nbr, count);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

}  // extern "C"
