// Batched brute-force k-nearest-neighbour graph — the step before EdgeConv's propagate
// (`torch_cluster.knn_graph` in the reference: README.md:149-150,
// examples/pointnet++.py:39-41, DynamicEdgeConv A.11).
//
// One thread per query point, 64 queries per CTA.  The candidates of the clouds the CTA
// touches stream through shared memory in tiles of 64 points; every thread keeps the k best
// (squared distance, index) pairs seen so far in a sorted list — in SHARED memory, laid out
// [rank][thread] so that the dynamically indexed insertion is bank-conflict free (round 1
// kept it in a per-thread array, i.e. a 512-byte local-memory stack frame) — and inserts a
// candidate only when it beats the current k-th (kept in a register).  Candidates are visited in ascending index and a candidate
// displaces only on a strictly smaller distance, so ties resolve to the lower index — the
// order of a stable sort by (distance, index).  Distances are fp32, accumulated over the
// coordinates in order with fused multiply-adds.
#include "stream_common.cuh"

namespace pygb {
namespace {

constexpr int kKnnThreads = 64;   // 32 x 1,024 points -> 512 CTAs
constexpr int kKnnMaxK = 64;

// QREG: the query lives in 4 registers (D <= 4, point clouds); else in shared memory with
// an odd row stride (conflict-free).
template <bool QREG>
__global__ void __launch_bounds__(kKnnThreads)
knn_kernel(const float* __restrict__ x, int64_t ldx, int D, const int64_t* __restrict__ batch,
           const int64_t* __restrict__ ptr, int64_t n, int k, int loop, int* __restrict__ nbr) {
  extern __shared__ float sm[];
  const int qs = D | 1;                             // odd stride
  float* sq = sm;                                   // [64][qs] queries
  float* sc = sm + kKnnThreads * qs;                // [64][D] candidates (read as broadcasts)
  const int t = threadIdx.x;
  const int64_t q0 = (int64_t)blockIdx.x * kKnnThreads;
  const int64_t q = q0 + t;
  const bool live = q < n;
  const int64_t q_last = min(q0 + kKnnThreads, n) - 1;
  // candidate range of this thread, and the union over the CTA (batch is sorted)
  const int64_t my_lo = live ? ptr[batch[q]] : 0, my_hi = live ? ptr[batch[q] + 1] : 0;
  const int64_t lo = ptr[batch[q0]], hi = ptr[batch[q_last] + 1];
  for (int i = t; i < kKnnThreads * D; i += kKnnThreads) {
    const int r = i / D, d = i - r * D;
    sq[r * qs + d] = q0 + r < n ? x[(q0 + r) * ldx + d] : 0.f;
  }
  __syncthreads();
  float* bd = sc + kKnnThreads * D + t;             // bd[r * 64]: rank r of this thread's list
  int* bi = reinterpret_cast<int*>(sc + kKnnThreads * D + kKnnThreads * k) + t;
  int have = 0;
  float worst = 3.402823466e38f;                    // k-th best so far (register copy of bd[k-1])
  const float* myq = sq + t * qs;
  float qr[4] = {0.f, 0.f, 0.f, 0.f};
  if (QREG)
    for (int d = 0; d < D; ++d) qr[d] = myq[d];
  for (int64_t c0 = lo; c0 < hi; c0 += kKnnThreads) {
    __syncthreads();
    const int nc = (int)min((int64_t)kKnnThreads, hi - c0);
    for (int i = t; i < nc * D; i += kKnnThreads) {
      const int r = i / D, d = i - r * D;
      sc[i] = x[(c0 + r) * ldx + d];
    }
    __syncthreads();
    if (!live) continue;
    for (int c = 0; c < nc; ++c) {
      const int64_t j = c0 + c;
      if (j < my_lo || j >= my_hi || (!loop && j == q)) continue;
      float dist = 0.f;
      const float* cc = sc + c * D;
      if (QREG) {
#pragma unroll
        for (int d = 0; d < 4; ++d) {
          if (d < D) {
            const float df = qr[d] - cc[d];
            dist = fmaf(df, df, dist);
          }
        }
      } else {
        for (int d = 0; d < D; ++d) {
          const float df = myq[d] - cc[d];
          dist = fmaf(df, df, dist);
        }
      }
      if (have == k && !(dist < worst)) continue;
      int pos = have < k ? have : k - 1;            // insertion: shift the worse entries down
      while (pos > 0 && dist < bd[(pos - 1) * kKnnThreads]) {
        bd[pos * kKnnThreads] = bd[(pos - 1) * kKnnThreads];
        bi[pos * kKnnThreads] = bi[(pos - 1) * kKnnThreads];
        --pos;
      }
      bd[pos * kKnnThreads] = dist;
      bi[pos * kKnnThreads] = (int)j;
      if (have < k) ++have;
      if (have == k) worst = bd[(k - 1) * kKnnThreads];
    }
  }
  if (live) {
    for (int r = 0; r < k; ++r) nbr[q * k + r] = r < have ? bi[r * kKnnThreads] : -1;
  }
}

}  // namespace
}  // namespace pygb

using namespace pygb;

extern "C" {

int pygb_knn_graph(const float* x, int64_t ldx, int64_t n, int D, const int64_t* batch,
                   const int64_t* ptr, int k, int loop, int32_t* nbr, pygb_stream_t s) {
  PYGB_REQUIRE(n >= 0 && D > 0 && k > 0 && ldx >= D && n < (1ll << 31) - 256,
               "pygb_knn_graph: bad sizes");
  if (k > kKnnMaxK || D > 128) {
    set_error("pygb_knn_graph: needs k <= %d and at most 128 coordinates", kKnnMaxK);
    return PYGB_E_UNSUPPORTED;
  }
  if (n == 0) return PYGB_OK;
  PYGB_REQUIRE(x && batch && ptr && nbr, "pygb_knn_graph: null pointer");
  const size_t smem = (size_t)kKnnThreads * ((D | 1) + D + 2 * k) * sizeof(float);
  const unsigned grid = (unsigned)ceil_div(n, kKnnThreads);
  if (D <= 4) {
    if (smem > 48 * 1024)
      PYGB_CUDA(cudaFuncSetAttribute(knn_kernel<true>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    knn_kernel<true><<<grid, kKnnThreads, smem, as_stream(s)>>>(x, ldx, D, batch, ptr, n, k, loop,
                                                                nbr);
  } else {
    if (smem > 48 * 1024)
      PYGB_CUDA(cudaFuncSetAttribute(knn_kernel<false>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    knn_kernel<false><<<grid, kKnnThreads, smem, as_stream(s)>>>(x, ldx, D, batch, ptr, n, k,
                                                                 loop, nbr);
  }
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

}  // extern "C"
