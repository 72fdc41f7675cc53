// K4 — graph build: self-loop handling, stable sort-by-destination into CSR,
// degrees and GCN normalisation.  All integer outputs are bit-exact with the
// CPU oracle (oracle/ref_cpu.py).  HBM-bound integer work:
// coalesced tiles, shared-memory ranking, integer atomics only (deterministic).
#include "common.cuh"

namespace pygb {
namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kItems = 16;
constexpr int kTile = kThreads * kItems;  // 4096 elements per CTA

// ---------------------------------------------------------------------------
// device-wide exclusive scan of int32 (three launches; out[n] = total)
// ---------------------------------------------------------------------------
__device__ __forceinline__ int warp_incl_scan(int v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(kFull, v, o);
    if (lane >= o) v += t;
  }
  return v;
}

// Block-wide exclusive scan of one value per thread; returns exclusive prefix,
// *total = block sum.  smem: kWarps ints.
__device__ __forceinline__ int block_excl_scan(int v, int* smem, int* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = warp_incl_scan(v, lane);
  if (lane == 31) smem[warp] = incl;
  __syncthreads();
  int wbase = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < kWarps; ++w) {
    int s = smem[w];
    if (w < warp) wbase += s;
    tot += s;
  }
  __syncthreads();
  *total = tot;
  return wbase + incl - v;
}

__global__ void __launch_bounds__(kThreads) scan_tile_sums(const int* __restrict__ in,
                                                           int64_t n, int* __restrict__ tsum) {
  __shared__ int smem[kWarps];
  const int64_t base = (int64_t)blockIdx.x * kTile;
  int s = 0;
#pragma unroll
  for (int k = 0; k < kItems; ++k) {
    int64_t i = base + (int64_t)k * kThreads + threadIdx.x;
    if (i < n) s += in[i];
  }
  int tot;
  block_excl_scan(s, smem, &tot);
  if (threadIdx.x == 0) tsum[blockIdx.x] = tot;
}

// single CTA: in-place exclusive scan of m ints; writes total to *total_out if non-null
__global__ void __launch_bounds__(1024) scan_single_block(int* __restrict__ data, int64_t m,
                                                          int* __restrict__ total_out) {
  __shared__ int wsum[32];
  __shared__ int carry_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int64_t base = 0; base < m; base += 1024) {
    int64_t i = base + threadIdx.x;
    int v = (i < m) ? data[i] : 0;
    int incl = warp_incl_scan(v, lane);
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int t = wsum[lane];
      int ti = warp_incl_scan(t, lane);
      wsum[lane] = ti - t;  // exclusive warp bases
    }
    __syncthreads();
    int carry = carry_s;
    int excl = carry + wsum[warp] + incl - v;
    if (i < m) data[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0 && total_out) *total_out = carry_s;
}

// out[i] = exclusive prefix; tile order: element (k, t) -> index base + k*kThreads + t
__global__ void __launch_bounds__(kThreads) scan_apply(const int* __restrict__ in, int64_t n,
                                                       const int* __restrict__ tsum,
                                                       int* __restrict__ out) {
  __shared__ int smem[kWarps];
  const int64_t base = (int64_t)blockIdx.x * kTile;
  int running = tsum[blockIdx.x];
#pragma unroll 4
  for (int k = 0; k < kItems; ++k) {
    int64_t i = base + (int64_t)k * kThreads + threadIdx.x;
    int v = (i < n) ? in[i] : 0;
    int tot;
    int ex = block_excl_scan(v, smem, &tot);
    if (i < n) out[i] = running + ex;
    running += tot;
  }
}

}  // namespace

// Exclusive scan in[0..n) -> out[0..n], out[n] = total.  ws_tiles: ceil(n/kTile) ints.
int exclusive_scan(const int* in, int* out, int64_t n, int* ws_tiles, cudaStream_t st) {
  int64_t nt = ceil_div(n, kTile);
  if (n == 0) {
    PYGB_CUDA(cudaMemsetAsync(out, 0, sizeof(int), st));
    return PYGB_OK;
  }
  scan_tile_sums<<<(unsigned)nt, kThreads, 0, st>>>(in, n, ws_tiles);
  PYGB_LAUNCH_CHECK();
  scan_single_block<<<1, 1024, 0, st>>>(ws_tiles, nt, out + n);
  PYGB_LAUNCH_CHECK();
  scan_apply<<<(unsigned)nt, kThreads, 0, st>>>(in, n, ws_tiles, out);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

int64_t exclusive_scan_tiles(int64_t n) { return ceil_div(n > 0 ? n : 1, kTile); }

namespace {

// ---------------------------------------------------------------------------
// self loops
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) count_loops_kernel(const int64_t* __restrict__ row,
                                                               const int64_t* __restrict__ col,
                                                               int64_t E,
                                                               unsigned long long* __restrict__ out) {
  int c = 0;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < E;
       i += (int64_t)gridDim.x * kThreads)
    c += (row[i] == col[i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(kFull, c, o);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, (unsigned long long)c);
}

// out[0] += edges with row == col, out[1] += edges with an id outside [0, N)
__global__ void __launch_bounds__(kThreads) scan_edges_kernel(const int64_t* __restrict__ row,
                                                              const int64_t* __restrict__ col,
                                                              int64_t E, int64_t N,
                                                              unsigned long long* __restrict__ out) {
  int c = 0, bad = 0;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < E;
       i += (int64_t)gridDim.x * kThreads) {
    const int64_t r = row[i], q = col[i];
    c += (r == q);
    bad += (r < 0 || r >= N || q < 0 || q >= N);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    c += __shfl_xor_sync(kFull, c, o);
    bad += __shfl_xor_sync(kFull, bad, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (c) atomicAdd(out, (unsigned long long)c);
    if (bad) atomicAdd(out + 1, (unsigned long long)bad);
  }
}

// per tile: number of kept (row != col) edges
__global__ void __launch_bounds__(kThreads) keep_count_kernel(const int64_t* __restrict__ row,
                                                              const int64_t* __restrict__ col,
                                                              int64_t E, int* __restrict__ tcount) {
  __shared__ int smem[kWarps];
  const int64_t base = (int64_t)blockIdx.x * kTile;
  int s = 0;
#pragma unroll
  for (int k = 0; k < kItems; ++k) {
    int64_t i = base + (int64_t)k * kThreads + threadIdx.x;
    if (i < E) s += (row[i] != col[i]);
  }
  int tot;
  block_excl_scan(s, smem, &tot);
  if (threadIdx.x == 0) tcount[blockIdx.x] = tot;
}

// order-preserving compaction of non-loop edges
__global__ void __launch_bounds__(kThreads)
compact_kernel(const int64_t* __restrict__ row, const int64_t* __restrict__ col,
               const float* __restrict__ w, int64_t E, const int* __restrict__ toff,
               int64_t Eout, int64_t* __restrict__ orow, int64_t* __restrict__ ocol,
               float* __restrict__ ow) {
  __shared__ int smem[kWarps];
  const int64_t base = (int64_t)blockIdx.x * kTile;
  int64_t running = toff[blockIdx.x];
  for (int k = 0; k < kItems; ++k) {
    int64_t i = base + (int64_t)k * kThreads + threadIdx.x;
    int64_t r = 0, c = 0;
    int keep = 0;
    if (i < E) {
      r = row[i];
      c = col[i];
      keep = (r != c);
    }
    int tot;
    int ex = block_excl_scan(keep, smem, &tot);
    if (keep) {
      int64_t o = running + ex;
      orow[o] = r;
      ocol[o] = c;
      if (ow) ow[o] = w[i];
    }
    running += tot;
  }
  (void)Eout;
}

__global__ void __launch_bounds__(kThreads) fill_i32_kernel(int* p, int64_t n, int v) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * kThreads)
    p[i] = v;
}

// last (highest id) self-loop edge per node  (Q1: last write wins)
__global__ void __launch_bounds__(kThreads) last_loop_kernel(const int64_t* __restrict__ row,
                                                             const int64_t* __restrict__ col,
                                                             int64_t E, int64_t N,
                                                             int* __restrict__ last) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < E;
       i += (int64_t)gridDim.x * kThreads) {
    int64_t r = row[i];
    if (r == col[i] && r >= 0 && r < N) atomicMax(&last[r], (int)i);
  }
}

__global__ void __launch_bounds__(kThreads)
append_loops_kernel(int64_t N, int64_t off, int64_t Eout, const int* __restrict__ last,
                    const float* __restrict__ w, float fill, int64_t* __restrict__ out_ei,
                    float* __restrict__ ow) {
  for (int64_t n = (int64_t)blockIdx.x * kThreads + threadIdx.x; n < N;
       n += (int64_t)gridDim.x * kThreads) {
    out_ei[off + n] = n;
    out_ei[Eout + off + n] = n;
    if (ow) ow[off + n] = (last && last[n] >= 0) ? w[last[n]] : fill;
  }
}

inline unsigned grid_for(int64_t n, int cap = 148 * 16) {
  int64_t b = ceil_div(n, kThreads);
  if (b < 1) b = 1;
  if (b > cap) b = cap;
  return (unsigned)b;
}

// ---------------------------------------------------------------------------
// stable LSD radix sort by key (8-bit digits) — the CSR build
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
make_keys_kernel(const int64_t* __restrict__ key, int64_t E, int64_t n_rows,
                 uint32_t* __restrict__ k32, int* __restrict__ cnt, int* __restrict__ status) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < E;
       i += (int64_t)gridDim.x * kThreads) {
    int64_t k = key[i];
    if (k < 0 || k >= n_rows) {
      if (status) *status = 1;
      k32[i] = (uint32_t)n_rows;  // sorts behind every valid row
    } else {
      k32[i] = (uint32_t)k;
      atomicAdd(&cnt[k], 1);
    }
  }
}

// hist[d * nblk + b] = number of keys in tile b whose digit is d
__global__ void __launch_bounds__(kThreads)
radix_hist_kernel(const uint32_t* __restrict__ keys, int64_t E, int shift, int nblk,
                  int* __restrict__ hist) {
  __shared__ int h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * kTile;
#pragma unroll
  for (int k = 0; k < kItems; ++k) {
    int64_t i = base + (int64_t)k * kThreads + threadIdx.x;
    if (i < E) atomicAdd(&h[(keys[i] >> shift) & 255], 1);
  }
  __syncthreads();
  hist[(int64_t)threadIdx.x * nblk + blockIdx.x] = h[threadIdx.x];
}

// Stable scatter.  Warp w owns the contiguous 512-element slice of the tile and
// walks it 32 keys at a time; __match_any_sync ranks equal digits inside the
// step, a per-warp counter row in shared memory carries the rank across steps,
// and a per-digit prefix over the warps (in warp order) makes the tile stable.
// The tile is then REORDERED IN SHARED MEMORY (position = digit offset inside the
// tile + rank) and leaves as runs of consecutive addresses per digit — whole
// sectors instead of one 4-byte store per key (1.77 -> 0.6 ms per pass at cfg5).
// Payloads: `vals` (the permutation; identity in the first pass) and, when the
// caller wants `col`, a second int32 payload read from `other64` in the first
// pass (range-checked there) and from `col_in` afterwards — this replaces the
// random 8-byte gather other[perm[p]] after the sort.
template <bool HAS_COL>
__global__ void __launch_bounds__(kThreads)
radix_scatter_kernel(const uint32_t* __restrict__ keys, const int* __restrict__ vals,
                     const int* __restrict__ col_in, const int64_t* __restrict__ other64,
                     int64_t n_cols, int* __restrict__ status, int64_t E, int shift, int nblk,
                     const int* __restrict__ gbase, uint32_t* __restrict__ keys_out,
                     int* __restrict__ vals_out, int* __restrict__ col_out) {
  __shared__ int wcnt[kWarps][256];
  __shared__ int dig_off[256];   // first tile-local position of digit d
  __shared__ int dig_gb[256];    // global position of the tile's first key of digit d
  __shared__ int scan_tmp[kWarps];
  extern __shared__ __align__(16) uint32_t s_dyn[];  // (HAS_COL ? 3 : 2) * kTile words
  uint32_t* s_key = s_dyn;
  int* s_val = reinterpret_cast<int*>(s_dyn + kTile);
  int* s_col = reinterpret_cast<int*>(s_dyn + 2 * kTile);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < kWarps * 256; i += kThreads) (&wcnt[0][0])[i] = 0;
  __syncthreads();
  const int64_t tbase = (int64_t)blockIdx.x * kTile;
  const int64_t wbase = tbase + (int64_t)warp * (32 * kItems);
  const unsigned lt = (1u << lane) - 1u;
  // every global read of the tile is issued up front (48 independent loads per thread in
  // flight) so that their latency is paid once, under the ranking
  uint32_t mykey[kItems];
  int myval[kItems], mycol[HAS_COL ? kItems : 1];
#pragma unroll
  for (int s = 0; s < kItems; ++s) {
    const int64_t i = wbase + s * 32 + lane;
    mykey[s] = i < E ? keys[i] : 0u;
  }
#pragma unroll
  for (int s = 0; s < kItems; ++s) {
    const int64_t i = wbase + s * 32 + lane;
    myval[s] = (vals && i < E) ? vals[i] : (int)i;
  }
  if (HAS_COL) {
    if (other64) {
      bool bad = false;
#pragma unroll
      for (int s = 0; s < kItems; ++s) {
        const int64_t i = wbase + s * 32 + lane;
        int64_t c64 = i < E ? other64[i] : 0;
        if (c64 < 0 || c64 >= n_cols) {
          bad = true;
          c64 = 0;  // keep later gathers in bounds; the caller raises on status
        }
        mycol[s] = (int)c64;
      }
      if (bad && status) *status = 2;
    } else {
#pragma unroll
      for (int s = 0; s < kItems; ++s) {
        const int64_t i = wbase + s * 32 + lane;
        mycol[s] = i < E ? col_in[i] : 0;
      }
    }
  }
  uint32_t rank16[kItems / 2];  // rank inside the warp slice, two per register
#pragma unroll
  for (int s = 0; s < kItems; ++s) {
    int64_t i = wbase + s * 32 + lane;
    bool valid = i < E;
    unsigned d = valid ? ((mykey[s] >> shift) & 255u) : 256u + lane;  // invalid lanes: unique groups
    unsigned peers = __match_any_sync(kFull, d);
    int leader = __ffs(peers) - 1;
    int c = 0;
    if (valid && lane == leader) {
      c = wcnt[warp][d];
      wcnt[warp][d] = c + __popc(peers);
    }
    c = __shfl_sync(kFull, c, leader);
    const uint32_t r = (uint32_t)(c + __popc(peers & lt));  // < 512
    if (s & 1) rank16[s >> 1] |= r << 16;
    else rank16[s >> 1] = r;
    __syncwarp();
  }
  __syncthreads();
  int tile_cnt;
  {  // thread d: per-warp counts -> offsets of each warp's keys inside the tile's digit-d group
    const int d = threadIdx.x;
    int running = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      int t = wcnt[w][d];
      wcnt[w][d] = running;
      running += t;
    }
    tile_cnt = running;
    dig_gb[d] = gbase[(int64_t)d * nblk + blockIdx.x];
  }
  int total;
  const int off = block_excl_scan(tile_cnt, scan_tmp, &total);  // syncs inside
  dig_off[threadIdx.x] = off;
  __syncthreads();
#pragma unroll
  for (int s = 0; s < kItems; ++s) {
    int64_t i = wbase + s * 32 + lane;
    if (i < E) {
      const unsigned d = (mykey[s] >> shift) & 255u;
      const int r = (int)((rank16[s >> 1] >> ((s & 1) * 16)) & 0xffffu);
      const int lp = dig_off[d] + wcnt[warp][d] + r;
      s_key[lp] = mykey[s];
      s_val[lp] = myval[s];
      if (HAS_COL) s_col[lp] = mycol[s];
    }
  }
  __syncthreads();
  // `total` valid keys, sorted by digit: consecutive threads -> consecutive global addresses
#pragma unroll
  for (int k = 0; k < kItems; ++k) {
    const int lp = k * kThreads + threadIdx.x;
    if (lp < total) {
      const uint32_t key = s_key[lp];
      const unsigned d = (key >> shift) & 255u;
      const int pos = dig_gb[d] + (lp - dig_off[d]);
      if (keys_out) keys_out[pos] = key;
      vals_out[pos] = s_val[lp];
      if (HAS_COL) col_out[pos] = s_col[lp];
    }
  }
}

__global__ void __launch_bounds__(kThreads)
permute_f32_kernel(const float* __restrict__ src, const int* __restrict__ perm, int64_t E,
                   float* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < E;
       i += (int64_t)gridDim.x * kThreads)
    out[i] = src[perm[i]];
}

// ---------------------------------------------------------------------------
// degree / GCN norm
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
degree_count_kernel(const int64_t* __restrict__ index, int64_t E, int64_t N,
                    unsigned long long* __restrict__ out64, int* __restrict__ out32) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < E;
       i += (int64_t)gridDim.x * kThreads) {
    int64_t k = index[i];
    if (k >= 0 && k < N) {
      if (out64) atomicAdd(&out64[k], 1ull);
      else atomicAdd(&out32[k], 1);
    }
  }
}

__global__ void __launch_bounds__(kThreads)
i32_to_f32_kernel(int* __restrict__ p, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * kThreads) {
    int v = p[i];
    reinterpret_cast<float*>(p)[i] = (float)v;
  }
}

// deg[j] = w[perm[p]] summed for p ascending (== ascending edge id: the CPU
// reference's scatter_add_ order) — sequential per row so the fp32 result is
// bit-identical to the oracle's.
__global__ void __launch_bounds__(kThreads)
weighted_degree_kernel(const float* __restrict__ w, const int* __restrict__ rowptr,
                       const int* __restrict__ perm, int64_t N, float* __restrict__ deg) {
  for (int64_t j = (int64_t)blockIdx.x * kThreads + threadIdx.x; j < N;
       j += (int64_t)gridDim.x * kThreads) {
    float s = 0.f;
    const int b = rowptr[j], e = rowptr[j + 1];
    for (int p = b; p < e; ++p) s = __fadd_rn(s, w[perm[p]]);
    deg[j] = s;
  }
}

__device__ __forceinline__ float inv_sqrt_deg(float d) {
  // deg.pow(-0.5) with inf -> 0 (A.9).  1/sqrt in round-to-nearest.
  float r = __frcp_rn(__fsqrt_rn(d));
  return (d == 0.f) ? 0.f : r;
}

__global__ void __launch_bounds__(kThreads)
gcn_norm_kernel(const int64_t* __restrict__ row, const int64_t* __restrict__ col,
                const float* __restrict__ w, int64_t E, const float* __restrict__ deg,
                float* __restrict__ norm) {
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < E;
       i += (int64_t)gridDim.x * kThreads) {
    float dr = inv_sqrt_deg(deg[row[i]]);
    float dc = inv_sqrt_deg(deg[col[i]]);
    float ww = w ? w[i] : 1.f;
    norm[i] = __fmul_rn(__fmul_rn(dr, ww), dc);
  }
}

// The same product, written directly in the slot order of a sorted graph: slot p of row i
// has the end points (i, col[p]); one warp per row.  Bit-identical to norm[perm[p]] for unit
// weights (fmul is commutative), without the random 4-byte gather.
__global__ void __launch_bounds__(kThreads)
gcn_norm_csr_kernel(const int* __restrict__ rowptr, const int* __restrict__ col,
                    const float* __restrict__ deg, int64_t n_rows, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t nw = (int64_t)gridDim.x * kWarps;
  for (int64_t r = (int64_t)blockIdx.x * kWarps + (threadIdx.x >> 5); r < n_rows; r += nw) {
    const int b = rowptr[r], e = rowptr[r + 1];
    if (b == e) continue;
    const float di = inv_sqrt_deg(deg[r]);
    for (int p = b + lane; p < e; p += 32) out[p] = __fmul_rn(inv_sqrt_deg(deg[col[p]]), di);
  }
}

}  // namespace
}  // namespace pygb

using namespace pygb;

extern "C" {

int pygb_count_self_loops(const int64_t* edge_index, int64_t E, int64_t* d_count,
                          pygb_stream_t s) {
  PYGB_REQUIRE(E >= 0 && d_count, "pygb_count_self_loops: bad arguments");
  cudaStream_t st = as_stream(s);
  PYGB_CUDA(cudaMemsetAsync(d_count, 0, sizeof(int64_t), st));
  if (E == 0) return PYGB_OK;
  count_loops_kernel<<<grid_for(E), kThreads, 0, st>>>(
      edge_index, edge_index + E, E, reinterpret_cast<unsigned long long*>(d_count));
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

int pygb_scan_edge_index(const int64_t* edge_index, int64_t E, int64_t N, int64_t* d_out,
                         pygb_stream_t s) {
  PYGB_REQUIRE(E >= 0 && N >= 0 && d_out, "pygb_scan_edge_index: bad arguments");
  cudaStream_t st = as_stream(s);
  PYGB_CUDA(cudaMemsetAsync(d_out, 0, 2 * sizeof(int64_t), st));
  if (E == 0) return PYGB_OK;
  scan_edges_kernel<<<grid_for(E), kThreads, 0, st>>>(
      edge_index, edge_index + E, E, N, reinterpret_cast<unsigned long long*>(d_out));
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

size_t pygb_apply_loops_workspace(int64_t E, int64_t N) {
  size_t tiles = (size_t)ceil_div(E > 0 ? E : 1, kTile) + 1;
  return align_up(tiles * sizeof(int), 256) + align_up((size_t)(N > 0 ? N : 1) * sizeof(int), 256) + 512;
}

int pygb_apply_loops(const int64_t* edge_index, int64_t E, int64_t N, int mode, int64_t n_loops,
                     const float* edge_weight, float fill_value, int64_t* out_edge_index,
                     float* out_edge_weight, void* ws, size_t ws_bytes, pygb_stream_t s) {
  PYGB_REQUIRE(E >= 0 && N >= 0 && mode >= 0 && mode <= 3, "pygb_apply_loops: bad arguments");
  PYGB_REQUIRE(edge_weight == nullptr || out_edge_weight != nullptr,
               "pygb_apply_loops: out_edge_weight missing");
  PYGB_REQUIRE(out_edge_weight == nullptr || edge_weight != nullptr || E == 0,
               "pygb_apply_loops: edge_weight missing");
  PYGB_REQUIRE(E < (1ll << 31) && N < (1ll << 31), "pygb_apply_loops: E, N must be < 2^31");
  cudaStream_t st = as_stream(s);
  if (ws_bytes < pygb_apply_loops_workspace(E, N)) {
    set_error("pygb_apply_loops: workspace too small");
    return PYGB_E_WORKSPACE;
  }
  Workspace W(ws, ws_bytes);
  const int64_t nt = ceil_div(E, kTile);
  int* toff = W.take<int>((size_t)nt + 1);
  int* last = W.take<int>((size_t)(N > 0 ? N : 1));
  const bool remove = (mode != PYGB_LOOPS_ADD);
  const bool add = (mode != PYGB_LOOPS_REMOVE);
  const int64_t keep = remove ? E - n_loops : E;
  const int64_t Eout = keep + (add ? N : 0);
  const int64_t* row = edge_index;
  const int64_t* col = edge_index + E;
  const float* w_in = edge_weight;
  float* w_out = out_edge_weight;

  if (remove) {
    if (E > 0) {
      keep_count_kernel<<<(unsigned)nt, kThreads, 0, st>>>(row, col, E, toff);
      PYGB_LAUNCH_CHECK();
      scan_single_block<<<1, 1024, 0, st>>>(toff, nt, nullptr);
      PYGB_LAUNCH_CHECK();
      compact_kernel<<<(unsigned)nt, kThreads, 0, st>>>(row, col, w_in, E, toff, Eout,
                                                       out_edge_index, out_edge_index + Eout,
                                                       w_out);
      PYGB_LAUNCH_CHECK();
    }
  } else if (E > 0) {
    PYGB_CUDA(cudaMemcpyAsync(out_edge_index, row, E * sizeof(int64_t), cudaMemcpyDeviceToDevice, st));
    PYGB_CUDA(cudaMemcpyAsync(out_edge_index + Eout, col, E * sizeof(int64_t),
                              cudaMemcpyDeviceToDevice, st));
    if (w_out)
      PYGB_CUDA(cudaMemcpyAsync(w_out, w_in, E * sizeof(float), cudaMemcpyDeviceToDevice, st));
  }
  if (add && N > 0) {
    const int* last_p = nullptr;
    if (mode == PYGB_LOOPS_ADD_REMAINING && w_out && n_loops > 0) {
      fill_i32_kernel<<<grid_for(N), kThreads, 0, st>>>(last, N, -1);
      PYGB_LAUNCH_CHECK();
      last_loop_kernel<<<grid_for(E), kThreads, 0, st>>>(row, col, E, N, last);
      PYGB_LAUNCH_CHECK();
      last_p = last;
    }
    append_loops_kernel<<<grid_for(N), kThreads, 0, st>>>(N, keep, Eout, last_p, w_in, fill_value,
                                                         out_edge_index, w_out);
    PYGB_LAUNCH_CHECK();
  }
  return PYGB_OK;
}

size_t pygb_csr_build_workspace(int64_t E, int64_t n_rows) {
  size_t e = (size_t)(E > 0 ? E : 1);
  size_t nblk = (size_t)ceil_div((int64_t)e, kTile);
  size_t hist = 256 * nblk;
  size_t bytes = 0;
  bytes += 2 * align_up(e * sizeof(uint32_t), 256);                      // keys ping-pong
  bytes += 2 * align_up(e * sizeof(int), 256);                          // vals pong, col pong
  bytes += 2 * align_up((hist + 1) * sizeof(int), 256);                 // hist + scanned
  bytes += align_up(((size_t)ceil_div((int64_t)hist + 1, kTile) + 1) * sizeof(int), 256);
  bytes += align_up(((size_t)n_rows + 2) * sizeof(int), 256);           // counts
  bytes += align_up(((size_t)ceil_div(n_rows + 1, kTile) + 1) * sizeof(int), 256);
  return bytes + 1024;
}

int pygb_csr_build(const int64_t* key, const int64_t* other, int64_t E, int64_t n_rows,
                   int64_t n_cols, int32_t* rowptr, int32_t* perm, int32_t* col,
                   int32_t* d_status, void* ws, size_t ws_bytes, pygb_stream_t s) {
  PYGB_REQUIRE(E >= 0 && n_rows >= 0 && rowptr, "pygb_csr_build: bad arguments");
  PYGB_REQUIRE(E < (1ll << 31) - kTile && n_rows < (1ll << 31) - 2,
               "pygb_csr_build: E and n_rows must be < 2^31");
  PYGB_REQUIRE(E == 0 || (key && perm), "pygb_csr_build: key/perm missing");
  PYGB_REQUIRE(!(other && !col), "pygb_csr_build: col missing");
  cudaStream_t st = as_stream(s);
  if (ws_bytes < pygb_csr_build_workspace(E, n_rows)) {
    set_error("pygb_csr_build: workspace too small (%zu < %zu)", ws_bytes,
              pygb_csr_build_workspace(E, n_rows));
    return PYGB_E_WORKSPACE;
  }
  Workspace W(ws, ws_bytes);
  const size_t e = (size_t)(E > 0 ? E : 1);
  const int nblk = (int)ceil_div((int64_t)e, kTile);
  const size_t hist_n = (size_t)256 * nblk;
  uint32_t* keysA = W.take<uint32_t>(e);
  uint32_t* keysB = W.take<uint32_t>(e);
  int* valsB = W.take<int>(e);
  int* colB = W.take<int>(e);
  int* hist = W.take<int>(hist_n + 1);
  int* hscan = W.take<int>(hist_n + 1);
  int* htiles = W.take<int>((size_t)ceil_div((int64_t)hist_n + 1, kTile) + 1);
  int* cnt = W.take<int>((size_t)n_rows + 2);
  int* ctiles = W.take<int>((size_t)ceil_div(n_rows + 1, kTile) + 1);
  if (!W.ok()) {
    set_error("pygb_csr_build: internal workspace accounting error");
    return PYGB_E_WORKSPACE;
  }

  PYGB_CUDA(cudaMemsetAsync(cnt, 0, ((size_t)n_rows + 1) * sizeof(int), st));
  if (E == 0) {
    PYGB_CUDA(cudaMemsetAsync(rowptr, 0, ((size_t)n_rows + 1) * sizeof(int), st));
    return PYGB_OK;
  }
  make_keys_kernel<<<grid_for(E), kThreads, 0, st>>>(key, E, n_rows, keysA, cnt, d_status);
  PYGB_LAUNCH_CHECK();
  int rc = exclusive_scan(cnt, rowptr, n_rows, ctiles, st);
  if (rc) return rc;

  int bits = 1;
  while ((1ll << bits) <= n_rows) ++bits;  // keys lie in [0, n_rows]
  const int passes = (bits + 7) / 8;
  // vals ping-pong arranged so that the last pass lands in `perm`.
  const uint32_t* kin = keysA;
  uint32_t* kout = keysB;
  if (other)  // 48 KB of staging + 10 KB of counters: above the default 48 KB per CTA
    PYGB_CUDA(cudaFuncSetAttribute(radix_scatter_kernel<true>,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)(3 * kTile * sizeof(int))));
  const int* vin = nullptr;  // first pass: identity
  const int* cin = nullptr;  // first pass: read (and range-check) other64
  for (int p = 0; p < passes; ++p) {
    const bool last = (p == passes - 1);
    int* vout = ((passes - 1 - p) % 2 == 0) ? perm : valsB;
    int* cout = ((passes - 1 - p) % 2 == 0) ? col : colB;
    radix_hist_kernel<<<nblk, kThreads, 0, st>>>(kin, E, 8 * p, nblk, hist);
    PYGB_LAUNCH_CHECK();
    rc = exclusive_scan(hist, hscan, (int64_t)hist_n, htiles, st);
    if (rc) return rc;
    if (other) {
      radix_scatter_kernel<true><<<nblk, kThreads, 3 * kTile * sizeof(int), st>>>(
          kin, vin, cin, p == 0 ? other : nullptr, n_cols, d_status, E, 8 * p, nblk, hscan,
          last ? nullptr : kout, vout, cout);
    } else {
      radix_scatter_kernel<false><<<nblk, kThreads, 2 * kTile * sizeof(int), st>>>(
          kin, vin, nullptr, nullptr, n_cols, d_status, E, 8 * p, nblk, hscan,
          last ? nullptr : kout, vout, nullptr);
    }
    PYGB_LAUNCH_CHECK();
    vin = vout;
    cin = cout;
    const uint32_t* t = kin;
    kin = kout;
    kout = const_cast<uint32_t*>(t);
  }
  return PYGB_OK;
}

int pygb_permute_f32(const float* src, const int32_t* perm, int64_t E, float* out,
                     pygb_stream_t s) {
  PYGB_REQUIRE(E >= 0, "pygb_permute_f32: bad arguments");
  if (E == 0) return PYGB_OK;
  permute_f32_kernel<<<grid_for(E), kThreads, 0, as_stream(s)>>>(src, perm, E, out);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

int pygb_degree(const int64_t* index, int64_t E, int64_t N, int out_is_float, void* out,
                pygb_stream_t s) {
  PYGB_REQUIRE(E >= 0 && N >= 0 && (out || N == 0), "pygb_degree: bad arguments");
  cudaStream_t st = as_stream(s);
  if (N == 0) return PYGB_OK;
  PYGB_CUDA(cudaMemsetAsync(out, 0, (size_t)N * (out_is_float ? 4 : 8), st));
  if (E > 0) {
    degree_count_kernel<<<grid_for(E), kThreads, 0, st>>>(
        index, E, N, out_is_float ? nullptr : reinterpret_cast<unsigned long long*>(out),
        out_is_float ? reinterpret_cast<int*>(out) : nullptr);
    PYGB_LAUNCH_CHECK();
  }
  if (out_is_float) {
    i32_to_f32_kernel<<<grid_for(N), kThreads, 0, st>>>(reinterpret_cast<int*>(out), N);
    PYGB_LAUNCH_CHECK();
  }
  return PYGB_OK;
}

int pygb_weighted_degree(const float* w, const int32_t* rowptr, const int32_t* perm, int64_t N,
                         float* deg, pygb_stream_t s) {
  PYGB_REQUIRE(N >= 0 && w && rowptr && perm, "pygb_weighted_degree: bad arguments");
  if (N == 0) return PYGB_OK;
  weighted_degree_kernel<<<grid_for(N), kThreads, 0, as_stream(s)>>>(w, rowptr, perm, N, deg);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

int pygb_gcn_norm(const int64_t* edge_index, const float* w, int64_t E, int64_t N,
                  const float* deg, float* norm, pygb_stream_t s) {
  PYGB_REQUIRE(E >= 0 && N >= 0, "pygb_gcn_norm: bad arguments");
  if (E == 0) return PYGB_OK;
  gcn_norm_kernel<<<grid_for(E), kThreads, 0, as_stream(s)>>>(edge_index, edge_index + E, w, E,
                                                             deg, norm);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

int pygb_gcn_norm_csr(const int32_t* rowptr, const int32_t* col, const float* deg,
                      int64_t n_rows, float* out, pygb_stream_t s) {
  PYGB_REQUIRE(n_rows >= 0, "pygb_gcn_norm_csr: bad arguments");
  if (n_rows == 0) return PYGB_OK;
  PYGB_REQUIRE(rowptr && col && deg && out, "pygb_gcn_norm_csr: null pointer");
  int64_t blocks = ceil_div(n_rows, kWarps);  // one row per warp: latency hidden by occupancy
  if (blocks > (1 << 30)) blocks = 1 << 30;
  gcn_norm_csr_kernel<<<(unsigned)blocks, kThreads, 0, as_stream(s)>>>(rowptr, col, deg, n_rows,
                                                                      out);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

}  // extern "C"
