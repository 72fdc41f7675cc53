// K1 v2 — streaming segmented gather-reduce (aggr = add | mean).
//
// Same result contract as seg_gather_reduce_kernel (seg_reduce.cu), organised
// for memory-level parallelism and a short instruction stream per edge:
//
//  * Work decomposition.  The E slots of the destination-sorted CSR are cut into
//    sub-items of Ts consecutive slots (equal work whatever the degree skew).  A
//    warp is 32/LPF independent lane groups; each group streams one sub-item at a
//    time (LPF lanes x 16 B cover a feature row, VPL vectors per lane), so narrow
//    rows keep all 32 lanes busy without any cross-group reduction.  Persistent
//    warps take items w, w+W, ... (an item = EPW consecutive sub-items).
//
//  * Row boundaries cost nothing in the hot loop: the index stream carries a flag
//    in bit 31 ("last slot of its row", pygb_seg_flag_idx) and the row ids come
//    from nz_rows (the non-empty rows in order), one entry prefetched ahead.
//    sub_info[s] = rank of the row that owns slot s*Ts | bit 31 if that row began
//    in an earlier sub-item.  A row cut by a sub-item boundary is reduced
//    piecewise into fp32 `partial` rows (at most two per sub-item) which the
//    fix-up kernel adds in slot order: no float atomics, run-to-run
//    deterministic, and a 17k-edge hub row is spread over ~35 streams.
//
//  * Staging.  For each step (one slot per lane group) the warp issues cp.async
//    (LDGSTS, 16 B per lane, L2-only) copies of the gathered feature rows and a
//    4-byte copy of the edge weight into a private shared-memory ring of NSTAGE
//    stages x G steps, NSTAGE-1 stages ahead of the FMAs.  Every lane later reads
//    exactly the bytes it copied (cp.async.wait_group is the only data sync);
//    flags and weights cross lanes and are ordered by __syncwarp.  ~8 KB in
//    flight per warp x 16-20 warps per SM, without spending registers on it.
//
// HBM-bound; algorithmic bytes per edge = 4 (col) + 4 (weight) + F*sizeof(T),
// plus one output row per destination.
#include <float.h>
#include <math.h>

#include <mutex>
#include <type_traits>

#include "stream_common.cuh"

namespace pygb {
namespace {

struct StreamParams {
  const void* x;
  int64_t ldx;
  int F;
  const int* rowptr;
  const int* idxf;  // index stream with the row-end flag in bit 31
  const float* ew;
  const float* bias;
  int64_t n_rows;
  int E;
  int mean;
  void* out;
  int64_t ldo;
  const int* sub_info;
  const int* nz_rows;
  int Ts;
  int n_sub;
  float* partial;  // [2*n_sub][Fp] fp32
  int* tail_row;   // [n_sub], pre-set to -1
  int Fp;
  // routed output (feature-sliced multi-GPU aggregation): row r is stored at
  // out_tab[r / blk_rows] + out_off + (r % blk_rows) * ldo — the table holds one
  // (possibly peer-GPU) base pointer per block of blk_rows destination rows.
  void* const* out_tab;
  int blk_rows;
  unsigned blk_magic;  // floor(2^32 / blk_rows) + 1
  int sub_rot;         // first sub-item this launch starts at (work order rotation)
  int max_ctas;        // > 0: cap of the persistent grid; < 0: leave |max_ctas| SMs to a
                       // concurrent kernel (grid = (SMs + max_ctas) x resident CTAs per SM)
  int64_t out_off;  // bytes
  int hot;          // the index stream carries a "hot source row" hint in bit 30 (pygb_seg_mark_hot):
                    // those gathers ask L2 to keep the line (evict_last)
};

template <typename T>
__device__ __forceinline__ T* out_row(const StreamParams& P, int64_t row) {
  if (P.out_tab) {
    // row / blk_rows without a division: the estimate is q or q + 1 for row < 2^31
    int q = (int)__umulhi((unsigned)row, P.blk_magic);
    if (q * P.blk_rows > (int)row) --q;
    char* base = reinterpret_cast<char*>(
        __ldg(reinterpret_cast<const unsigned long long*>(P.out_tab) + q));
    return reinterpret_cast<T*>(base + P.out_off) + (int64_t)((int)row - q * P.blk_rows) * P.ldo;
  }
  return static_cast<T*>(P.out) + row * P.ldo;
}

// WEIGHTED: an edge weight stream exists (16-byte aligned).  EXACT: a feature row
// is exactly LPF*VPL 16-byte units, so no lane is idle and no unit test is needed.
template <typename T, int LPF, int VPL, bool WEIGHTED, bool EXACT>
__global__ void __launch_bounds__(kStreamWarps * 32)
seg_stream_kernel(const StreamParams P) {
  using C = StreamCfg<T, LPF, VPL>;
  constexpr int EPW = C::EPW, KV = C::KV, G = C::G, NSTAGE = C::NSTAGE;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int fl = lane & (LPF - 1), es = lane / LPF;
  // per-lane base inside a stage: data at +0, flags at flag_off, weights after them
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem_raw) +
                         (uint32_t)(warp * C::WARP_B) + (uint32_t)(lane * 16);
  const uint32_t meta = (uint32_t)(G * C::DATA_B + es * G * 4) - (uint32_t)(lane * 16);
  const int F = P.F;
  const int units = (F + KV - 1) / KV;
  bool act[VPL];
#pragma unroll
  for (int v = 0; v < VPL; ++v) act[v] = EXACT || (fl + v * LPF < units);
  const int W = gridDim.x * kStreamWarps;
  const char* xb = static_cast<const char*>(P.x) + fl * 16;
  asm volatile("" : "+l"(xb));  // keep the per-lane base in registers
  const uint32_t ldxb = (uint32_t)(P.ldx * sizeof(T));
  // L2 eviction priority of the gathered rows: rows of frequently gathered sources (bit 30 of the
  // index stream, when the caller marked them) are kept in preference to the rest
  const int hot_bit = P.hot ? 0x40000000 : 0;
  const int idx_mask = P.hot ? 0x3fffffff : 0x7fffffff;
  const uint64_t pol_hot = l2_policy_evict_last();
  const uint64_t pol_norm = P.hot == 2 ? l2_policy_evict_first() : l2_policy_evict_normal();
  const int* __restrict__ idxf = P.idxf;
  const float* __restrict__ ew = P.ew;
  const int* __restrict__ nz_rows = P.nz_rows;
  const int E = P.E, Ts = P.Ts;
  const int n_items = (P.n_sub + EPW - 1) / EPW;
  const int grp_lane0 = es * LPF;
  const bool lead = fl == 0;

  float acc[VPL][KV];
#pragma unroll
  for (int v = 0; v < VPL; ++v)
#pragma unroll
    for (int i = 0; i < KV; ++i) acc[v][i] = 0.f;

  // rotated start (routed output): ranks begin at different row blocks so that no peer
  // receives from everybody at once
  const int item_rot = (P.sub_rot / EPW) % (n_items > 0 ? n_items : 1);
  for (int it = blockIdx.x + gridDim.x * warp; it < n_items; it += W) {
    int item = it + item_rot;
    if (item >= n_items) item -= n_items;
    // ---- per lane-group stream state ------------------------------------------------
    const int s = item * EPW + es;
    const bool live = s < P.n_sub;
    const int pg0 = live ? s * Ts : 0;
    const int n_g = live ? min(Ts, E - pg0) : 0;
    const int info = live ? __ldg(P.sub_info + s) : 0;
    int rank = info & 0x7fffffff;
    bool left_open = info < 0;
    int row_cur = live ? __ldg(nz_rows + rank) : 0;
    int row_nxt = live ? __ldg(nz_rows + rank + 1) : 0;
    int kstart = 0;
    const int* __restrict__ my_idx = idxf + pg0 + fl;
    const float* __restrict__ my_w = ew + pg0;
    int idx_cur = (fl < n_g) ? __ldg(my_idx) : 0;
    int idx_nxt = (LPF + fl < n_g) ? __ldg(my_idx + LPF) : 0;
    // every lane group of the warp has all Ts slots (true except at the very end of the CSR)
    const bool all_full = (int64_t)(item * EPW + EPW) * Ts <= (int64_t)E;

    auto flush = [&](int k) {  // slot k closes the current row of this lane group
      if (!left_open) {
        const float inv = (float)(k + 1 - kstart);
        T* orow = out_row<T>(P, row_cur);
#pragma unroll
        for (int v = 0; v < VPL; ++v)
          if (act[v]) store_row<T, KV>(acc[v], (fl + v * LPF) * KV, F, inv, P.mean, P.bias, orow);
      } else {  // the row began in an earlier sub-item: head piece
        float* pr = P.partial + (int64_t)(2 * s) * P.Fp;
#pragma unroll
        for (int v = 0; v < VPL; ++v)
          if (act[v]) store_partial<KV>(acc[v], pr + (fl + v * LPF) * KV);
      }
#pragma unroll
      for (int v = 0; v < VPL; ++v)
#pragma unroll
        for (int i = 0; i < KV; ++i) acc[v][i] = 0.f;
      left_open = false;
      kstart = k + 1;
      ++rank;
      row_cur = row_nxt;
      row_nxt = __ldg(nz_rows + rank + 1);
    };

    auto run = [&](auto full_tag) {
      constexpr bool FULL = decltype(full_tag)::value;
      const int nsteps = FULL ? Ts : __shfl_sync(kFull, n_g, 0);  // group 0 has the most slots
      const int nstages = (nsteps + G - 1) / G;

      auto issue_stage = [&](int st) {
        const uint32_t sb = sbase + (uint32_t)((st & (NSTAGE - 1)) * C::STAGE_B);
        const int k0 = st * G;
        if (st > 0 && (k0 & (LPF - 1)) == 0) {  // next batch of LPF index entries
          idx_cur = idx_nxt;
          idx_nxt = (FULL ? k0 + LPF < Ts : k0 + LPF + fl < n_g) ? __ldg(my_idx + k0 + LPF) : 0;
        }
        int craw[G];
        const char* src[G];
        const int lane0 = grp_lane0 + (k0 & (LPF - 1));
#pragma unroll
        for (int g = 0; g < G; ++g) {
          craw[g] = __shfl_sync(kFull, idx_cur, lane0 + g);
          src[g] = xb + (uint64_t)(uint32_t)(craw[g] & idx_mask) * ldxb;
        }
#pragma unroll
        for (int g = 0; g < G; ++g) {
          if (FULL || k0 + g < n_g) {
#pragma unroll
            for (int v = 0; v < VPL; ++v)
              if (act[v])
                cp_async16_hint(sb + (uint32_t)(g * C::DATA_B + v * 512), src[g] + v * LPF * 16,
                                (craw[g] & hot_bit) ? pol_hot : pol_norm);
          }
        }
        if (lead) {
          sts_words<G>(sb + meta, craw);
          if (WEIGHTED) {
            if (FULL) {
              cp_async_small<G * 4>(sb + meta + C::META_B, my_w + k0);
            } else {
#pragma unroll
              for (int g = 0; g < G; ++g)
                if (k0 + g < n_g) cp_async_small<4>(sb + meta + C::META_B + g * 4, my_w + k0 + g);
            }
          }
        }
        cp_async_commit();
      };

      auto consume_stage = [&](int st) {
        const uint32_t sb = sbase + (uint32_t)((st & (NSTAGE - 1)) * C::STAGE_B);
        const int k0 = st * G;
        uint32_t flags[G], wbits[G];
        lds_words<G>(sb + meta, flags);
        if (WEIGHTED) lds_words<G>(sb + meta + C::META_B, wbits);
#pragma unroll
        for (int g = 0; g < G; ++g) {
          if (FULL || k0 + g < n_g) {
            const float w = WEIGHTED ? __uint_as_float(wbits[g]) : 1.f;
#pragma unroll
            for (int v = 0; v < VPL; ++v) {
              if (act[v]) {
                const uint4 qv = lds128(sb + (uint32_t)(g * C::DATA_B + v * 512));
                float f[KV];
                Elem<T>::unpack(qv, f);
#pragma unroll
                for (int i = 0; i < KV; ++i) acc[v][i] = fmaf(w, f[i], acc[v][i]);
              }
            }
            if ((int)flags[g] < 0) flush(k0 + g);
          }
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

    // ---- a row that continues in the next sub-item: head or tail piece ---------------
    if (kstart < n_g) {
      float* pr = P.partial + (int64_t)(2 * s + (left_open ? 0 : 1)) * P.Fp;
#pragma unroll
      for (int v = 0; v < VPL; ++v)
        if (act[v]) store_partial<KV>(acc[v], pr + (fl + v * LPF) * KV);
      if (lead && !left_open) P.tail_row[s] = row_cur;
#pragma unroll
      for (int v = 0; v < VPL; ++v)
#pragma unroll
        for (int i = 0; i < KV; ++i) acc[v][i] = 0.f;
    }
  }
  cp_async_wait<0>();
}

// Adds the pieces of rows that were cut by sub-item boundaries (slot order), and
// writes the rows that have no slot at all.
template <typename T>
__global__ void __launch_bounds__(256)
seg_stream_fixup_kernel(const StreamParams P) {
  constexpr int KV = Elem<T>::kVec;
  const int lane = threadIdx.x & 31;
  const int gw = (blockIdx.x * 256 + threadIdx.x) >> 5;
  const int nw = (gridDim.x * 256) >> 5;
  const int F = P.F, Fp = P.Fp;
  for (int c = gw; c < P.n_sub; c += nw) {
    const int row = P.tail_row[c];
    if (row < 0) continue;
    const int beg = P.rowptr[row], end = P.rowptr[row + 1];
    const int c_last = (end - 1) / P.Ts;
    const float inv = (float)max(end - beg, 1);
    for (int f0 = lane * KV; f0 < Fp; f0 += 32 * KV) {
      float acc[KV];
      const float* pt = P.partial + (int64_t)(2 * c + 1) * Fp + f0;
#pragma unroll
      for (int i = 0; i < KV; ++i) acc[i] = pt[i];
      for (int cc = c + 1; cc <= c_last; ++cc) {
        const float* ph = P.partial + (int64_t)(2 * cc) * Fp + f0;
#pragma unroll
        for (int i = 0; i < KV; ++i) acc[i] += ph[i];
      }
      store_row<T, KV>(acc, f0, F, inv, P.mean, P.bias, out_row<T>(P, row));
    }
  }
  // rows without any slot: sum = 0 (+ bias)
  for (int64_t base = (int64_t)gw * 32; base < P.n_rows; base += (int64_t)nw * 32) {
    const int64_t row = base + lane;
    const bool empty = row < P.n_rows && P.rowptr[row] == P.rowptr[row + 1];
    unsigned mask = __ballot_sync(kFull, empty);
    while (mask) {
      const int l = __ffs(mask) - 1;
      mask &= mask - 1;
      T* orow = out_row<T>(P, base + l);
      for (int f = lane; f < F; f += 32)
        Elem<T>::store(orow + f, P.bias ? __ldg(P.bias + f) : 0.f);
    }
  }
}

// ---- plan kernels ------------------------------------------------------------------
__global__ void __launch_bounds__(256)
row_nonempty_kernel(const int* __restrict__ rowptr, int64_t n_rows, int* __restrict__ flag) {
  for (int64_t r = (int64_t)blockIdx.x * 256 + threadIdx.x; r < n_rows;
       r += (int64_t)gridDim.x * 256)
    flag[r] = rowptr[r + 1] > rowptr[r];
}

__global__ void __launch_bounds__(256)
nz_rows_kernel(const int* __restrict__ rowptr, const int* __restrict__ rank, int64_t n_rows,
               int* __restrict__ nz_rows) {
  for (int64_t r = (int64_t)blockIdx.x * 256 + threadIdx.x; r < n_rows;
       r += (int64_t)gridDim.x * 256)
    if (rowptr[r + 1] > rowptr[r]) nz_rows[rank[r]] = (int)r;
}

__global__ void __launch_bounds__(256)
sub_info_kernel(const int* __restrict__ rowptr, const int* __restrict__ rank, int n_rows, int Ts,
                int n_sub, int* __restrict__ sub_info) {
  for (int c = blockIdx.x * 256 + threadIdx.x; c < n_sub; c += gridDim.x * 256) {
    const int target = c * Ts;  // the row r with rowptr[r] <= target < rowptr[r+1]
    int lo = 0, hi = n_rows;    // invariant: rowptr[lo] <= target < rowptr[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (__ldg(rowptr + mid) <= target) lo = mid;
      else hi = mid;
    }
    sub_info[c] = rank[lo] | (rowptr[lo] < target ? (int)0x80000000 : 0);
  }
}

__global__ void __launch_bounds__(256)
flag_idx_kernel(const int* __restrict__ rowptr, int64_t n_rows, int* __restrict__ idxf) {
  for (int64_t r = (int64_t)blockIdx.x * 256 + threadIdx.x; r < n_rows;
       r += (int64_t)gridDim.x * 256) {
    const int e = rowptr[r + 1];
    if (e > rowptr[r]) idxf[e - 1] |= (int)0x80000000;
  }
}

// bit 30 on every slot whose source row is gathered at least `threshold` times
__global__ void __launch_bounds__(256)
mark_hot_kernel(int* __restrict__ idxf, int64_t E, const int* __restrict__ freq, int threshold) {
  for (int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x; e < E; e += (int64_t)gridDim.x * 256) {
    const int v = idxf[e];
    if (__ldg(freq + (v & 0x3fffffff)) >= threshold) idxf[e] = v | 0x40000000;
  }
}

template <typename T, int LPF, int VPL, bool WEIGHTED, bool EXACT>
int launch_stream_k(const StreamParams& P, cudaStream_t st) {
  using C = StreamCfg<T, LPF, VPL>;
  auto kern = seg_stream_kernel<T, LPF, VPL, WEIGHTED, EXACT>;
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
      attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::CTA_B);
      if (attr_err == cudaSuccess)
        attr_err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, kern,
                                                                 kStreamWarps * 32, C::CTA_B);
      if (ctas_per_sm < 1) ctas_per_sm = 1;
      if (attr_err == cudaSuccess && dev >= 0 && dev < 64) ctas_cache[dev] = ctas_per_sm;
    }
  }
  PYGB_CUDA(attr_err);
  const int64_t n_items = ceil_div(P.n_sub, C::EPW);
  int64_t grid = (int64_t)sm_count() * ctas_per_sm;
  const int64_t need = ceil_div(n_items, kStreamWarps);
  if (grid > need) grid = need;
  if (P.max_ctas > 0 && grid > P.max_ctas) grid = P.max_ctas;
  if (P.max_ctas < 0) {
    int64_t sms_left = (int64_t)sm_count() + P.max_ctas;
    if (sms_left < 1) sms_left = 1;
    if (grid > sms_left * ctas_per_sm) grid = sms_left * ctas_per_sm;
  }
  if (grid < 1) grid = 1;
  kern<<<(unsigned)grid, kStreamWarps * 32, C::CTA_B, st>>>(P);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

template <typename T, int LPF, int VPL>
int launch_stream_one(const StreamParams& P, cudaStream_t st) {
  constexpr int KV = Elem<T>::kVec;
  const bool exact = (P.F % KV == 0) && (P.F / KV == LPF * VPL);
  if (P.ew != nullptr)
    return exact ? launch_stream_k<T, LPF, VPL, true, true>(P, st)
                 : launch_stream_k<T, LPF, VPL, true, false>(P, st);
  return exact ? launch_stream_k<T, LPF, VPL, false, true>(P, st)
               : launch_stream_k<T, LPF, VPL, false, false>(P, st);
}

// Lane-group shape for a row of `units` 16-byte vectors: LPF lanes x VPL vectors per
// lane.  Other shapes with the same capacity (e.g. 16x2 or 8x4 instead of 32x1) were
// measured within 2 % on the products graph (profiles/r01_k1_shapes.md), so only the
// widest-lane shape of every capacity is instantiated.
template <typename T>
int launch_stream(const StreamParams& P, cudaStream_t st) {
  constexpr int KV = Elem<T>::kVec;
  const int units = (P.F + KV - 1) / KV;
  int rc;
  if (units <= 4) rc = launch_stream_one<T, 4, 1>(P, st);
  else if (units <= 8) rc = launch_stream_one<T, 8, 1>(P, st);
  else if (units <= 16) rc = launch_stream_one<T, 16, 1>(P, st);
  else if (units <= 32) rc = launch_stream_one<T, 32, 1>(P, st);
  else if (units <= 64) rc = launch_stream_one<T, 32, 2>(P, st);
  else if (units <= 96) rc = launch_stream_one<T, 32, 3>(P, st);
  else rc = launch_stream_one<T, 32, 4>(P, st);
  if (rc) return rc;
  int64_t work = (int64_t)P.n_sub * 32;
  if (work < P.n_rows) work = P.n_rows;
  seg_stream_fixup_kernel<T><<<grid1d(work), 256, 0, st>>>(P);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

}  // namespace

int exclusive_scan(const int* in, int* out, int64_t n, int* ws_tiles, cudaStream_t st);
int64_t exclusive_scan_tiles(int64_t n);

}  // namespace pygb

using namespace pygb;

extern "C" {

int pygb_seg_plan(int64_t E, int32_t* Ts_out, int64_t* n_sub_out) {
  PYGB_REQUIRE(E >= 0 && E < (1ll << 31) && Ts_out && n_sub_out, "pygb_seg_plan: bad arguments");
  // >= ~16 sub-items per resident lane-group stream for balance; Ts a power of two in [32, 512]
  const int64_t streams = (int64_t)sm_count() * 20;
  int Ts = 512;
  while (Ts > 32 && E / Ts < streams * 16) Ts >>= 1;
  *Ts_out = Ts;
  *n_sub_out = ceil_div(E, Ts);
  return PYGB_OK;
}

size_t pygb_seg_tables_workspace(int64_t n_rows) {
  return 2 * align_up(((size_t)n_rows + 2) * sizeof(int), 256) +
         align_up(((size_t)exclusive_scan_tiles(n_rows + 1) + 2) * sizeof(int), 256) + 512;
}

int pygb_seg_tables(const int32_t* rowptr, int64_t n_rows, int64_t E, int32_t Ts, int64_t n_sub,
                    int32_t* nz_rows, int32_t* sub_info, void* ws, size_t ws_bytes,
                    pygb_stream_t s) {
  PYGB_REQUIRE(n_rows >= 0 && E >= 0 && Ts > 0 && n_sub == ceil_div(E, Ts) &&
                   n_rows < (1ll << 31) - 64 && E < (1ll << 31) - 1024,
               "pygb_seg_tables: bad arguments");
  if (n_rows == 0) return PYGB_OK;
  PYGB_REQUIRE(rowptr && nz_rows && (n_sub == 0 || sub_info), "pygb_seg_tables: null pointer");
  if (ws_bytes < pygb_seg_tables_workspace(n_rows)) {
    set_error("pygb_seg_tables: workspace too small");
    return PYGB_E_WORKSPACE;
  }
  cudaStream_t st = as_stream(s);
  Workspace Wk(ws, ws_bytes);
  int* flag = Wk.take<int>((size_t)n_rows + 2);
  int* rank = Wk.take<int>((size_t)n_rows + 2);
  int* tiles = Wk.take<int>((size_t)exclusive_scan_tiles(n_rows + 1) + 2);
  row_nonempty_kernel<<<grid1d(n_rows), 256, 0, st>>>(rowptr, n_rows, flag);
  PYGB_LAUNCH_CHECK();
  int rc = exclusive_scan(flag, rank, n_rows, tiles, st);
  if (rc) return rc;
  // nz_rows has n_rows + 2 entries; the tail past the last non-empty row stays 0
  PYGB_CUDA(cudaMemsetAsync(nz_rows, 0, ((size_t)n_rows + 2) * sizeof(int), st));
  nz_rows_kernel<<<grid1d(n_rows), 256, 0, st>>>(rowptr, rank, n_rows, nz_rows);
  PYGB_LAUNCH_CHECK();
  if (n_sub > 0) {
    sub_info_kernel<<<grid1d(n_sub), 256, 0, st>>>(rowptr, rank, (int)n_rows, Ts, (int)n_sub,
                                                  sub_info);
    PYGB_LAUNCH_CHECK();
  }
  return PYGB_OK;
}

int pygb_seg_flag_idx(const int32_t* rowptr, int64_t n_rows, const int32_t* idx, int64_t E,
                      int32_t* idx_flagged, pygb_stream_t s) {
  PYGB_REQUIRE(n_rows >= 0 && E >= 0, "pygb_seg_flag_idx: bad arguments");
  if (E == 0 || n_rows == 0) return PYGB_OK;
  PYGB_REQUIRE(rowptr && idx && idx_flagged, "pygb_seg_flag_idx: null pointer");
  cudaStream_t st = as_stream(s);
  if (idx != idx_flagged)
    PYGB_CUDA(cudaMemcpyAsync(idx_flagged, idx, (size_t)E * sizeof(int), cudaMemcpyDeviceToDevice,
                              st));
  flag_idx_kernel<<<grid1d(n_rows), 256, 0, st>>>(rowptr, n_rows, idx_flagged);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

static int64_t padded_f(int dtype, int64_t F) {
  const int kv = dtype == PYGB_F32 ? 4 : 8;
  return (F + kv - 1) / kv * kv;
}

size_t pygb_seg_stream_workspace(int dtype, int64_t n_sub, int64_t F) {
  const int64_t fblk = 128 * (dtype == PYGB_F32 ? 4 : 8);
  if (F > fblk) F = fblk;
  return align_up((size_t)(2 * n_sub) * padded_f(dtype, F) * sizeof(float), 256) +
         align_up((size_t)n_sub * sizeof(int), 256) + 256;
}

static int seg_stream_impl(int reduce, int dtype, const void* x, int64_t ldx, int64_t F,
                           const int32_t* rowptr, const int32_t* idx_flagged, const float* edge_w,
                           const float* bias, int64_t n_rows, int64_t E, const int32_t* nz_rows,
                           const int32_t* sub_info, int32_t Ts, int64_t n_sub, void* out,
                           void* const* out_tab, int64_t blk_rows, int64_t sub_rot, int64_t ldo,
                           int64_t max_ctas, void* ws, size_t ws_bytes, pygb_stream_t s,
                           int hot = 0) {
  PYGB_REQUIRE(reduce == PYGB_SUM || reduce == PYGB_MEAN,
               "pygb_seg_gather_reduce_stream: bad reduce %d", reduce);
  PYGB_REQUIRE(dtype == PYGB_F32 || dtype == PYGB_BF16,
               "pygb_seg_gather_reduce_stream: bad dtype %d", dtype);
  PYGB_REQUIRE(F >= 0 && n_rows >= 0 && E >= 0 && E < (1ll << 31) - 1024 &&
                   n_rows < (1ll << 31) - 64 && ldo >= F && F < (1ll << 24),
               "pygb_seg_gather_reduce_stream: bad sizes");
  PYGB_REQUIRE(Ts >= 32 && (Ts & (Ts - 1)) == 0 && n_sub == ceil_div(E, Ts),
               "pygb_seg_gather_reduce_stream: bad plan (Ts must be a power of two >= 32)");
  if (n_rows == 0 || F == 0) return PYGB_OK;
  PYGB_REQUIRE(rowptr && (out || out_tab) && (E == 0 || (x && idx_flagged && nz_rows && sub_info)),
               "pygb_seg_gather_reduce_stream: null pointer");
  PYGB_REQUIRE(sub_rot >= 0 && sub_rot <= n_sub, "pygb_seg_gather_reduce_stream_routed: bad sub_rot");
  PYGB_REQUIRE(!out_tab || (blk_rows > 1 && blk_rows < (1ll << 31)),
               "pygb_seg_gather_reduce_stream_routed: bad blk_rows");
  const int esz = dtype == PYGB_F32 ? 4 : 2;
  const int kv = 16 / esz;
  PYGB_REQUIRE(aligned16(x) && (!out || aligned16(out)) && ldx % kv == 0 && ldo % kv == 0 &&
                   ldx >= padded_f(dtype, F) && ldo >= padded_f(dtype, F) &&
                   ldx * esz < (1ll << 31) && aligned16(edge_w),
               "pygb_seg_gather_reduce_stream: x/out rows (and edge_w) must be 16-byte aligned, "
               "rows padded to 16 bytes");
  if (ws_bytes < pygb_seg_stream_workspace(dtype, n_sub, F)) {
    set_error("pygb_seg_gather_reduce_stream: workspace too small");
    return PYGB_E_WORKSPACE;
  }
  cudaStream_t st = as_stream(s);
  // feature blocks of at most 128 16-byte units (2 KB) per row
  const int64_t fblk = 128 * kv;
  const int64_t Fp_max = padded_f(dtype, F < fblk ? F : fblk);
  Workspace Wk(ws, ws_bytes);
  float* partial = Wk.take<float>((size_t)(2 * n_sub) * Fp_max);
  int* tail_row = Wk.take<int>((size_t)n_sub);
  for (int64_t f0 = 0; f0 < F; f0 += fblk) {
    const int64_t fb = F - f0 < fblk ? F - f0 : fblk;
    StreamParams P;
    P.x = static_cast<const char*>(x) + f0 * esz;
    P.ldx = ldx;
    P.F = (int)fb;
    P.rowptr = rowptr;
    P.idxf = idx_flagged;
    P.ew = edge_w;
    P.bias = bias ? bias + f0 : nullptr;
    P.n_rows = n_rows;
    P.E = (int)E;
    P.mean = reduce == PYGB_MEAN;
    P.out = out ? static_cast<char*>(out) + f0 * esz : nullptr;
    P.out_tab = out_tab;
    P.blk_rows = (int)blk_rows;
    P.blk_magic = out_tab ? (unsigned)((1ull << 32) / (uint64_t)blk_rows + 1) : 0u;
    P.sub_rot = (int)sub_rot;
    P.max_ctas = (int)(max_ctas > -(1 << 20) && max_ctas < (1 << 30) ? max_ctas : 0);
    P.out_off = f0 * esz;
    P.hot = hot;
    if (hot) {
      const char* e = getenv("PYGB_K1_COLD_FIRST");   // probe: unmarked gathers evict_first
      if (e && atoi(e) != 0) P.hot = 2;
    }
    P.ldo = ldo;
    P.sub_info = sub_info;
    P.nz_rows = nz_rows;
    P.Ts = Ts;
    P.n_sub = (int)n_sub;
    P.partial = partial;
    P.tail_row = tail_row;
    P.Fp = (int)padded_f(dtype, fb);
    if (n_sub > 0) PYGB_CUDA(cudaMemsetAsync(tail_row, 0xff, (size_t)n_sub * sizeof(int), st));
    int rc = dtype == PYGB_F32 ? launch_stream<float>(P, st) : launch_stream<__nv_bfloat16>(P, st);
    if (rc) return rc;
  }
  return PYGB_OK;
}

int pygb_seg_gather_reduce_stream(int reduce, int dtype, const void* x, int64_t ldx, int64_t F,
                                  const int32_t* rowptr, const int32_t* idx_flagged,
                                  const float* edge_w, const float* bias, int64_t n_rows,
                                  int64_t E, const int32_t* nz_rows, const int32_t* sub_info,
                                  int32_t Ts, int64_t n_sub, void* out, int64_t ldo, void* ws,
                                  size_t ws_bytes, pygb_stream_t s) {
  PYGB_REQUIRE(out || n_rows == 0 || F == 0, "pygb_seg_gather_reduce_stream: null pointer");
  return seg_stream_impl(reduce, dtype, x, ldx, F, rowptr, idx_flagged, edge_w, bias, n_rows, E,
                         nz_rows, sub_info, Ts, n_sub, out, nullptr, 0, 0, ldo, 0, ws, ws_bytes, s);
}

int pygb_seg_gather_reduce_stream_routed(int reduce, int dtype, const void* x, int64_t ldx,
                                         int64_t F, const int32_t* rowptr,
                                         const int32_t* idx_flagged, const float* edge_w,
                                         const float* bias, int64_t n_rows, int64_t E,
                                         const int32_t* nz_rows, const int32_t* sub_info,
                                         int32_t Ts, int64_t n_sub, void* const* out_tab,
                                         int64_t blk_rows, int64_t sub_rot, int64_t ldo,
                                         int64_t max_ctas, void* ws, size_t ws_bytes,
                                         pygb_stream_t s) {
  PYGB_REQUIRE(out_tab || n_rows == 0 || F == 0,
               "pygb_seg_gather_reduce_stream_routed: null pointer table");
  return seg_stream_impl(reduce, dtype, x, ldx, F, rowptr, idx_flagged, edge_w, bias, n_rows, E,
                         nz_rows, sub_info, Ts, n_sub, nullptr, out_tab, blk_rows, sub_rot, ldo,
                         max_ctas, ws, ws_bytes, s);
}

// pygb_seg_gather_reduce_stream with an index stream that carries pygb_seg_mark_hot's hints
int pygb_seg_gather_reduce_stream_hinted(int reduce, int dtype, const void* x, int64_t ldx,
                                         int64_t F, const int32_t* rowptr,
                                         const int32_t* idx_flagged, const float* edge_w,
                                         const float* bias, int64_t n_rows, int64_t E,
                                         const int32_t* nz_rows, const int32_t* sub_info,
                                         int32_t Ts, int64_t n_sub, void* out, int64_t ldo,
                                         void* ws, size_t ws_bytes, pygb_stream_t s) {
  PYGB_REQUIRE(out || n_rows == 0 || F == 0, "pygb_seg_gather_reduce_stream_hinted: null pointer");
  return seg_stream_impl(reduce, dtype, x, ldx, F, rowptr, idx_flagged, edge_w, bias, n_rows, E,
                         nz_rows, sub_info, Ts, n_sub, out, nullptr, 0, 0, ldo, 0, ws, ws_bytes, s,
                         1);
}

// pygb_seg_gather_reduce_stream_routed with an index stream that carries the hints
int pygb_seg_gather_reduce_stream_routed_hinted(int reduce, int dtype, const void* x, int64_t ldx,
                                                int64_t F, const int32_t* rowptr,
                                                const int32_t* idx_flagged, const float* edge_w,
                                                const float* bias, int64_t n_rows, int64_t E,
                                                const int32_t* nz_rows, const int32_t* sub_info,
                                                int32_t Ts, int64_t n_sub, void* const* out_tab,
                                                int64_t blk_rows, int64_t sub_rot, int64_t ldo,
                                                int64_t max_ctas, void* ws, size_t ws_bytes,
                                                pygb_stream_t s) {
  PYGB_REQUIRE(out_tab || n_rows == 0 || F == 0,
               "pygb_seg_gather_reduce_stream_routed_hinted: null pointer table");
  return seg_stream_impl(reduce, dtype, x, ldx, F, rowptr, idx_flagged, edge_w, bias, n_rows, E,
                         nz_rows, sub_info, Ts, n_sub, nullptr, out_tab, blk_rows, sub_rot, ldo,
                         max_ctas, ws, ws_bytes, s, 1);
}

int pygb_seg_mark_hot(int32_t* idx_flagged, int64_t E, const int32_t* freq, int64_t n_src,
                      int32_t threshold, pygb_stream_t s) {
  PYGB_REQUIRE(E >= 0 && E < (1ll << 31) && n_src >= 0 && n_src <= (1ll << 30),
               "pygb_seg_mark_hot: bad sizes (source ids must fit 30 bits)");
  if (E == 0) return PYGB_OK;
  PYGB_REQUIRE(idx_flagged && freq, "pygb_seg_mark_hot: null pointer");
  mark_hot_kernel<<<grid1d(E), 256, 0, as_stream(s)>>>(idx_flagged, E, freq, threshold);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

}  // extern "C"

// ---- no-allocation variant ---------------------------------------------------------------
// The partial rows and tail table of the fix-up pass live in a file-scope device array instead
// of a caller-provided workspace, for callers that cannot allocate (CUDA graph capture, serving
// loops).  Every GPU that loads the library holds the array until exit.  In host code the
// array's NAME is only a stand-in: the device address comes from cudaGetSymbolAddress, looked
// up once per device with that device current.
static __device__ __align__(256) unsigned char g_seg_static_ws[PYGB_SEG_STREAM_STATIC_WS_BYTES];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_ws_mu;
void* g_static_ws_ptr[kMaxDevices] = {};

int static_ws_for_current_device(void** ptr) {
  int dev = -1;
  PYGB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("pygb_seg_gather_reduce_stream_static: device ordinal %d out of range", dev);
    return PYGB_E_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(g_static_ws_mu);
  if (!g_static_ws_ptr[dev]) {
    void* p = nullptr;
    PYGB_CUDA(cudaGetSymbolAddress(&p, g_seg_static_ws));
    if (!p || (reinterpret_cast<uintptr_t>(p) & 255) != 0) {
      set_error("pygb_seg_gather_reduce_stream_static: bad symbol address %p", p);
      return PYGB_E_CUDA;
    }
    g_static_ws_ptr[dev] = p;
  }
  *ptr = g_static_ws_ptr[dev];
  return PYGB_OK;
}
}  // namespace

extern "C" int pygb_seg_gather_reduce_stream_static(
    int reduce, int dtype, const void* x, int64_t ldx, int64_t F, const int32_t* rowptr,
    const int32_t* idx_flagged, const float* edge_w, const float* bias, int64_t n_rows, int64_t E,
    const int32_t* nz_rows, const int32_t* sub_info, int32_t Ts, int64_t n_sub, void* out,
    int64_t ldo, pygb_stream_t s) {
  PYGB_REQUIRE(out || n_rows == 0 || F == 0, "pygb_seg_gather_reduce_stream_static: null pointer");
  PYGB_REQUIRE((dtype == PYGB_F32 || dtype == PYGB_BF16) && F >= 0 && n_sub >= 0 &&
                   n_sub < (1ll << 31),
               "pygb_seg_gather_reduce_stream_static: bad dtype / sizes");
  const size_t need = pygb_seg_stream_workspace(dtype, n_sub, F);
  if (need > (size_t)PYGB_SEG_STREAM_STATIC_WS_BYTES) {
    set_error("pygb_seg_gather_reduce_stream_static: needs %zu bytes of scratch for n_sub=%lld, "
              "F=%lld but the built-in array holds %zu; use pygb_seg_gather_reduce_stream with a "
              "caller-provided workspace (or a plan with a larger Ts)",
              need, (long long)n_sub, (long long)F, (size_t)PYGB_SEG_STREAM_STATIC_WS_BYTES);
    return PYGB_E_WORKSPACE;
  }
  void* ws = nullptr;
  int rc = static_ws_for_current_device(&ws);
  if (rc) return rc;
  // seg_stream_impl resets the tail table (0xff) on the stream before every launch and the
  // fix-up pass reads only the partial rows that launch wrote: nothing else needs clearing
  return seg_stream_impl(reduce, dtype, x, ldx, F, rowptr, idx_flagged, edge_w, bias, n_rows, E,
                         nz_rows, sub_info, Ts, n_sub, out, nullptr, 0, 0, ldo, 0, ws,
                         (size_t)PYGB_SEG_STREAM_STATIC_WS_BYTES, s);
}
