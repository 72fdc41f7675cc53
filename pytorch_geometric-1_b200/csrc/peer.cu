// Peer-memory plumbing and the column scatter of the feature-sliced multi-GPU
// layout.
//
// One process per GPU.  Every rank allocates one arena with pygb_peer_alloc,
// exports its CUDA IPC handle, and opens the handles of the other ranks of the
// box; kernels then store straight into peer HBM over NVLink/NVSwitch:
//   * the aggregation kernel writes each finished row slice into the arena of
//     the rank that owns the row (pygb_seg_gather_reduce_stream_routed), and
//   * cols_scatter below redistributes a rank's rows [n, P*Fc] column block by
//     column block to the P ranks (row layout -> column layout).
// There is no spinning on flags: ranks synchronise with a stream-ordered NCCL
// collective issued by the host layer after these kernels.
#include <string.h>

#include "common.cuh"
#include "stream_common.cuh"

namespace pygb {
namespace {

// src [n_rows, lds] -> for q < n_parts, k < n_dup:
//   dst_tab[q*n_dup + k][(row0 + r) * ldd + c] = act(src[r, q*Fc + c])
// One thread per 16-byte unit; a group of Fc/KV consecutive threads writes one
// contiguous Fc-wide row slice of one destination rank.
template <typename T, bool RELU>
__global__ void __launch_bounds__(256)
cols_scatter_kernel(const T* __restrict__ src, int64_t lds, int64_t n_rows, int upr /*units/row*/,
                    int ups /*units per slice*/, int n_dup, void* const* __restrict__ dst_tab,
                    int64_t row0, int64_t ldd) {
  constexpr int KV = Elem<T>::kVec;
  const int64_t total = n_rows * upr;
  for (int64_t u = (int64_t)blockIdx.x * 256 + threadIdx.x; u < total;
       u += (int64_t)gridDim.x * 256) {
    const int64_t r = u / upr;
    const int cu = (int)(u - r * upr);
    const int q = cu / ups, c = cu - q * ups;
    uint4 v = ldg_nc_v4(src + r * lds + (int64_t)cu * KV);
    if (RELU) {
      float f[KV];
      Elem<T>::unpack(v, f);
#pragma unroll
      for (int i = 0; i < KV; ++i) f[i] = f[i] < 0.f ? 0.f : f[i];  // NaN stays NaN (torch.relu)
      v = Elem<T>::pack(f);
    }
    for (int k = 0; k < n_dup; ++k) {
      T* d = reinterpret_cast<T*>(
                 __ldg(reinterpret_cast<const unsigned long long*>(dst_tab) + q * n_dup + k)) +
             (row0 + r) * ldd + (int64_t)c * KV;
      *reinterpret_cast<uint4*>(d) = v;
    }
  }
}

// One-block barrier over peer memory.  Rank r holds a block of `world` 64-bit flags; thread q
// of rank r stores `epoch` (release, system scope) into flag r of rank q, then waits until flag
// q of its own block reaches `epoch` (acquire, system scope).  Epochs only grow, so a flag
// never has to be reset.  The kernels launched before the barrier on this stream have
// completed, hence their stores into peer memory precede the flag store; the kernels launched
// after it start only once every rank has arrived.  A rank that waits longer than timeout_ns
// traps (the process fails loudly instead of hanging the GPU).
__global__ void peer_barrier_kernel(unsigned long long* const* __restrict__ flag_tab, int rank,
                                    int world, unsigned long long epoch,
                                    unsigned long long timeout_ns) {
  const int q = threadIdx.x;
  if (q < world) {
    __threadfence_system();
    unsigned long long* remote = flag_tab[q] + rank;
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(remote), "l"(epoch) : "memory");
    const unsigned long long* mine = flag_tab[rank] + q;
    unsigned long long t0, t1, v;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
    for (;;) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(mine) : "memory");
      if (v >= epoch) break;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
      if (t1 - t0 > timeout_ns) {
        printf("pygb_peer_barrier: rank %d waited %llu ns for rank %d (epoch %llu, saw %llu)\n",
               rank, timeout_ns, q, epoch, v);
        __trap();
      }
      __nanosleep(64);
    }
  }
  __syncthreads();
}

}  // namespace
}  // namespace pygb

using namespace pygb;

extern "C" {

int pygb_peer_alloc(size_t bytes, void** ptr) {
  PYGB_REQUIRE(ptr && bytes > 0, "pygb_peer_alloc: bad arguments");
  PYGB_CUDA(cudaMalloc(ptr, bytes));
  PYGB_CUDA(cudaMemset(*ptr, 0, bytes));
  return PYGB_OK;
}

int pygb_peer_free(void* ptr) {
  if (ptr) PYGB_CUDA(cudaFree(ptr));
  return PYGB_OK;
}

int pygb_peer_export(const void* ptr, void* handle, size_t handle_bytes) {
  PYGB_REQUIRE(ptr && handle && handle_bytes >= sizeof(cudaIpcMemHandle_t),
               "pygb_peer_export: handle buffer must hold %zu bytes", sizeof(cudaIpcMemHandle_t));
  cudaIpcMemHandle_t h;
  PYGB_CUDA(cudaIpcGetMemHandle(&h, const_cast<void*>(ptr)));
  memcpy(handle, &h, sizeof(h));
  return PYGB_OK;
}

int pygb_peer_open(const void* handle, size_t handle_bytes, void** ptr) {
  PYGB_REQUIRE(ptr && handle && handle_bytes >= sizeof(cudaIpcMemHandle_t),
               "pygb_peer_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  PYGB_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return PYGB_OK;
}

int pygb_peer_close(void* ptr) {
  if (ptr) PYGB_CUDA(cudaIpcCloseMemHandle(ptr));
  return PYGB_OK;
}

int pygb_peer_barrier(void* const* flag_tab, int32_t rank, int32_t world, int64_t epoch,
                      int64_t timeout_ns, pygb_stream_t s) {
  PYGB_REQUIRE(flag_tab && world >= 1 && world <= 64 && rank >= 0 && rank < world && epoch > 0 &&
                   timeout_ns > 0,
               "pygb_peer_barrier: bad arguments (world <= 64, epoch > 0)");
  peer_barrier_kernel<<<1, 64, 0, as_stream(s)>>>(
      reinterpret_cast<unsigned long long* const*>(flag_tab), rank, world,
      (unsigned long long)epoch, (unsigned long long)timeout_ns);
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

int pygb_cols_scatter(int dtype, const void* src, int64_t lds, int64_t n_rows, int64_t Fc,
                      int32_t n_parts, int32_t n_dup, void* const* dst_tab, int64_t dst_row0,
                      int64_t ldd, int relu, pygb_stream_t s) {
  PYGB_REQUIRE(dtype == PYGB_F32 || dtype == PYGB_BF16, "pygb_cols_scatter: bad dtype %d", dtype);
  const int esz = dtype == PYGB_F32 ? 4 : 2, kv = 16 / esz;
  PYGB_REQUIRE(n_rows >= 0 && Fc > 0 && n_parts > 0 && n_dup > 0 && dst_row0 >= 0 && Fc % kv == 0 &&
                   lds % kv == 0 && ldd % kv == 0 && lds >= Fc * n_parts && ldd >= Fc &&
                   Fc * n_parts < (1ll << 24),
               "pygb_cols_scatter: widths must be multiples of 16 bytes (Fc=%lld, lds=%lld, "
               "ldd=%lld)", (long long)Fc, (long long)lds, (long long)ldd);
  if (n_rows == 0) return PYGB_OK;
  PYGB_REQUIRE(src && dst_tab && aligned16(src), "pygb_cols_scatter: null or unaligned pointer");
  const int ups = (int)(Fc / kv), upr = ups * n_parts;
  cudaStream_t st = as_stream(s);
  const unsigned grid = grid1d(n_rows * upr);
#define PYGB_SCATTER(TT)                                                                       \
  do {                                                                                         \
    if (relu)                                                                                  \
      cols_scatter_kernel<TT, true><<<grid, 256, 0, st>>>(static_cast<const TT*>(src), lds,    \
                                                          n_rows, upr, ups, n_dup, dst_tab,    \
                                                          dst_row0, ldd);                      \
    else                                                                                       \
      cols_scatter_kernel<TT, false><<<grid, 256, 0, st>>>(static_cast<const TT*>(src), lds,   \
                                                           n_rows, upr, ups, n_dup, dst_tab,   \
                                                           dst_row0, ldd);                     \
  } while (0)
  if (dtype == PYGB_F32) PYGB_SCATTER(float);
  else PYGB_SCATTER(__nv_bfloat16);
#undef PYGB_SCATTER
  PYGB_LAUNCH_CHECK();
  return PYGB_OK;
}

}  // extern "C"
