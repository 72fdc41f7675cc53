// Shared pieces of the streaming kernels (seg_stream.cu, gat_stream.cu): cp.async /
// shared-memory PTX wrappers, the ring geometry and the row epilogues.
#pragma once
#include "common.cuh"

namespace pygb {
namespace {

constexpr int kStreamWarps = 4;  // warps per CTA; every warp is an independent stream

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
// the same copy with an L2 eviction-priority hint (createpolicy: evict_last keeps a line in L2
// in preference to lines filled with evict_normal)
__device__ __forceinline__ void cp_async16_hint(uint32_t dst, const void* src, uint64_t policy) {
  asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(dst), "l"(src),
               "l"(policy)
               : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_normal() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
  asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t a) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"(a));
  return r;
}
__device__ __forceinline__ void sts32(uint32_t a, int v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
template <int G>
__device__ __forceinline__ void lds_words(uint32_t a, uint32_t* o) {
  if (G == 4) {
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(o[0]), "=r"(o[1]), "=r"(o[2]), "=r"(o[3])
                 : "r"(a));
  } else if (G == 2) {
    asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(o[0]), "=r"(o[1]) : "r"(a));
  } else {
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(o[0]) : "r"(a));
  }
}

template <typename T, int LPF, int VPL>
struct StreamCfg {
  static constexpr int EPW = 32 / LPF;
  static constexpr int KV = Elem<T>::kVec;
  static constexpr int NSTAGE = 4;                             // power of two
  static constexpr int G = VPL >= 4 ? 1 : (VPL >= 2 ? 2 : 4);  // steps per stage
  static constexpr int DATA_B = VPL * 512;                     // one step
  static constexpr int META_B = (EPW * G * 4 + 15) / 16 * 16;  // flags, then weights
  static constexpr int STAGE_B = G * DATA_B + 2 * META_B;
  static constexpr int WARP_B = NSTAGE * STAGE_B;
  static constexpr int CTA_B = WARP_B * kStreamWarps;
};

// Final value of a finished row: mean, bias, conversion, store.
template <typename T, int KV>
__device__ __forceinline__ void store_row(float* acc, int f0, int F, float cnt, int mean,
                                          const float* __restrict__ bias, T* orow) {
  if (mean) {  // scatter_mean divides (torch-scatter 2.0.4: out.div_(count)), so do we
#pragma unroll
    for (int i = 0; i < KV; ++i) acc[i] = __fdiv_rn(acc[i], cnt);
  }
  if (bias) {
#pragma unroll
    for (int i = 0; i < KV; ++i)
      if (f0 + i < F) acc[i] += __ldg(bias + f0 + i);
  }
  if (f0 + KV <= F) {
    *reinterpret_cast<uint4*>(orow + f0) = Elem<T>::pack(acc);
  } else {
#pragma unroll
    for (int i = 0; i < KV; ++i)
      if (f0 + i < F) Elem<T>::store(orow + f0 + i, acc[i]);
  }
}

template <int KV>
__device__ __forceinline__ void store_partial(const float* acc, float* dst) {
#pragma unroll
  for (int i = 0; i < KV; i += 4)
    *reinterpret_cast<float4*>(dst + i) = make_float4(acc[i], acc[i + 1], acc[i + 2], acc[i + 3]);
}

template <int G>
__device__ __forceinline__ void sts_words(uint32_t a, const int* v) {
  if (G == 4) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(v[0]), "r"(v[1]),
                 "r"(v[2]), "r"(v[3])
                 : "memory");
  } else if (G == 2) {
    asm volatile("st.shared.v2.u32 [%0], {%1,%2};" ::"r"(a), "r"(v[0]), "r"(v[1]) : "memory");
  } else {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v[0]) : "memory");
  }
}
template <int BYTES>
__device__ __forceinline__ void cp_async_small(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(dst), "l"(src), "n"(BYTES)
               : "memory");
}


inline int sm_count() {
  static int sms[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (sms[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    sms[dev] = n;
  }
  return sms[dev];
}

inline unsigned grid1d(int64_t n) {
  int64_t b = ceil_div(n, 256);
  const int64_t cap = (int64_t)sm_count() * 8;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (unsigned)b;
}

}  // namespace
}  // namespace pygb
