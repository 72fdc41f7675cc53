// Probe: achievable HBM bandwidth of row gathers (no reduction logic), to find the
// hardware ceiling for K1's access pattern.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include <algorithm>
#include <random>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ uint4 ldg_nc(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

// Each warp reads rows idx[k] for k in its contiguous chunk; ROWS rows in flight; row = VPL*512 bytes (or LPF lanes*16)
template <int LPF, int VPL, int U>
__global__ void __launch_bounds__(256) gather_kernel(const char* __restrict__ x, int64_t row_bytes, const int* __restrict__ idx, int64_t n_idx, int* sink) {
  constexpr int EPW = 32 / LPF;
  const int lane = threadIdx.x & 31, fl = lane % LPF, es = lane / LPF;
  const int64_t gw = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 5;
  const int64_t nw = ((int64_t)gridDim.x * 256) >> 5;
  const int64_t per = (n_idx + nw - 1) / nw;
  const int64_t beg = gw * per, end = min(beg + per, n_idx);
  uint32_t acc = 0;
  for (int64_t base = beg; base < end; base += 32) {
    const int n = (int)min((int64_t)32, end - base);
    const int my = lane < n ? idx[base + lane] : 0;
    for (int j = 0; j < n; j += EPW * U) {
      uint4 v[U][VPL];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int k = j + u * EPW + es;
        const int c = __shfl_sync(0xffffffffu, my, k & 31);
#pragma unroll
        for (int q = 0; q < VPL; ++q) {
          if (k < n) v[u][q] = ldg_nc(x + (int64_t)c * row_bytes + (q * LPF + fl) * 16);
          else v[u][q] = make_uint4(0, 0, 0, 0);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int q = 0; q < VPL; ++q) acc ^= v[u][q].x ^ v[u][q].y ^ v[u][q].z ^ v[u][q].w;
    }
  }
  if (acc == 0x12345678u) *sink = 1;
}

template <int LPF, int VPL, int U>
float run(const char* x, int64_t row_bytes, const int* idx, int64_t n_idx, int* sink, int grid) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 2; ++i) gather_kernel<LPF, VPL, U><<<grid, 256>>>(x, row_bytes, idx, n_idx, sink);
  cudaEventRecord(e0);
  for (int i = 0; i < 5; ++i) gather_kernel<LPF, VPL, U><<<grid, 256>>>(x, row_bytes, idx, n_idx, sink);
  cudaEventRecord(e1);
  CK(cudaEventSynchronize(e1));
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  return ms / 5;
}

int main(int argc, char** argv) {
  const int64_t n_rows = 2449029;
  const int64_t n_idx = 64449029;
  int* d_idx; int* sink;
  CK(cudaMalloc(&d_idx, n_idx * 4)); CK(cudaMalloc(&sink, 4));
  std::vector<int> h(n_idx);
  std::mt19937_64 rng(1);
  char* x; CK(cudaMalloc(&x, n_rows * 1024)); CK(cudaMemset(x, 1, n_rows * 1024));
  for (int mode = 0; mode < 3; ++mode) {
    // 0 uniform random, 1 sequential (k % n_rows), 2 random but sorted within windows of 4096 (locality)
    if (mode == 0) for (int64_t i = 0; i < n_idx; ++i) h[i] = (int)(rng() % n_rows);
    if (mode == 1) for (int64_t i = 0; i < n_idx; ++i) h[i] = (int)(i % n_rows);
    if (mode == 2) { for (int64_t i = 0; i < n_idx; ++i) h[i] = (int)(rng() % n_rows); for (int64_t i = 0; i + 65536 <= n_idx; i += 65536) std::sort(h.begin() + i, h.begin() + i + 65536); }
    CK(cudaMemcpy(d_idx, h.data(), n_idx * 4, cudaMemcpyHostToDevice));
    const char* mname[] = {"uniform", "sequential", "sorted64k"};
    for (int grid : {148 * 4, 148 * 8}) {
      float ms;
      ms = run<32, 2, 4>(x, 1024, d_idx, n_idx, sink, grid);
      printf("%s grid=%d row=1024B U=4 : %.3f ms  %.1f GB/s\n", mname[mode], grid, ms, n_idx * 1028.0 / ms / 1e6);
      ms = run<32, 2, 8>(x, 1024, d_idx, n_idx, sink, grid);
      printf("%s grid=%d row=1024B U=8 : %.3f ms  %.1f GB/s\n", mname[mode], grid, ms, n_idx * 1028.0 / ms / 1e6);
      ms = run<32, 1, 8>(x, 512, d_idx, n_idx, sink, grid);
      printf("%s grid=%d row=512B U=8 : %.3f ms  %.1f GB/s\n", mname[mode], grid, ms, n_idx * 516.0 / ms / 1e6);
      ms = run<16, 1, 8>(x, 256, d_idx, n_idx, sink, grid);
      printf("%s grid=%d row=256B U=8 : %.3f ms  %.1f GB/s\n", mname[mode], grid, ms, n_idx * 260.0 / ms / 1e6);
      ms = run<8, 1, 8>(x, 128, d_idx, n_idx, sink, grid);
      printf("%s grid=%d row=128B U=8 : %.3f ms  %.1f GB/s\n", mname[mode], grid, ms, n_idx * 132.0 / ms / 1e6);
      fflush(stdout);
    }
  }
  return 0;
}
