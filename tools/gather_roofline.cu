// Random-row gather bandwidth of the HBM system as a function of the row size: the ceiling a
// paged-KV gather with page_size 1 can reach when each token row (all local kv heads of one slot)
// is R bytes at a random slot.  Plain 16-byte loads, full occupancy, 8 loads in flight per thread:
// no shared memory, no compute -- this is the memory system's number, not a kernel's.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/bin/gather_roofline tools/gather_roofline.cu
//   tools/bin/gather_roofline            -> one JSON line per row size
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <numeric>
#include <random>
#include <vector>

__global__ void gather_kernel(const uint4* __restrict__ pool, const int32_t* __restrict__ rows,
                              int64_t n_rows, int units_per_row_shift, uint4* __restrict__ sink) {
  const int64_t n_units = n_rows << units_per_row_shift;
  const int umask = (1 << units_per_row_shift) - 1;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  uint4 acc = make_uint4(0, 0, 0, 0);
  int64_t u = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  for (; u + 7 * stride < n_units; u += 8 * stride) {
    uint4 v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int64_t uu = u + k * stride;
      const int64_t row = rows[uu >> units_per_row_shift];
      const uint4* src = pool + (row << units_per_row_shift) + (uu & umask);
      asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                   : "=r"(v[k].x), "=r"(v[k].y), "=r"(v[k].z), "=r"(v[k].w)
                   : "l"(src));
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      acc.x ^= v[k].x; acc.y ^= v[k].y; acc.z ^= v[k].z; acc.w ^= v[k].w;
    }
  }
  if ((acc.x ^ acc.y ^ acc.z ^ acc.w) == 0x12345678u) sink[0] = acc;  // never true in practice
}

int main(int argc, char** argv) {
  const size_t pool_bytes = argc > 1 ? strtoull(argv[1], nullptr, 10) << 20 : (size_t)4096 << 20;
  const size_t read_bytes = pool_bytes / 2;  // every launch reads half the pool, rows distinct
  uint4 *pool, *sink;
  cudaMalloc(&pool, pool_bytes);
  cudaMalloc(&sink, 64);
  cudaMemset(pool, 1, pool_bytes);
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  std::mt19937_64 rng(1234);
  for (int sequential = 0; sequential < 2; ++sequential) {
    for (int R = 64; R <= 4096; R *= 2) {
      const int64_t pool_rows = pool_bytes / R;
      const int64_t n = read_bytes / R;
      std::vector<int32_t> perm(pool_rows);
      std::iota(perm.begin(), perm.end(), 0);
      if (!sequential) std::shuffle(perm.begin(), perm.end(), rng);
      int32_t* d_rows;
      cudaMalloc(&d_rows, n * 4);
      cudaMemcpy(d_rows, perm.data(), n * 4, cudaMemcpyHostToDevice);
      int shift = 0;
      while ((16 << shift) < R) ++shift;
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0);
      cudaEventCreate(&e1);
      float best = 1e30f;
      for (int it = 0; it < 6; ++it) {
        cudaEventRecord(e0);
        gather_kernel<<<sms * 8, 256>>>(pool, d_rows, n, shift, sink);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (it > 0) best = std::min(best, ms);
      }
      if (cudaGetLastError() != cudaSuccess) { printf("{\"error\": \"launch failed\"}\n"); return 1; }
      printf("{\"row_bytes\": %d, \"order\": \"%s\", \"gb_per_s\": %.1f, \"gb_per_s_with_index\": %.1f, \"ms\": %.4f, \"read_mb\": %zu}\n",
             R, sequential ? "sequential" : "random", n * (double)R / best / 1e6, n * (double)(R + 4) / best / 1e6, best,
             read_bytes >> 20);
      cudaFree(d_rows);
    }
  }
  return 0;
}
