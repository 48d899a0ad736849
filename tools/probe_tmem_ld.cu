// Probe (sm_100a): tcgen05.ld throughput per SM.  W warps of one CTA (warp w reads TMEM lane quarter w % 4) each
// issue `iters` x (tcgen05.ld 32x32b.x32 + wait::ld) = 4 KB per instruction; reports clk per instruction and the
// aggregate bytes / clk for W = 4 (one warp per sub-partition), 8 and 16, alone and with a tensor-pipe MMA stream
// absent.  Answers: is the S tile's read (64 KB per 128 x 128 fp32 tile) 1024 clk (64 B/clk per SM) or 256 clk
// (64 B/clk per sub-partition)?
// Tooling only: nvcc -gencode arch=compute_100a,code=sm_100a -o tools/bin/probe_tmem_ld tools/probe_tmem_ld.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

#define R16(r, o) r[o + 0], r[o + 1], r[o + 2], r[o + 3], r[o + 4], r[o + 5], r[o + 6], r[o + 7], r[o + 8], \
    r[o + 9], r[o + 10], r[o + 11], r[o + 12], r[o + 13], r[o + 14], r[o + 15]

__device__ __forceinline__ void ld32(uint32_t taddr, uint32_t (&u)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,"
      "%29,%30,%31}, [%32];"
      : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]), "=r"(u[8]),
        "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15]), "=r"(u[16]),
        "=r"(u[17]), "=r"(u[18]), "=r"(u[19]), "=r"(u[20]), "=r"(u[21]), "=r"(u[22]), "=r"(u[23]), "=r"(u[24]),
        "=r"(u[25]), "=r"(u[26]), "=r"(u[27]), "=r"(u[28]), "=r"(u[29]), "=r"(u[30]), "=r"(u[31])
      : "r"(taddr));
}

// mode 0: ld + wait each time (dependent); mode 1: two lds in flight per wait
__global__ void __launch_bounds__(512, 1) probe(int iters, int mode, long long* out, uint32_t* sink) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = slot;
  const uint32_t base = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * 64 % 512);
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  if (mode == 0) {
    for (int i = 0; i < iters; ++i) {
      uint32_t u[32];
      ld32(base + (i & 1) * 32, u);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int j = 0; j < 32; ++j) acc ^= u[j];
    }
  } else {
    for (int i = 0; i < iters; i += 2) {
      uint32_t u[32], v[32];
      ld32(base, u);
      ld32(base + 32, v);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int j = 0; j < 32; ++j) acc ^= u[j] ^ v[j];
    }
  }
  const long long t1 = clock64();
  if ((threadIdx.x & 31) == 0) out[blockIdx.x * 16 + warp] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
  }
}

int main() {
  long long* out;
  uint32_t* sink;
  cudaMalloc(&out, 148 * 16 * sizeof(long long));
  cudaMalloc(&sink, 148 * 512 * 4);
  const int iters = 4096;
  for (int mode = 0; mode < 2; ++mode)
    for (int warps : {1, 4, 8, 16}) {
      long long h[16];
      for (int rep = 0; rep < 2; ++rep) {
        probe<<<1, warps * 32>>>(iters, mode, out, sink);
        if (cudaDeviceSynchronize() != cudaSuccess) {
          printf("error: %s\n", cudaGetErrorString(cudaGetLastError()));
          return 1;
        }
      }
      cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
      long long mx = 0;
      for (int w = 0; w < warps; ++w) mx = h[w] > mx ? h[w] : mx;
      printf("{\"mode\": \"%s\", \"warps\": %d, \"clk_per_ld_per_warp\": %.1f, \"sm_bytes_per_clk\": %.1f}\n",
             mode == 0 ? "ld+wait" : "2 lds per wait", warps, (double)mx / iters,
             (double)warps * iters * 4096.0 / (double)mx);
    }
  return 0;
}
