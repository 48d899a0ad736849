// Probe (sm_100a): how fast can W warps of one SM sub-partition run the softmax inner loop of extend_tc_kernel --
// per element FFMA (scale), MUFU.EX2, FADD (row sum), half an F2FP (pack) -- on E register-resident values per
// thread?  Prints clk per pass for W = 1, 2, 4 warps per sub-partition (x 4 sub-partitions) and E = 128 / 64.
// Tooling only: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/probe_exp_rate tools/probe_exp_rate.cu
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

template <int E>
__global__ void __launch_bounds__(512, 1) probe(int iters, float c, float mc, long long* out, uint32_t* sink) {
  float s[E];
#pragma unroll
  for (int j = 0; j < E; ++j) s[j] = -0.01f * (float)((threadIdx.x * 7 + j) & 63);
  uint32_t acc = 0;
  float l = 0.f;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
    for (int j = 0; j < E; j += 2) {
      const float p0 = ex2(fmaf(s[j], c, -mc));
      const float p1 = ex2(fmaf(s[j + 1], c, -mc));
      sum0 += p0;
      sum1 += p1;
      acc ^= pack(p0, p1);
    }
    l += sum0 + sum1;
    mc += 1e-7f * l;   // the next pass depends on this one (as the next tile's max does)
  }
  const long long t1 = clock64();
  if ((threadIdx.x & 31) == 0) out[threadIdx.x >> 5] = t1 - t0;
  sink[threadIdx.x] = acc + __float_as_uint(l);
}

int main() {
  long long* out;
  uint32_t* sink;
  cudaMalloc(&out, 16 * sizeof(long long));
  cudaMalloc(&sink, 512 * 4);
  const int iters = 2000;
  for (int e : {128, 64})
    for (int w : {1, 2, 4}) {
      long long h[16];
      for (int rep = 0; rep < 2; ++rep) {
        if (e == 128) probe<128><<<1, w * 128>>>(iters, 1.3f, 0.5f, out, sink);
        else probe<64><<<1, w * 128>>>(iters, 1.3f, 0.5f, out, sink);
        if (cudaDeviceSynchronize() != cudaSuccess) {
          printf("error %s\n", cudaGetErrorString(cudaGetLastError()));
          return 1;
        }
      }
      cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
      long long mx = 0;
      for (int i = 0; i < 4 * w; ++i) mx = h[i] > mx ? h[i] : mx;
      printf("{\"elements_per_thread\": %d, \"warps_per_subpartition\": %d, \"clk_per_pass\": %.0f, \"clk_per_mufu_per_subpartition\": %.2f}\n",
             e, w, (double)mx / iters, (double)mx / iters / (e * w));
    }
  return 0;
}
