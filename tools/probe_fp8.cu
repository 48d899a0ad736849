// Probes (sm_100a): (1) fragment layout of ldmatrix.m16n16.trans.b8, (2) issue rate of
// mma.sync m16n8k32 e4m3 vs m16n8k16 f16.  Tooling only.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__global__ void probe_ldsm(uint32_t* out) {
  __shared__ __align__(128) uint8_t m[2][16][16];
  for (int i = threadIdx.x; i < 512; i += 32) (&m[0][0][0])[i] = (uint8_t)i;  // matrix 0: r*16+c, matrix 1: 256 + ...
  __syncwarp();
  uint32_t addr = (uint32_t)__cvta_generic_to_shared(&m[0][threadIdx.x & 15][0]);
  uint32_t r0, r1;
  asm volatile("ldmatrix.sync.aligned.m16n16.x1.trans.shared.b8 {%0,%1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(addr));
  out[threadIdx.x * 2] = r0;
  out[threadIdx.x * 2 + 1] = r1;
  uint32_t addr2 = (uint32_t)__cvta_generic_to_shared(&m[threadIdx.x >> 4][threadIdx.x & 15][0]);
  uint32_t q0, q1, q2, q3;
  asm volatile("ldmatrix.sync.aligned.m16n16.x2.trans.shared.b8 {%0,%1,%2,%3}, [%4];" : "=r"(q0), "=r"(q1), "=r"(q2), "=r"(q3) : "r"(addr2));
  out[64 + threadIdx.x * 4] = q0;
  out[64 + threadIdx.x * 4 + 1] = q1;
  out[64 + threadIdx.x * 4 + 2] = q2;
  out[64 + threadIdx.x * 4 + 3] = q3;
}

template <int MODE>
__global__ void rate(float* out, int iters) {
  float d[4][4] = {};
  uint32_t a0 = threadIdx.x, a1 = 1, a2 = 2, a3 = 3, b0 = 5, b1 = 7;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (MODE == 0)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(d[k][0]), "+f"(d[k][1]), "+f"(d[k][2]), "+f"(d[k][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
      else
        asm volatile("mma.sync.aligned.m16n8k32.row.col.f32.e4m3.e4m3.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(d[k][0]), "+f"(d[k][1]), "+f"(d[k][2]), "+f"(d[k][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
  }
  float s = 0;
  for (int k = 0; k < 4; ++k) for (int e = 0; e < 4; ++e) s += d[k][e];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  uint32_t* d; cudaMalloc(&d, 4096);
  probe_ldsm<<<1, 32>>>(d);
  uint32_t h[192];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("x1: lane: r0 r1 (bytes = row*16+col of the smem matrix)\n");
  for (int l = 0; l < 32; ++l) printf("lane %2d: %08x %08x\n", l, h[2 * l], h[2 * l + 1]);
  printf("x2 (matrix 1 rows carry +256 -> wraps mod 256, same byte pattern):\n");
  for (int l = 0; l < 32; ++l) printf("lane %2d: %08x %08x %08x %08x\n", l, h[64 + 4 * l], h[64 + 4 * l + 1], h[64 + 4 * l + 2], h[64 + 4 * l + 3]);
  float* o; cudaMalloc(&o, 148 * 256 * 4);
  for (int mode = 0; mode < 2; ++mode) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (mode == 0) rate<0><<<148, 256>>>(o, iters); else rate<1><<<148, 256>>>(o, iters);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
    }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double mmas = 148.0 * 8 * iters * 4;
    printf("mode %d (%s): %.3f ms, %.1f G mma/s, %.1f TFLOP/s, %.2f clk/mma/SM-subpartition @1.965GHz\n", mode, mode ? "e4m3 m16n8k32" : "f16 m16n8k16", ms,
           mmas / ms / 1e6, mmas * (mode ? 16 * 8 * 32 * 2 : 16 * 8 * 16 * 2) / ms / 1e9, ms * 1e-3 * 1.965e9 / (iters * 4.0 * 2));
  }
  printf("err: %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
