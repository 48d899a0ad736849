// Probe (sm_100a): fragment layouts the transposed decode consumer relies on -- movmatrix.trans.b16 and
// ldmatrix.m16n16.x2.trans.b8 -- printed as (row, col) of the source.  Tooling only.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
__global__ void probe(uint32_t* out) {
  const int lane = threadIdx.x;
  // movmatrix: thread (g, j) holds row g, cols 2j, 2j+1 of an 8x8 b16 matrix; element value = row * 8 + col
  const int g = lane >> 2, j = lane & 3;
  uint32_t x = (uint32_t)(g * 8 + 2 * j) | ((uint32_t)(g * 8 + 2 * j + 1) << 16), y;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(y) : "r"(x));
  out[lane] = y;
  __shared__ __align__(128) uint8_t m[32][16];   // 32 rows (tokens) x 16 bytes; value = row * 16 + col (mod 256)
  for (int i = lane; i < 512; i += 32) (&m[0][0])[i] = (uint8_t)i;
  __syncwarp();
  uint32_t addr = (uint32_t)__cvta_generic_to_shared(&m[lane][0]);
  uint32_t q0, q1, q2, q3;
  asm volatile("ldmatrix.sync.aligned.m16n16.x2.trans.shared.b8 {%0,%1,%2,%3}, [%4];" : "=r"(q0), "=r"(q1), "=r"(q2), "=r"(q3) : "r"(addr));
  out[32 + lane * 4] = q0; out[32 + lane * 4 + 1] = q1; out[32 + lane * 4 + 2] = q2; out[32 + lane * 4 + 3] = q3;
}
int main() {
  uint32_t* d; cudaMalloc(&d, 4096);
  probe<<<1, 32>>>(d);
  uint32_t h[160];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("movmatrix.trans: lane (g,j) -> low = r*8+c, high\n");
  for (int l = 0; l < 32; ++l) printf("lane %2d (g%d j%d): lo (r%d,c%d) hi (r%d,c%d)\n", l, l >> 2, l & 3, (h[l] & 0xffff) / 8, (h[l] & 0xffff) % 8, (h[l] >> 16) / 8, (h[l] >> 16) % 8);
  printf("ldmatrix m16n16.x2.trans.b8, rows 0-15 = matrix 0 (lanes 0-15 give the addresses), rows 16-31 = matrix 1; bytes as (row,col) of matrix:\n");
  for (int l = 0; l < 32; ++l) {
    printf("lane %2d (g%d j%d):", l, l >> 2, l & 3);
    for (int r = 0; r < 4; ++r) {
      printf("  q%d:", r);
      for (int b = 0; b < 4; ++b) { int v = (h[32 + 4 * l + r] >> (8 * b)) & 0xff; printf(" (%d,%d)", v / 16, v % 16); }
    }
    printf("\n");
  }
  printf("err: %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
