// Probe (sm_100a): pins the tcgen05 encodings the extend kernel relies on, against a CPU matmul.
//   test 1  SS MMA: D[128x128] = A[128xK] . B[128xK]^T, both K-major, SWIZZLE_128B, K = 128 (two
//           64-element slabs), descriptor advanced by 32 B per K = 16 step
//   test 2  TS MMA: D[128x128] = P[128xK] . V[Kx128], P packed bf16x2 in TMEM (lane = row, column c =
//           elements 2c, 2c + 1), V MN-major SWIZZLE_128B (token rows of 128 B per 64-wide slab)
// Tooling only: nvcc -gencode arch=compute_100a,code=sm_100a -o tools/bin/probe_umma tools/probe_umma.cu
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= 1ull << 46;  // descriptor version (sm_100)
  d |= 2ull << 61;  // SWIZZLE_128B
  return d;
}

__device__ __forceinline__ uint32_t make_idesc(int M, int N, int a_major, int b_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_major << 15) | ((uint32_t)b_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p, e;\n\telect.sync _|e, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a_tmem, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p, e;\n\telect.sync _|e, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "@e tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d), "r"(a_tmem), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit(uint32_t bar) {
  asm volatile("{\n\t.reg .pred e;\n\telect.sync _|e, 0xffffffff;\n\t@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  while (!ok) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  }
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%32], "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31};"
      ::"r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
        "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
        "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]), "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

constexpr int SLAB = 128 * 128;  // bytes: 128 rows of 128 B
template <int VARIANT>
__device__ __forceinline__ uint32_t time_variant(uint32_t tmem, uint8_t* sA, uint8_t* sB, uint8_t* sV, uint64_t* bar,
                                                 uint32_t par, float* D2, int t) {
  constexpr bool ts = VARIANT == 2 || VARIANT == 3, mn = VARIANT & 1;
  constexpr int nn = VARIANT == 6 ? 256 : (VARIANT == 7 ? 64 : 128);
  constexpr bool alt = VARIANT == 4 || VARIANT == 5;
  const uint32_t idesc = make_idesc(128, nn, 0, mn);
  const uint64_t a_base = make_desc(smem_u32(sA), 16, 1024);
  const uint64_t b_base = mn ? make_desc(smem_u32(sV), SLAB, 1024) : make_desc(smem_u32(sB), 16, 1024);
  const long long c0 = clock64();
#pragma unroll 1
  for (int it = 0; it < 8; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t koff = (k >> 2) * SLAB + (k & 3) * 32;
      const uint64_t bd = b_base + (mn ? (k * 2048) >> 4 : koff >> 4);
      const uint32_t dd = tmem + ((alt && (k & 1)) ? 384 : 128);
      if (ts) mma_ts(dd, tmem + 256 + k * 8, bd, idesc, 1);
      else mma_ss(dd, a_base + (koff >> 4), bd, idesc, 1);
    }
  }
  const long long c1 = clock64();
  commit(smem_u32(bar));
  mbar_wait(smem_u32(bar), par);
  const long long c2 = clock64();
  if (t == 0) {
    D2[128 * 128 + VARIANT * 2] = (float)(c1 - c0);
    D2[128 * 128 + VARIANT * 2 + 1] = (float)(c2 - c0);
  }
  return par ^ 1;
}

constexpr int M = 128, N = 128, K = 128;

// A, B: [128][K] row-major bf16 (K-major).  P: [128][K].  V: [K tokens][N d] row-major.
// mode (timing section only): warps 4-15 interfere while warp 0 times the MMAs -- bit 0: tcgen05.ld x32 loops on
// the S columns (what the softmax warps do), bit 1: 16-byte shared-memory stores (what the loaders' cp.async does),
// bit 2: MUFU.EX2 loops
__global__ void __launch_bounds__(512) probe(const __nv_bfloat16* A, const __nv_bfloat16* B, const __nv_bfloat16* P,
                                             const __nv_bfloat16* V, float* D1, float* D2, int mode) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                 // 2 slabs
  uint8_t* sB = smem + 2 * SLAB;      // 2 slabs
  uint8_t* sV = smem + 4 * SLAB;      // 2 slabs (d 0-63, d 64-127), rows = tokens
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  __shared__ volatile int done_flag;
  __shared__ uint4 scratch[512];
  const int t = threadIdx.x, warp = t >> 5;
  if (t == 0) done_flag = 0;
  if (t >= 128) {
    // interferers: wait for the set-up, then run until warp 0 has finished timing
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();   // matches the set-up barrier below
    __syncthreads();   // test 2 staging barrier
    __syncthreads();   // timing start
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tmem_base_s + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t acc = 0;
    float f = (float)t;
    while (mode && !done_flag) {
      if (mode & 1) {
        uint32_t r[32];
        tmem_ld32(tm + (warp & 4 ? 32 : 0), r);
        for (int j = 0; j < 32; ++j) acc ^= r[j];
      }
      if (mode & 2) {
        for (int j = 0; j < 8; ++j) scratch[(t + 32 * j) & 511] = make_uint4(acc, t, j, 0);
      }
      if (mode & 4) {
        for (int j = 0; j < 32; ++j) f = exp2f(f * 0.001f);
      }
    }
    if (acc == 0x12345 || f == 3.f) D1[0] = 1.f;
    __syncthreads();   // end
    return;
  }

  // K-major swizzled fill: row r, 16-byte chunk c (0..15): slab c / 8, chunk (c % 8) ^ (r % 8)
  for (int i = t; i < 128 * 16; i += 128) {
    const int r = i >> 4, c = i & 15;
    const uint32_t off = (c >> 3) * SLAB + r * 128 + (((c & 7) ^ (r & 7)) << 4);
    *(uint4*)(sA + off) = *(const uint4*)((const uint8_t*)A + (size_t)r * K * 2 + c * 16);
    *(uint4*)(sB + off) = *(const uint4*)((const uint8_t*)B + (size_t)r * K * 2 + c * 16);
    *(uint4*)(sV + off) = *(const uint4*)((const uint8_t*)V + (size_t)r * N * 2 + c * 16);  // r = token, c = d chunk
  }
  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base_s)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = __shfl_sync(0xffffffffu, tmem_base_s, 0);
  const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);

  // ---- test 1: SS ----
  if (warp == 0) {
    const uint32_t idesc = make_idesc(M, N, 0, 0);
    for (int k = 0; k < K / 16; ++k) {
      const uint32_t off = (k >> 2) * SLAB + (k & 3) * 32;
      mma_ss(tmem, make_desc(smem_u32(sA) + off, 16, 1024), make_desc(smem_u32(sB) + off, 16, 1024), idesc, k > 0);
    }
    commit(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int c0 = 0; c0 < N; c0 += 32) {
    uint32_t r[32];
    tmem_ld32(lane_addr + c0, r);
    for (int j = 0; j < 32; ++j) D1[(size_t)t * N + c0 + j] = __uint_as_float(r[j]);
  }

  // ---- test 2: TS.  P packed into TMEM columns 256.. (row = this thread) ----
  for (int c0 = 0; c0 < K / 2; c0 += 32) {
    uint32_t r[32];
    for (int j = 0; j < 32; ++j) r[j] = *(const uint32_t*)(P + (size_t)t * K + 2 * (c0 + j));
    tmem_st32(lane_addr + 256 + c0, r);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (warp == 0) {
    const uint32_t idesc = make_idesc(M, N, 0, 1);  // B = V is MN-major
    for (int k = 0; k < K / 16; ++k) {
      // 16 tokens per step = two 8-row groups of 1024 B; the two 64-wide d slabs are SLAB apart
      mma_ts(tmem + 128, tmem + 256 + k * 8, make_desc(smem_u32(sV) + k * 2048, SLAB, 1024), idesc, k > 0);
    }
    commit(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 1);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int c0 = 0; c0 < N; c0 += 32) {
    uint32_t r[32];
    tmem_ld32(lane_addr + 128 + c0, r);
    for (int j = 0; j < 32; ++j) D2[(size_t)t * N + c0 + j] = __uint_as_float(r[j]);
  }
  // ---- timing: 64 back-to-back MMAs (M=128, N=128, K=16 each) per variant, one commit at the end ----
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (warp == 0) {
    uint32_t par = 0;
    par = time_variant<0>(tmem, sA, sB, sV, &bar, par, D2, t);
    par = time_variant<1>(tmem, sA, sB, sV, &bar, par, D2, t);
    par = time_variant<2>(tmem, sA, sB, sV, &bar, par, D2, t);
    par = time_variant<3>(tmem, sA, sB, sV, &bar, par, D2, t);
    par = time_variant<4>(tmem, sA, sB, sV, &bar, par, D2, t);
    par = time_variant<5>(tmem, sA, sB, sV, &bar, par, D2, t);
    par = time_variant<6>(tmem, sA, sB, sV, &bar, par, D2, t);
    par = time_variant<7>(tmem, sA, sB, sV, &bar, par, D2, t);
    done_flag = 1;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

int main() {
  std::vector<__nv_bfloat16> A(M * K), B(N * K), P(M * K), V(K * N);
  srand(1);
  auto rnd = [] { return __float2bfloat16((float)(rand() % 17 - 8) / 8.0f); };
  for (auto& x : A) x = rnd();
  for (auto& x : B) x = rnd();
  for (auto& x : P) x = rnd();
  for (auto& x : V) x = rnd();
  __nv_bfloat16 *dA, *dB, *dP, *dV;
  float *d1, *d2;
  cudaMalloc(&dA, A.size() * 2); cudaMalloc(&dB, B.size() * 2); cudaMalloc(&dP, P.size() * 2); cudaMalloc(&dV, V.size() * 2);
  cudaMalloc(&d1, M * N * 4); cudaMalloc(&d2, M * N * 4 + 64);
  cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dP, P.data(), P.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dV, V.data(), V.size() * 2, cudaMemcpyHostToDevice);
  const int smem = 6 * SLAB + 1024;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaError_t e = cudaSuccess;
  for (int mode = 0; mode < 8 && e == cudaSuccess; ++mode) {
    probe<<<1, 512, smem>>>(dA, dB, dP, dV, d1, d2, mode);
    e = cudaDeviceSynchronize();
    float tm[16];
    cudaMemcpy(tm, d2 + M * N, sizeof(tm), cudaMemcpyDeviceToHost);
    printf("interference mode %d (1 = tcgen05.ld, 2 = st.shared, 4 = MUFU): clk / MMA  SS K-major %.1f | TS MN-major %.1f | SS N=64 %.1f | SS N=256 %.1f\n",
           mode, tm[1] / 64, tm[7] / 64, tm[15] / 64, tm[13] / 64);
  }
  probe<<<1, 512, smem>>>(dA, dB, dP, dV, d1, d2, 0);
  e = cudaDeviceSynchronize();
  printf("kernel: %s\n", cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  std::vector<float> h1(M * N), h2(M * N);
  cudaMemcpy(h1.data(), d1, M * N * 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(h2.data(), d2, M * N * 4, cudaMemcpyDeviceToHost);
  double e1 = 0, e2 = 0;
  for (int i = 0; i < M; ++i)
    for (int j = 0; j < N; ++j) {
      double s1 = 0, s2 = 0;
      for (int k = 0; k < K; ++k) {
        s1 += (double)__bfloat162float(A[i * K + k]) * __bfloat162float(B[j * K + k]);
        s2 += (double)__bfloat162float(P[i * K + k]) * __bfloat162float(V[k * N + j]);
      }
      e1 = fmax(e1, fabs(s1 - h1[i * N + j]));
      e2 = fmax(e2, fabs(s2 - h2[i * N + j]));
    }
  {
    float tm[16];
    cudaMemcpy(tm, d2 + M * N, sizeof(tm), cudaMemcpyDeviceToHost);
    const char* nm[8] = {"SS B K-major", "SS B MN-major", "TS B K-major", "TS B MN-major", "SS K alt-acc", "SS MN alt-acc", "SS K N=256", "SS MN N=64"};
    for (int v = 0; v < 8; ++v) printf("timing %-14s: issue %.0f clk, done %.0f clk for 64 MMAs (%.1f clk / MMA)\n", nm[v], tm[2 * v], tm[2 * v + 1], tm[2 * v + 1] / 64);
  }
  printf("test1 SS K-major x K-major SW128: max err %.3e %s\n", e1, e1 < 1e-3 ? "PASS" : "FAIL");
  printf("test2 TS P(tmem) x V MN-major SW128: max err %.3e %s\n", e2, e2 < 1e-3 ? "PASS" : "FAIL");
  return (e1 < 1e-3 && e2 < 1e-3) ? 0 : 2;
}
