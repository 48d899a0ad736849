// K5: extend (prefill with cached prefix) attention over the paged KV pool (sm_100a).
//
// Replaces the reference's Triton extend kernel
// (python/sglang/srt/layers/attention/triton_ops/extend_attention.py:41-303) and computes the math
// of the torch_native oracle (layers/attention/torch_native_backend.py:27-110).
//
// One CTA = (request, kv head, q tile).  A q tile is 128 MMA rows: 128 / Gp tokens x Gp q-head
// slots (Gp = group size rounded up to a power of two), so all q heads that share the kv head
// ride in the M dimension and every K/V row is fetched once per tile.  8 warps x 16 rows.
// K/V rows (prefix through the index list, then the extend tokens through out_cache_loc) are
// gathered 64 tokens at a time with 16-byte cp.async into a 3-stage ring of padded rows
// (bank-conflict-free ldmatrix); slot ids are fetched two tiles ahead.  S = Q.K^T and O += P.V
// on mma.sync.m16n8k16 with fp32 accumulators, online softmax in registers (quad shuffles).
#pragma once

#include <math_constants.h>

#include <type_traits>

#include "decode_kernel.cuh"

namespace b200attn {

constexpr int kExtWarps = 8;
constexpr int kExtThreads = kExtWarps * 32;
constexpr int kExtRows = kExtWarps * 16;  // MMA rows per CTA
constexpr int kExtBN = 64;                // kv tokens per stage
constexpr int kExtStages = 3;

struct ExtendParams {
  const void* q;
  int64_t q_st, q_sh;  // elements
  void* o;
  int64_t o_st, o_sh;
  float* lse;
  const uint8_t* k_pool;
  const uint8_t* v_pool;
  int64_t k_slot_bytes, v_slot_bytes;
  const int32_t* qo_indptr;
  const int32_t* kv_indptr;
  const int32_t* kv_indices;
  const int32_t* req_to_token;
  int64_t r2t_stride;
  const void* req_pool_indices;
  const void* prefix_lens;
  const int64_t* extend_loc;
  const uint8_t* custom_mask;
  const void* mask_indptr;
  int32_t idx64, mask64;
  int32_t hq, group, gp_shift;
  int32_t causal, window, skip_prefix_mask;
  float scale_log2;  // sm_scale * log2(e)
  float sm_scale;
  float logit_cap;
  float v_scale;
  int32_t batch;
  int32_t e5m2;      // fp8 pools: 1 = float8_e5m2 bytes (the byte is the upper half of an f16), 0 = e4m3fn
};

__host__ __device__ inline int extend_smem_bytes(int d) {
  return kExtStages * 2 * kExtBN * (d * 2 + kRowPad);
}

// 16 fp8 bytes -> 16 f16 values (exact), in element order; e5m2 (uniform flag): byte permutes instead of cvt.
__device__ __forceinline__ void e4m3x16_to_f16(const uint4& r, uint4& lo, uint4& hi, bool e5m2 = false) {
  if (e5m2) {
    lo.x = __byte_perm(r.x, 0u, 0x1404);
    lo.y = __byte_perm(r.x, 0u, 0x3424);
    lo.z = __byte_perm(r.y, 0u, 0x1404);
    lo.w = __byte_perm(r.y, 0u, 0x3424);
    hi.x = __byte_perm(r.z, 0u, 0x1404);
    hi.y = __byte_perm(r.z, 0u, 0x3424);
    hi.z = __byte_perm(r.w, 0u, 0x1404);
    hi.w = __byte_perm(r.w, 0u, 0x3424);
    return;
  }
  lo.x = e4m3x2_to_f16x2(r.x);
  lo.y = e4m3x2_to_f16x2(r.x >> 16);
  lo.z = e4m3x2_to_f16x2(r.y);
  lo.w = e4m3x2_to_f16x2(r.y >> 16);
  hi.x = e4m3x2_to_f16x2(r.z);
  hi.y = e4m3x2_to_f16x2(r.z >> 16);
  hi.z = e4m3x2_to_f16x2(r.w);
  hi.w = e4m3x2_to_f16x2(r.w >> 16);
}

// Two q elements -> the MMA operand type: unchanged, or (fp8 pool with bf16 q) f16 clamped to its range,
// the same conversion the decode kernel applies to q.
template <typename QT, bool TO_HALF>
__device__ __forceinline__ uint32_t q_pair(uint32_t x) {
  if constexpr (TO_HALF && !std::is_same<QT, __half>::value) {
    const float lo = fminf(fmaxf(__uint_as_float(x << 16), -65504.f), 65504.f);
    const float hi = fminf(fmaxf(__uint_as_float(x & 0xffff0000u), -65504.f), 65504.f);
    return pack2<true>(lo, hi);
  } else {
    return x;
  }
}

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// FP8: the pool holds float8_e4m3fn bytes.  Rows travel as stored (1 B / element) into the upper half
// of their shared-memory row and are widened in place to f16 (exact) once landed; both MMAs then run in
// f16 (q converted and clamped like the decode kernel does), k_scale rides in the softmax scale and
// v_scale in the final normalisation (FlashInfer's semantics, flashinfer_backend.py:560-567).
// CAP: logit cap compiled in (as a run-time flag the compiler if-converts it and evaluates tanh for every
// element of every tile, cap or not).
template <typename QT, int D, bool FP8, bool CAP>
__global__ void __launch_bounds__(kExtThreads, 1) extend_kernel(const ExtendParams p) {
  constexpr bool OUT_HALF = std::is_same<QT, __half>::value;
  constexpr bool HALF = OUT_HALF || FP8;     // MMA operand type
  constexpr int ESZ = FP8 ? 1 : 2;           // pool element bytes
  constexpr int KSTEPS = D / 16;
  constexpr int NTILES = D / 8;
  constexpr int RS = D * 2 + kRowPad;        // smem row stride, bytes
  constexpr int CH = D * ESZ / 16;           // 16-byte chunks per pool row
  constexpr int RAW_OFF = FP8 ? D : 0;       // fp8 bytes land in the upper half of the row
  constexpr int RPP = kExtThreads / CH;      // rows per loader pass
  constexpr int PASSES = (kExtBN + RPP - 1) / RPP;
  // head_dim 80 / 96 (10 / 12 chunks per row): the threads past RPP * CH idle and the last pass is ragged
  constexpr bool EXACT = kExtThreads % CH == 0 && kExtBN % RPP == 0;
  static_assert(D % 16 == 0 && (D / 8) % 2 == 0, "head_dim: whole k-steps, n8 tiles in pairs");
  static_assert(!FP8 || EXACT, "fp8 pools: head dims 64 / 128 / 256");
  constexpr int TENSOR_BYTES = kExtBN * RS;
  constexpr int STAGE_BYTES = 2 * TENSOR_BYTES;
  constexpr int NT = kExtBN / 8;             // n8 tiles of S per stage

  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t stage0 = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y;

  const int q0 = p.qo_indptr[b];
  const int Le = p.qo_indptr[b + 1] - q0;
  const int bm_tok = kExtRows >> p.gp_shift;
  const int ntq = (Le + bm_tok - 1) / bm_tok;
  const int t = ntq - 1 - static_cast<int>(blockIdx.x);  // the longest (last) tiles are scheduled first
  if (t < 0) return;
  const int tok0 = t * bm_tok;
  const int kv0 = p.kv_indptr[b];
  const int Lp = p.kv_indptr[b + 1] - kv0;
  const int32_t* pidx;
  if (p.kv_indices) {
    pidx = p.kv_indices + kv0;
  } else {
    const int64_t req = p.idx64 ? static_cast<const int64_t*>(p.req_pool_indices)[b]
                                : static_cast<const int32_t*>(p.req_pool_indices)[b];
    const int64_t pl = p.idx64 ? static_cast<const int64_t*>(p.prefix_lens)[b]
                               : static_cast<const int32_t*>(p.prefix_lens)[b];
    pidx = p.req_to_token + req * p.r2t_stride + (pl - Lp);
  }
  const bool use_mask = p.custom_mask != nullptr;
  const int tile_tok_end = tok0 + bm_tok < Le ? tok0 + bm_tok : Le;
  const int ext_end = (p.causal && !use_mask) ? tile_tok_end : Le;
  const int Ltot = Lp + ext_end;
  const int ntiles = (Ltot + kExtBN - 1) / kExtBN;
  const int64_t mask_row_len = static_cast<int64_t>(Lp) + Le;
  const uint8_t* mask = nullptr;
  if (use_mask) {
    const int64_t ms = p.mask64 ? static_cast<const int64_t*>(p.mask_indptr)[b]
                                : static_cast<const int32_t*>(p.mask_indptr)[b];
    mask = p.custom_mask + ms;
  }

  // ------------------------------ loader ------------------------------
  const int ld_row = threadIdx.x / CH, ld_chunk = threadIdx.x % CH;
  const uint8_t* k_head = p.k_pool + static_cast<int64_t>(h) * D * ESZ + ld_chunk * 16;
  const uint8_t* v_head = p.v_pool + static_cast<int64_t>(h) * D * ESZ + ld_chunk * 16;
  int32_t nslot[PASSES];
  auto fetch_slots = [&](int tile) {
#pragma unroll
    for (int i = 0; i < PASSES; ++i) {
      const int j = tile * kExtBN + ld_row + i * RPP;
      int32_t s = -1;
      if (j < Ltot) s = j < Lp ? pidx[j] : static_cast<int32_t>(p.extend_loc[q0 + j - Lp]);
      nslot[i] = s;
    }
  };
  // (ragged loader only) this thread copies row ld_row + i * RPP of a stage
  auto ld_active = [&](int i) { return EXACT || (ld_row < RPP && ld_row + i * RPP < kExtBN); };
  auto issue = [&](int stage) {
    const uint32_t dst = stage0 + stage * STAGE_BYTES + ld_row * RS + RAW_OFF + ld_chunk * 16;
#pragma unroll
    for (int i = 0; i < PASSES; ++i) {
      if (!ld_active(i)) continue;
      const bool ok = nslot[i] >= 0;
      const int64_t s = ok ? nslot[i] : 0;
      cp_async16_zfill(dst + i * RPP * RS, k_head + s * p.k_slot_bytes, ok);
      cp_async16_zfill(dst + i * RPP * RS + TENSOR_BYTES, v_head + s * p.v_slot_bytes, ok);
    }
  };

  fetch_slots(0);
  issue(0);
  cp_async_commit();
  fetch_slots(1);
  if (1 < ntiles) issue(1);
  cp_async_commit();
  fetch_slots(2);

  // ------------------------------ this warp's rows ------------------------------
  const int gq = lane >> 2, jq = lane & 3;
  const int mat = lane >> 3, mr = lane & 7;
  const int gmask = (1 << p.gp_shift) - 1;
  const int R0 = warp * 16 + gq, R1 = R0 + 8;
  const int i0 = tok0 + (R0 >> p.gp_shift), i1 = tok0 + (R1 >> p.gp_shift);
  const int hd0 = R0 & gmask, hd1 = R1 & gmask;
  const bool row0 = hd0 < p.group && i0 < Le;
  const bool row1 = hd1 < p.group && i1 < Le;
  const int i_min_warp = tok0 + ((warp * 16) >> p.gp_shift);
  constexpr bool cap = CAP;
  const float c = cap ? kLog2e : p.scale_log2;
  const float cap_pre = cap ? p.sm_scale / p.logit_cap : 0.f;

  uint32_t qf[KSTEPS][4];
  {
    const QT* qr0 = reinterpret_cast<const QT*>(p.q) + static_cast<int64_t>(q0 + i0) * p.q_st +
                    static_cast<int64_t>(h * p.group + hd0) * p.q_sh;
    const QT* qr1 = reinterpret_cast<const QT*>(p.q) + static_cast<int64_t>(q0 + i1) * p.q_st +
                    static_cast<int64_t>(h * p.group + hd1) * p.q_sh;
#pragma unroll
    for (int kk = 0; kk < KSTEPS; ++kk) {
      const int d0 = kk * 16 + 2 * jq;
      qf[kk][0] = row0 ? q_pair<QT, FP8>(*reinterpret_cast<const uint32_t*>(qr0 + d0)) : 0u;
      qf[kk][1] = row0 ? q_pair<QT, FP8>(*reinterpret_cast<const uint32_t*>(qr0 + d0 + 8)) : 0u;
      qf[kk][2] = row1 ? q_pair<QT, FP8>(*reinterpret_cast<const uint32_t*>(qr1 + d0)) : 0u;
      qf[kk][3] = row1 ? q_pair<QT, FP8>(*reinterpret_cast<const uint32_t*>(qr1 + d0 + 8)) : 0u;
    }
  }

  float o[NTILES][4];
#pragma unroll
  for (int n = 0; n < NTILES; ++n) o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;
  float m0 = -CUDART_INF_F, m1 = -CUDART_INF_F, l0 = 0.f, l1 = 0.f;

  const uint32_t k_lane = mr * RS + mat * 16;
  const uint32_t v_lane = TENSOR_BYTES + ((mat & 1) * 8 + mr) * RS + (mat >> 1) * 16;

  for (int n = 0; n < ntiles; ++n) {
    cp_async_wait<1>();
    int stage = n - (n / kExtStages) * kExtStages;
    const uint32_t sbase = stage0 + stage * STAGE_BYTES;
    uint4 raw[FP8 ? PASSES : 1][2];
    if constexpr (FP8) {
      // this thread's own raw bytes of tile n have landed: lift them out before anyone widens over them
      const uint32_t src = sbase + ld_row * RS + RAW_OFF + ld_chunk * 16;
#pragma unroll
      for (int i = 0; i < PASSES; ++i) {
        raw[i][0] = lds128(src + i * RPP * RS);
        raw[i][1] = lds128(src + i * RPP * RS + TENSOR_BYTES);
      }
    }
    __syncthreads();  // tile n has landed for every thread; everyone is done with tile n - 1
    if (n + 2 < ntiles) {
      int st = n + 2;
      st -= (st / kExtStages) * kExtStages;
      issue(st);
    }
    cp_async_commit();
    fetch_slots(n + 3);
    if constexpr (FP8) {
      const uint32_t dst = sbase + ld_row * RS + ld_chunk * 32;
#pragma unroll
      for (int i = 0; i < PASSES; ++i) {
#pragma unroll
        for (int kv = 0; kv < 2; ++kv) {
          uint4 lo, hi;
          e4m3x16_to_f16(raw[i][kv], lo, hi, p.e5m2 != 0);
          sts128(dst + i * RPP * RS + kv * TENSOR_BYTES, lo);
          sts128(dst + i * RPP * RS + kv * TENSOR_BYTES + 16, hi);
        }
      }
      __syncthreads();  // widened tile visible to every warp
    }

    // ---- S = Q.K^T (16 rows x 64 tokens) ----
    float s[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
      const uint32_t ka = sbase + k_lane + nt * 8 * RS;
#pragma unroll
      for (int k2 = 0; k2 < (KSTEPS + 1) / 2; ++k2) {
        uint32_t kf[4];
        ldsm_x4(ka + k2 * 64, kf);   // (head_dim 80: the last load's upper two matrices read the row pad; unused)
        mma16816<HALF>(s[nt], qf[2 * k2][0], qf[2 * k2][2], qf[2 * k2][1], qf[2 * k2][3], kf[0], kf[1]);
        if (2 * k2 + 1 < KSTEPS)
          mma16816<HALF>(s[nt], qf[2 * k2 + 1][0], qf[2 * k2 + 1][2], qf[2 * k2 + 1][1], qf[2 * k2 + 1][3],
                         kf[2], kf[3]);
      }
    }
    if constexpr (cap) {
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) s[nt][e] = p.logit_cap * tanhf(s[nt][e] * cap_pre);
    }

    // ---- masks (extend_attention.py:166-184 prefix part, :245-264 extend part) ----
    const int j0 = n * kExtBN;
    const int j_end = j0 + kExtBN;
    const bool need_mask = j_end > Ltot || (p.window > 0 && j0 < Lp) ||
                           (use_mask && (!p.skip_prefix_mask || j_end > Lp)) ||
                           (!use_mask && p.causal && j_end - 1 - Lp > i_min_warp);
    if (need_mask) {
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int j = j0 + nt * 8 + 2 * jq + (e & 1);
          const int i = e < 2 ? i0 : i1;
          const bool rv = e < 2 ? row0 : row1;
          bool vis = rv && j < Ltot;
          if (j < Lp) {
            if (p.window > 0) vis = vis && (i <= j + p.window);
            if (use_mask && !p.skip_prefix_mask && vis) vis = mask[i * mask_row_len + j] != 0;
          } else if (use_mask) {
            if (vis) vis = mask[i * mask_row_len + j] != 0;
          } else if (p.causal) {
            vis = vis && (j - Lp <= i);
          }
          if (!vis) s[nt][e] = -CUDART_INF_F;
        }
      }
    }

    // ---- online softmax: rows gq (elements 0, 1) and gq + 8 (elements 2, 3) ----
    float mx0 = -CUDART_INF_F, mx1 = -CUDART_INF_F;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
      mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
    const float mr0 = mn0 == -CUDART_INF_F ? 0.f : mn0;
    const float mr1 = mn1 == -CUDART_INF_F ? 0.f : mn1;
    const float corr0 = fast_exp2((m0 - mr0) * c);
    const float corr1 = fast_exp2((m1 - mr1) * c);
    m0 = mn0;
    m1 = mn1;
    const float mc0 = mr0 * c, mc1 = mr1 * c;
    float sum0 = 0.f, sum1 = 0.f;
    uint32_t pa[NT][2];  // packed P: [nt][0] = row gq, [nt][1] = row gq + 8
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const float p0 = fast_exp2(fmaf(s[nt][0], c, -mc0));
      const float p1 = fast_exp2(fmaf(s[nt][1], c, -mc0));
      const float p2 = fast_exp2(fmaf(s[nt][2], c, -mc1));
      const float p3 = fast_exp2(fmaf(s[nt][3], c, -mc1));
      sum0 += p0 + p1;
      sum1 += p2 + p3;
      pa[nt][0] = pack2<HALF>(p0, p1);
      pa[nt][1] = pack2<HALF>(p2, p3);
    }
    l0 = fmaf(l0, corr0, sum0);
    l1 = fmaf(l1, corr1, sum1);
    if (__any_sync(0xffffffffu, corr0 != 1.f || corr1 != 1.f)) {
#pragma unroll
      for (int nn = 0; nn < NTILES; ++nn) {
        o[nn][0] *= corr0;
        o[nn][1] *= corr0;
        o[nn][2] *= corr1;
        o[nn][3] *= corr1;
      }
    }

    // ---- O += P.V ----  (rows past Ltot were zero-filled: p = 0 times v = 0)
#pragma unroll
    for (int kt = 0; kt < kExtBN / 16; ++kt) {
      const uint32_t va = sbase + v_lane + kt * 16 * RS;
#pragma unroll
      for (int n2 = 0; n2 < NTILES / 2; ++n2) {
        uint32_t vf[4];
        ldsm_x4_trans(va + n2 * 32, vf);
        mma16816<HALF>(o[2 * n2], pa[2 * kt][0], pa[2 * kt][1], pa[2 * kt + 1][0], pa[2 * kt + 1][1], vf[0], vf[1]);
        mma16816<HALF>(o[2 * n2 + 1], pa[2 * kt][0], pa[2 * kt][1], pa[2 * kt + 1][0], pa[2 * kt + 1][1], vf[2],
                       vf[3]);
      }
    }
  }
  cp_async_wait<0>();

  // ------------------------------ epilogue ------------------------------
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float inv0 = l0 > 0.f ? p.v_scale / l0 : 0.f;
  const float inv1 = l1 > 0.f ? p.v_scale / l1 : 0.f;
  if (row0) {
    QT* orow = reinterpret_cast<QT*>(p.o) + static_cast<int64_t>(q0 + i0) * p.o_st +
               static_cast<int64_t>(h * p.group + hd0) * p.o_sh;
#pragma unroll
    for (int nn = 0; nn < NTILES; ++nn)
      *reinterpret_cast<uint32_t*>(orow + nn * 8 + 2 * jq) = pack2<OUT_HALF>(o[nn][0] * inv0, o[nn][1] * inv0);
    if (p.lse && jq == 0)
      p.lse[static_cast<int64_t>(q0 + i0) * p.hq + h * p.group + hd0] =
          l0 > 0.f ? m0 * c * kLn2 + logf(l0) : -CUDART_INF_F;
  }
  if (row1) {
    QT* orow = reinterpret_cast<QT*>(p.o) + static_cast<int64_t>(q0 + i1) * p.o_st +
               static_cast<int64_t>(h * p.group + hd1) * p.o_sh;
#pragma unroll
    for (int nn = 0; nn < NTILES; ++nn)
      *reinterpret_cast<uint32_t*>(orow + nn * 8 + 2 * jq) = pack2<OUT_HALF>(o[nn][2] * inv1, o[nn][3] * inv1);
    if (p.lse && jq == 0)
      p.lse[static_cast<int64_t>(q0 + i1) * p.hq + h * p.group + hd1] =
          l1 > 0.f ? m1 * c * kLn2 + logf(l1) : -CUDART_INF_F;
  }
}

}  // namespace b200attn
