// K5, small-q variant: extend / verify attention for batches in which no request has more than 128
// (token, q head of the group) rows -- decode-like EXTEND batches and TARGET_VERIFY (8 draft tokens x 4 heads).
//
// Same contract, tile machinery and numerics as extend_tc_kernel.cuh (tcgen05 + TMEM, the reference's
// python/sglang/srt/layers/attention/triton_ops/extend_attention.py:41-303).  Two things differ:
//   * KV ping-pong: both 128-row accumulators take the SAME q rows and alternate over the KV tiles (sub-tile
//     v owns tiles n = v mod 2; stages are consumed in the order K0 K1 V0 K2 V1 ... K(N-1) V(N-2) V(N-1)),
//     each with its own running max / sum / O, merged in the epilogue: an in-CTA 2-way split-KV.
//   * persistent: grid = min(items, SMs) CTAs, item = (request, kv head), assigned round-robin.  Every role
//     loops over its CTA's items with running barrier phases, so the loaders run ahead into the next item
//     while the softmax warps still finish the current one: TMEM allocation, barrier setup, the launch gap
//     between CTAs, the prologue's dependent round trips and the pipeline fill are paid once per CTA, not
//     once per item (per item they cost ~20 k of ~125 k clk on the verify shape; tools/trace_tc_verify.py).
//     Q is double-buffered (the two 32 KB sub-tile buffers of the general kernel).
#pragma once

#include "extend_tc_kernel.cuh"

namespace b200attn {

// -DB200ATTN_TC_TRACE: CTA 0 appends (tag << 48 | clock64) events of four threads (softmax sub-tile 0 / 1,
// loader, MMA issuer) to the lse buffer, 2048 events per thread (tools/trace_tc_smallq.py decodes them).
#ifdef B200ATTN_TC_TRACE
#define SQ_TR(tag) do { if (trp && trn < 2047) trp[trn++] = (static_cast<long long>(tag) << 48) | (clock64() & 0xffffffffffffll); } while (0)
#else
#define SQ_TR(tag) do { } while (0)
#endif

constexpr int kSqBarBytes = 11264;  // mbarriers + row max / row sum / merge exchange + 4 slot-id buffers
constexpr int kSqSmemBytes = kTcQBytes + kTcSlots * kTcSlotBytes + kSqBarBytes + 1024;

// fp8 pools: launched with three more warps (21-23) that widen every landed stage to f16 in place, as in the one-shot
// kernel (extend_tc_kernel.cuh, WIDEN) -- the loaders only issue copies.
template <typename QT, int D, bool FP8, bool CAP>
__global__ void __launch_bounds__(FP8 ? kTcLaunchThreads : kTcThreads, 1) extend_tc_smallq_kernel(const ExtendParams p) {
  constexpr bool WIDEN = FP8 && kTcWidenWarps;
  static_assert(D == 128 || (D == 64 && !FP8), "head_dim 128, or 64 over a 16-bit pool");
  constexpr bool NARROW = FP8 || D == 64;
  constexpr int OH = D / 2;
  constexpr bool OUT_HALF = std::is_same<QT, __half>::value;
  constexpr bool HALF = OUT_HALF || FP8;
  constexpr uint32_t IDESC_QK = umma_idesc(kTcBM, kTcBN, HALF ? 0 : 1, 0);
  constexpr uint32_t IDESC_PV = umma_idesc(kTcBM, D, HALF ? 0 : 1, 1);

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t q_base = smem0;  // two Q buffers of kTcSlotBytes, item k uses buffer k & 1
  const uint32_t kv_base = smem0 + kTcQBytes;
  const uint32_t bars = kv_base + kTcSlots * kTcSlotBytes;
  const uint32_t bar_kv_full = bars;                  // [kTcSlots]
  const uint32_t bar_kv_empty = bars + 8 * kTcSlots;  // [kTcSlots]
  const uint32_t bar_q_full = bars + 16 * kTcSlots;   // [2]
  const uint32_t bar_q_empty = bar_q_full + 16;       // [2]
  const uint32_t bar_s_full = bar_q_empty + 16;       // [2]
  const uint32_t bar_p_full = bar_s_full + 16;        // [2]
  const uint32_t bar_o_full = bar_p_full + 16;
  const uint32_t bar_o_empty = bar_o_full + 8;
  const uint32_t bar_raw_full = bar_o_empty + 8;      // [kTcSlots] fp8: raw bytes landed (loaders -> widening warps)
  const uint32_t tmem_slot = bar_raw_full + 8 * kTcSlots;
  float* xch = reinterpret_cast<float*>(smem_raw + (bars + 1024 - smem_u32(smem_raw)));  // [parity][sub][half][row]
  float* lxch = xch + 2 * 2 * 2 * kTcBM;                                                   // [sub][half][row]
  int32_t* sid = reinterpret_cast<int32_t*>(lxch + 2 * 2 * kTcBM);                         // [4][kTcBN]
  float* mxch = reinterpret_cast<float*>(sid + 4 * kTcBN);                                 // [item parity][2][row]
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int hkv = p.hq / p.group;
  const int items = p.batch * hkv;
  const bool use_mask = p.custom_mask != nullptr;

  // One item's geometry, CTA-uniform; all loads are independent (one round trip).
  struct Geo {
    int b, h, q0, Le, Lp, Ltot, N;
    const int32_t* pidx;
  };
  auto geo = [&](int w) {
    Geo g;
    g.b = w / hkv;
    g.h = w - g.b * hkv;
    g.q0 = p.qo_indptr[g.b];
    const int q1 = p.qo_indptr[g.b + 1];
    const int kv0 = p.kv_indptr[g.b];
    const int kv1 = p.kv_indptr[g.b + 1];
    int64_t req_row = 0, pl_full = 0;
    if (!p.kv_indices) {
      req_row = p.idx64 ? static_cast<const int64_t*>(p.req_pool_indices)[g.b]
                        : static_cast<const int32_t*>(p.req_pool_indices)[g.b];
      pl_full = p.idx64 ? static_cast<const int64_t*>(p.prefix_lens)[g.b]
                        : static_cast<const int32_t*>(p.prefix_lens)[g.b];
    }
    g.Le = q1 - g.q0;
    g.Lp = kv1 - kv0;
    g.pidx = p.kv_indices ? p.kv_indices + kv0 : p.req_to_token + req_row * p.r2t_stride + (pl_full - g.Lp);
    g.Ltot = g.Lp + g.Le;  // one q tile: every new token is visible to some row
    g.N = g.Le > 0 ? (g.Ltot + kTcBN - 1) / kTcBN : 0;  // N == 0: the item is skipped by every role
    return g;
  };

  if (threadIdx.x == 0) {
    for (int s = 0; s < kTcSlots; ++s) {
      mbar_init(bar_kv_full + 8 * s, WIDEN ? 96 : 128);
      mbar_init(bar_kv_empty + 8 * s, 1);
      mbar_init(bar_raw_full + 8 * s, 128);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_q_full + 8 * i, 128);
      mbar_init(bar_q_empty + 8 * i, 1);
      mbar_init(bar_s_full + 8 * i, 1);
      mbar_init(bar_p_full + 8 * i, 256);
    }
    mbar_init(bar_o_full, 1);
    mbar_init(bar_o_empty, kTcSoftmaxWarps * 32);
    fence_barrier_init();
  }
  if (warp == kTcMmaWarp) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(tmem_slot));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot_ptr, 0);

#ifdef B200ATTN_TC_TRACE
  long long* trp = nullptr;
  int trn = 0;
  if (blockIdx.x == 0) {
    const int role = threadIdx.x == 0 ? 0 : threadIdx.x == 256 ? 1 : threadIdx.x == kTcLoaderWarp0 * 32 ? 2
                     : threadIdx.x == kTcMmaWarp * 32 ? 3 : -1;
    if (role >= 0) trp = reinterpret_cast<long long*>(p.lse) + role * 2048;
  }
#endif
  if (warp < kTcSoftmaxWarps) {
    // =============================== softmax warps ===============================
    // warp w: sub-tile w / 8 (= parity of the KV tiles it sees), column half (w / 4) & 1, TMEM lane quarter w & 3
    const int sub = warp >> 3;
    const int half = (warp >> 2) & 1;
    const int r = (warp & 3) * 32 + lane;
    const int itok = r >> p.gp_shift;
    const int hd = r & ((1 << p.gp_shift) - 1);
    const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t tS = tmem + lane_off + sub * kTcBN + half * 64;
    const uint32_t tO = tmem + lane_off + 2 * kTcBN + sub * kTcD + half * OH;
    constexpr bool cap = CAP;
    const float c = cap ? kLog2e : p.scale_log2;
    const float cap_pre = cap ? p.sm_scale / p.logit_cap : 0.f;
    float* xmine = xch + (sub * 2 + half) * kTcBM + r;
    float* xother = xch + (sub * 2 + (half ^ 1)) * kTcBM + r;
    int it_run = 0;  // S tiles this sub-tile has consumed since the kernel started (barrier phases)
    int k = 0;       // non-empty items this CTA has finished
    for (int w = blockIdx.x; w < items; w += gridDim.x) {
      const Geo g = geo(w);
      if (g.N == 0) continue;
      const int Lp = g.Lp, Le = g.Le, Ltot = g.Ltot, N = g.N;
      const bool rv = hd < p.group && itok < Le;
      // a warp whose 32 rows are all padding only keeps the barriers moving
      const bool dead = (warp & 3) * 32 >= (Le << p.gp_shift);
      const uint8_t* mrow = nullptr;
      if (use_mask) {
        const int64_t ms = p.mask64 ? static_cast<const int64_t*>(p.mask_indptr)[g.b]
                                    : static_cast<const int32_t*>(p.mask_indptr)[g.b];
        mrow = p.custom_mask + ms + (rv ? itok : 0) * (static_cast<int64_t>(Lp) + Le);
      }
      float m = -CUDART_INF_F, l = 0.f;
      int it = 0;
      SQ_TR(1);
      for (int n = sub; n < N; n += 2, ++it, ++it_run) {
        mbar_wait_wd(bar_s_full + 8 * sub, it_run & 1);
        tc_fence_after();
        SQ_TR(2);
        if (dead) {
          named_bar_sync(1 + sub, 256);
          mbar_arrive(bar_p_full + 8 * sub);
          continue;
        }
        const int j0 = n * kTcBN, j_end = j0 + kTcBN;
        const bool need_mask = j_end > Ltot || (p.window > 0 && j0 < Lp) ||
                               (use_mask && (!p.skip_prefix_mask || j_end > Lp)) ||
                               (!use_mask && p.causal && j_end - 1 - Lp > 0);
        auto tile_body = [&](auto masked_tag) {
          constexpr bool MASKED = decltype(masked_tag)::value;
          auto load_chunk = [&](int cc, float (&s)[32]) {
            tmem_ld32<0>(tS + cc, s);
            tmem_wait_ld();
            if constexpr (cap) {
#pragma unroll
              for (int j = 0; j < 32; ++j) s[j] = p.logit_cap * tanhf(s[j] * cap_pre);
            }
            if constexpr (MASKED) {
              const int hi = !rv ? 0 : ((p.causal && !use_mask) ? min(Ltot, Lp + itok + 1) : Ltot);
              const int lo = p.window > 0 ? min(max(itok - p.window, 0), Lp) : 0;
              const unsigned span = hi > lo ? static_cast<unsigned>(hi - lo) : 0u;
              const int mask_from = !use_mask ? 0x7fffffff : (p.skip_prefix_mask ? Lp : 0);
              const int jb = j0 + half * 64 + cc;
#pragma unroll
              for (int jj = 0; jj < 32; ++jj)
                s[jj] = static_cast<unsigned>(jb + jj - lo) < span ? s[jj] : -CUDART_INF_F;
              if (jb + 32 > mask_from) {
#pragma unroll
                for (int jj = 0; jj < 32; ++jj) {
                  const int j = jb + jj;
                  if (j >= mask_from && j < hi && mrow[j] == 0) s[jj] = -CUDART_INF_F;
                }
              }
            }
          };
          float mx0 = -CUDART_INF_F, mx1 = -CUDART_INF_F, mx2 = -CUDART_INF_F, mx3 = -CUDART_INF_F;
#pragma unroll
          for (int cc = 0; cc < 64; cc += 32) {
            float s[32];
            load_chunk(cc, s);
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              mx0 = fmaxf(mx0, s[j]);
              mx1 = fmaxf(mx1, s[j + 1]);
              mx2 = fmaxf(mx2, s[j + 2]);
              mx3 = fmaxf(mx3, s[j + 3]);
            }
          }
          float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
          const int par = (it_run & 1) * (2 * 2 * kTcBM);
          xmine[par] = mx;
          named_bar_sync(1 + sub, 256);
          mx = fmaxf(mx, xother[par]);
          const bool grow = mx > m && (m == -CUDART_INF_F || (mx - m) * c > kTcRescaleLog2);
          if (__any_sync(0xffffffffu, grow)) {
            float corr = 1.f;
            if (grow) {
              corr = fast_exp2((m - mx) * c);
              m = mx;
              l *= corr;
            }
            if (it > 0) {
#pragma unroll
              for (int cc = 0; cc < OH; cc += 32) {
                float o[32];
                tmem_ld32<0>(tO + cc, o);
                tmem_wait_ld();
#pragma unroll
                for (int j = 0; j < 32; ++j) o[j] *= corr;
                tmem_st32<0>(tO + cc, o);
              }
            }
          }
          const float mc = (m == -CUDART_INF_F ? 0.f : m) * c;
          float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
          for (int cc = 0; cc < 64; cc += 32) {
            float s[32];
            load_chunk(cc, s);
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              const float p0 = fast_exp2(fmaf(s[j], c, -mc));
              const float p1 = fast_exp2(fmaf(s[j + 1], c, -mc));
              sum0 += p0;
              sum1 += p1;
              pk[j >> 1] = pack2<HALF>(p0, p1);
            }
            tmem_st16(tS + (cc >> 1), pk);
          }
          l += sum0 + sum1;
          tmem_wait_st();
          tc_fence_before();
          mbar_arrive(bar_p_full + 8 * sub);
          SQ_TR(3);
        };
        if (need_mask) tile_body(std::true_type{});
        else tile_body(std::false_type{});
      }
      // ---- epilogue: row sum over both halves, sub-tile 1 hands (max, sum) to sub-tile 0, which merges the
      //      two accumulators and writes the rows ----
      SQ_TR(4);
      lxch[(sub * 2 + half) * kTcBM + r] = l;
      named_bar_sync(1 + sub, 256);
      l += lxch[(sub * 2 + (half ^ 1)) * kTcBM + r];
      float* mx = mxch + (k & 1) * (2 * kTcBM);
      if (sub == 1 && half == 0) {
        mx[r] = m == -CUDART_INF_F ? -CUDART_INF_F : m * c;
        mx[kTcBM + r] = l;
      }
      named_bar_sync(5, 512);
      SQ_TR(5);
      if (sub == 0 && !dead) {
        const float m0c = m == -CUDART_INF_F ? -CUDART_INF_F : m * c;
        const float m1c = mx[r], l1 = mx[kTcBM + r];
        const float mm = fmaxf(m0c, m1c);
        const float a0 = (l > 0.f) ? fast_exp2(m0c - mm) : 0.f;
        const float a1 = (l1 > 0.f) ? fast_exp2(m1c - mm) : 0.f;
        l = a0 * l + a1 * l1;
        const float mlog2 = mm == -CUDART_INF_F ? 0.f : mm;
        mbar_wait_wd(bar_o_full, k & 1);
        tc_fence_after();
        SQ_TR(6);
        const float inv = l > 0.f ? p.v_scale / l : 0.f;
        QT* orow = reinterpret_cast<QT*>(p.o) + static_cast<int64_t>(g.q0 + (rv ? itok : 0)) * p.o_st +
                   static_cast<int64_t>(g.h * p.group + (rv ? hd : 0)) * p.o_sh + half * OH;
#pragma unroll
        for (int cc = 0; cc < OH; cc += 32) {
          float o[32];
          tmem_ld32<0>(tO + cc, o);
          tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 32; ++j) o[j] = a0 > 0.f ? o[j] * a0 : 0.f;
#pragma unroll
          for (int c16 = 0; c16 < 32; c16 += 16) {
            float o1[16];
            tmem_ld16(tO + kTcD + cc + c16, o1);
            tmem_wait_ld();
#pragma unroll
            for (int j = 0; j < 16; ++j) o[c16 + j] = a1 > 0.f ? fmaf(o1[j], a1, o[c16 + j]) : o[c16 + j];
          }
          if (rv) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              uint4 v;
              v.x = pack2<OUT_HALF>(o[j] * inv, o[j + 1] * inv);
              v.y = pack2<OUT_HALF>(o[j + 2] * inv, o[j + 3] * inv);
              v.z = pack2<OUT_HALF>(o[j + 4] * inv, o[j + 5] * inv);
              v.w = pack2<OUT_HALF>(o[j + 6] * inv, o[j + 7] * inv);
              *reinterpret_cast<uint4*>(orow + cc + j) = v;
            }
          }
        }
#ifndef B200ATTN_TC_TRACE
        if (rv && half == 0 && p.lse)
          p.lse[static_cast<int64_t>(g.q0 + itok) * p.hq + g.h * p.group + hd] =
              l > 0.f ? mlog2 * kLn2 + logf(l) : -CUDART_INF_F;
#endif
      }
      // both accumulators have been read (or never will be): the MMA warp may start the next item's P.V
      tc_fence_before();
      mbar_arrive(bar_o_empty);
      SQ_TR(7);
      ++k;
    }
  } else if (warp < kTcMmaWarp) {
    // =============================== loaders (4 warps) ===============================
    const int tid = threadIdx.x - kTcLoaderWarp0 * 32;  // 0..127
    const int chunk = tid & 15, rbase = tid >> 4;       // 256-byte rows: row = rbase + 8 * i
    const uint32_t sw_col = static_cast<uint32_t>(chunk >> 3) * kTcSlab;
    const int chunk8 = tid & 7, rb8 = tid >> 3;         // 128-byte rows (fp8, head_dim 64): row = rb8 + 16 * i
    const int gmask = (1 << p.gp_shift) - 1;
    const int sid_pos = NARROW ? (tid & 15) * 8 + (tid >> 4) : (tid & 7) * 16 + (tid >> 3);
    // stage st of this thread is complete: (fp8: widen it in place,) hand it to the MMA warp; code bit 1 = the
    // stage is an item's first (its group also carries the item's Q), bit 0 = that item's Q buffer
    auto publish = [&](int st, uint32_t code) {
      if constexpr (WIDEN) {
        if (code & 2u) {
          fence_proxy_async();
          mbar_arrive(bar_q_full + 8 * (code & 1u));
        }
        mbar_arrive(bar_raw_full + 8 * (st % kTcSlots));
        return;
      }
      if constexpr (FP8) {
        const uint32_t slot0 = kv_base + (st % kTcSlots) * kTcSlotBytes;
        uint4 raw[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) raw[i] = lds128(slot0 + kTcSlab + (rb8 + 16 * i) * 128 + chunk8 * 16);
        __syncwarp();  // a row's eight raw chunks sit in eight adjacent lanes and its outputs only overwrite that row: no block barrier
        const uint32_t dslab = slot0 + static_cast<uint32_t>(chunk8 >> 2) * kTcSlab;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int row = rb8 + 16 * i;
          uint4 lo, hi;
          e4m3x16_to_f16(raw[i], lo, hi, p.e5m2 != 0);
          const uint32_t rowa = dslab + row * 128;
          sts128(rowa + ((((2 * chunk8) & 7) ^ (row & 7)) << 4), lo);
          sts128(rowa + ((((2 * chunk8 + 1) & 7) ^ (row & 7)) << 4), hi);
        }
      }
      fence_proxy_async();
      if (code & 2u) mbar_arrive(bar_q_full + 8 * (code & 1u));
      mbar_arrive(bar_kv_full + 8 * (st % kTcSlots));
    };
    // ---- slot-id stream: a cursor that runs three KV tiles ahead of the copies, across item boundaries ----
    int pf_w = blockIdx.x, pf_tile = 0;
    Geo pf;
    pf.N = 0;
    auto pf_seek = [&]() {  // first non-empty item at or after pf_w
      for (; pf_w < items; pf_w += gridDim.x) {
        pf = geo(pf_w);
        if (pf.N > 0) return;
      }
      pf.N = 0;
    };
    // Copies this thread's slot id of the cursor's tile into id buffer `buf` with a 4-byte cp.async (the id never
    // sits in a register: held in one across a stage it was spilled, and the spill store waited for the load --
    // ~1500 clk per tile with the whole loader stalled), then advances the cursor.  extend_loc is int64: the low
    // word is the slot id.
    auto pf_issue = [&](int buf) {
      if (pf.N > 0) {
        const int j = pf_tile * kTcBN + tid;
        int32_t* dst = sid + buf * kTcBN + sid_pos;
        if (j < pf.Ltot) {
          const void* src = j < pf.Lp ? static_cast<const void*>(pf.pidx + j)
                                      : static_cast<const void*>(p.extend_loc + (pf.q0 + j - pf.Lp));
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
        } else {
          *dst = -1;
        }
        if (++pf_tile == pf.N) {
          pf_tile = 0;
          pf_w += gridDim.x;
          pf_seek();
        }
      }
    };
    pf_seek();
    pf_issue(0);
    pf_issue(1);
    cp_async_commit();
    cp_async_wait<0>();  // once per kernel; later id copies ride in the K stage groups, two tiles ahead
    int gst = 0;        // stages issued so far (ring position)
    int vt0 = 0;        // virtual index of the current item's tile 0
    uint32_t hist = 0;  // publish codes of the stages issued so far, 2 bits each, newest lowest
    int k = 0;
    for (int w = blockIdx.x; w < items; w += gridDim.x) {
      const Geo g = geo(w);
      if (g.N == 0) continue;
      const int N = g.N, S = 2 * N;
      const size_t head_off = static_cast<size_t>(g.h) * D * (FP8 ? 1 : 2) + (NARROW ? chunk8 : chunk) * 16;
      const uint8_t* k_head = p.k_pool + head_off;
      const uint8_t* v_head = p.v_pool + head_off;
      for (int s = 0; s < S; ++s, ++gst) {
        if (gst >= kTcLag) {
          cp_async_wait<kTcLag - 1>();
          publish(gst - kTcLag, (hist >> (2 * (kTcLag - 1))) & 3u);
        }
        SQ_TR(10);
        const int slot = gst % kTcSlots;
        mbar_wait_wd(bar_kv_empty + 8 * slot, (static_cast<uint32_t>(gst / kTcSlots) & 1u) ^ 1u);
        SQ_TR(11);
        // stage -> (tensor, tile): K0 K1 V0 K2 V1 ... K(N-1) V(N-2) V(N-1)
        bool is_v = false;
        int tile = 0;
        if (s > 0) {
          is_v = s == S - 1 || !(s & 1);
          tile = s == S - 1 ? N - 1 : ((s & 1) ? (s + 1) >> 1 : (s >> 1) - 1);
        }
        if (!is_v) {
          named_bar_sync(3, 128);  // slot ids of this tile are visible
          // ids of virtual tile T + 2 replace those of T - 2, whose last reader -- V(T - 2) -- every loader
          // thread issued before it arrived at this barrier (also across item boundaries, where two V stages
          // are adjacent); the copy rides in this stage's group, complete for every thread (wait_group, then
          // the barrier) before K(T + 2), four stages on, reads the ids
          pf_issue((vt0 + tile + 2) & 3);
        }
        if (s == 0) {
          // the item's Q rows ride in the group of its first K stage
          const int qb = k & 1;
          mbar_wait_wd(bar_q_empty + 8 * qb, (static_cast<uint32_t>(k >> 1) & 1u) ^ 1u);
          SQ_TR(13);
          constexpr int QSTEP = D == 64 ? 16 : 8;
          const int qchunk = D == 64 ? chunk8 : chunk, qrb = D == 64 ? rb8 : rbase;
          for (int i = 0; i < kTcBM / QSTEP; ++i) {
            const int R = qrb + QSTEP * i;
            const int itok = R >> p.gp_shift, hd = R & gmask;
            const bool ok = hd < p.group && itok < g.Le;
            const QT* src = reinterpret_cast<const QT*>(p.q) +
                            (ok ? static_cast<int64_t>(g.q0 + itok) * p.q_st + static_cast<int64_t>(g.h * p.group + hd) * p.q_sh : 0) +
                            qchunk * 8;
            const uint32_t dst = q_base + qb * kTcSlotBytes + (D == 64 ? 0u : sw_col) + R * 128 + (((qchunk & 7) ^ (R & 7)) << 4);
            if constexpr (FP8 && !OUT_HALF) {
              uint4 v = make_uint4(0u, 0u, 0u, 0u);
              if (ok) v = *reinterpret_cast<const uint4*>(src);
              v.x = q_pair<QT, true>(v.x);
              v.y = q_pair<QT, true>(v.y);
              v.z = q_pair<QT, true>(v.z);
              v.w = q_pair<QT, true>(v.w);
              sts128(dst, v);
            } else {
              cp_async16_zfill(dst, src, ok);
            }
          }
        }
        const uint8_t* base = is_v ? v_head : k_head;
        const uint32_t sb = static_cast<uint32_t>(is_v ? p.v_slot_bytes : p.k_slot_bytes);  // < 2^32 (host check): one 32 x 32 -> 64 mad per row address
        const int32_t* sidt = sid + ((vt0 + tile) & 3) * kTcBN;
        if constexpr (NARROW) {
          const int4* idv = reinterpret_cast<const int4*>(sidt + rb8 * 8);
          const int4 i0 = idv[0], i1 = idv[1];
          const int32_t ids[8] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w};
          const uint32_t dst0 = kv_base + slot * kTcSlotBytes + (FP8 ? kTcSlab + chunk8 * 16 : 0);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int32_t sl = ids[i];
            const bool ok = sl >= 0;
            const int row = rb8 + 16 * i;
            cp_async16_zfill(dst0 + row * 128 + (FP8 ? 0 : ((chunk8 ^ (row & 7)) << 4)), base + static_cast<uint64_t>(static_cast<uint32_t>(ok ? sl : 0)) * sb, ok);
          }
        } else {
          // two batches of 8 rows: ids of a batch in registers before its copies (an asm volatile copy orders
          // the next shared load behind it), and half the addresses live at once -- this loop runs at the
          // kernel's 80-register budget
          const int4* idv = reinterpret_cast<const int4*>(sidt + rbase * 16);
          // row & 7 == rbase & 7 for every row of this thread: one swizzled column offset
          const uint32_t dst0 = kv_base + slot * kTcSlotBytes + sw_col + rbase * 128 + (((chunk & 7) ^ (rbase & 7)) << 4);
#pragma unroll
          for (int hb = 0; hb < 2; ++hb) {
            const int4 i0 = idv[2 * hb], i1 = idv[2 * hb + 1];
            const int32_t ids[8] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int32_t sl = ids[i];
              const bool ok = sl >= 0;
              cp_async16_zfill(dst0 + (hb * 8 + i) * (8 * 128), base + static_cast<uint64_t>(static_cast<uint32_t>(ok ? sl : 0)) * sb, ok);
            }
          }
        }
        cp_async_commit();
        SQ_TR(12);
        hist = (hist << 2) | (s == 0 ? 2u | static_cast<uint32_t>(k & 1) : 0u);
      }
      vt0 += N;
      ++k;
    }
    // drain: the last min(gst, kTcLag) stages
    cp_async_wait<0>();
    for (int st = (gst > kTcLag ? gst - kTcLag : 0); st < gst; ++st) publish(st, (hist >> (2 * (gst - 1 - st))) & 3u);
  } else if (warp == kTcMmaWarp) {
    // =============================== MMA issuer (warp 20) ===============================
    const uint64_t dq = umma_desc_sw128(q_base, 16, 1024);
    const uint64_t dk = umma_desc_sw128(kv_base, 16, 1024);
    const uint64_t dv = umma_desc_sw128(kv_base, kTcSlab, 1024);
    auto issue_qk = [&](int sub, int slot, int qb) {
      const uint64_t a0 = dq + static_cast<uint32_t>(qb * (kTcSlotBytes >> 4));
      const uint64_t b0 = dk + static_cast<uint32_t>(slot * (kTcSlotBytes >> 4));
      const uint32_t d = tmem + sub * kTcBN;
#pragma unroll
      for (int kk = 0; kk < D / 16; ++kk) {
        const uint32_t off = ((kk >> 2) * kTcSlab + (kk & 3) * 32) >> 4;
        umma_ss(d, a0 + off, b0 + off, IDESC_QK, kk > 0);
      }
    };
    auto issue_pv = [&](int sub, int slot, bool acc) {
      const uint64_t v0 = dv + static_cast<uint32_t>(slot * (kTcSlotBytes >> 4));
      const uint32_t d = tmem + 2 * kTcBN + sub * kTcD, a = tmem + sub * kTcBN;
      umma_ts(d, a, v0, IDESC_PV, acc ? 1u : 0u);
#pragma unroll
      for (int kk = 1; kk < kTcBN / 16; ++kk)
        umma_ts(d, a + kk * 8 + (kk >= 4 ? 32 : 0), v0 + kk * (2048 >> 4), IDESC_PV, 1u);
    };
    auto full_wait = [&](int st) {
      mbar_wait_wd(bar_kv_full + 8 * (st % kTcSlots), static_cast<uint32_t>(st / kTcSlots) & 1u);
    };
    int gst = 0;             // ring position of the current item's stage 0
    uint32_t pph = 0;        // bit v: parity of the next p_full phase of sub-tile v
    int k = 0;
    for (int w = blockIdx.x; w < items; w += gridDim.x) {
      const Geo g = geo(w);
      if (g.N == 0) continue;
      const int N = g.N;
      const int qb = k & 1;
      SQ_TR(20);
      // one chain per item: S0(0) S1(1) | PV_v(j) S_v(j + 2) with v = j & 1 -- every stage feeds one sub-tile
      mbar_wait_wd(bar_q_full + 8 * qb, static_cast<uint32_t>(k >> 1) & 1u);
      full_wait(gst);
      tc_fence_after();
      SQ_TR(21);
      issue_qk(0, gst % kTcSlots, qb);
      umma_commit(bar_s_full);
      umma_commit(bar_kv_empty + 8 * (gst % kTcSlots));
      if (N >= 2) {
        full_wait(gst + 1);
        tc_fence_after();
        issue_qk(1, (gst + 1) % kTcSlots, qb);
        umma_commit(bar_s_full + 8);
        umma_commit(bar_kv_empty + 8 * ((gst + 1) % kTcSlots));
      }
      if (N <= 2) umma_commit(bar_q_empty + 8 * qb);  // the item's last Q.K^T has been issued
      for (int j = 0; j < N; ++j) {
        const int v = j & 1;
        const int vst = gst + (j < N - 1 ? 2 * j + 2 : 2 * N - 1);
        full_wait(vst);
        SQ_TR(22);
        mbar_wait_wd(bar_p_full + 8 * v, (pph >> v) & 1u);
        pph ^= 1u << v;
        // the first P.V of an item overwrites an accumulator: the previous item's epilogue must have read both
        if (j == 0 && k > 0) mbar_wait_wd(bar_o_empty, static_cast<uint32_t>(k - 1) & 1u);
        tc_fence_after();
        SQ_TR(23);
        issue_pv(v, vst % kTcSlots, j >= 2);
        umma_commit(bar_kv_empty + 8 * (vst % kTcSlots));
        if (j + 2 < N) {
          const int kst = gst + 2 * j + 3;
          full_wait(kst);
          tc_fence_after();
          issue_qk(v, kst % kTcSlots, qb);
          umma_commit(bar_s_full + 8 * v);
          umma_commit(bar_kv_empty + 8 * (kst % kTcSlots));
          if (j + 3 == N) umma_commit(bar_q_empty + 8 * qb);
        }
      }
      umma_commit(bar_o_full);
      SQ_TR(25);
      gst += 2 * N;
      ++k;
    }
    tc_fence_before();
  } else if constexpr (WIDEN) {
    // =============================== fp8 widening warps (21-23) ===============================
    // (same mapping as extend_tc_kernel's: thread = (row group, 16-byte chunk), a row's eight chunks in eight
    //  adjacent lanes, __syncwarp between a row's loads and its stores)
    const int wt = threadIdx.x - (kTcMmaWarp + 1) * 32;  // 0..95
    const int chunk8 = wt & 7, rb = wt >> 3;
    int gst = 0;
    for (int w = blockIdx.x; w < items; w += gridDim.x) {
      const Geo g = geo(w);
      for (int s2 = 0; s2 < 2 * g.N; ++s2, ++gst) {
        const int slot = gst % kTcSlots;
        mbar_wait_wd(bar_raw_full + 8 * slot, static_cast<uint32_t>(gst / kTcSlots) & 1u);
        const uint32_t slot0 = kv_base + slot * kTcSlotBytes;
        const uint32_t dslab = slot0 + static_cast<uint32_t>(chunk8 >> 2) * kTcSlab;
#pragma unroll
        for (int i0 = 0; i0 < 12; i0 += 4) {
          uint4 raw[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int row = rb + 12 * (i0 + i);
            raw[i] = row < kTcBN ? lds128(slot0 + kTcSlab + row * 128 + chunk8 * 16) : make_uint4(0u, 0u, 0u, 0u);
          }
          __syncwarp();
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int row = rb + 12 * (i0 + i);
            if (row < kTcBN) {
              uint4 lo, hi;
              e4m3x16_to_f16(raw[i], lo, hi, p.e5m2 != 0);
              const uint32_t rowa = dslab + row * 128;
              sts128(rowa + ((((2 * chunk8) & 7) ^ (row & 7)) << 4), lo);
              sts128(rowa + ((((2 * chunk8 + 1) & 7) ^ (row & 7)) << 4), hi);
            }
          }
        }
        fence_proxy_async();
        mbar_arrive(bar_kv_full + 8 * slot);
      }
    }
  }
  __syncthreads();
  if (warp == kTcMmaWarp) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
  }
}

}  // namespace b200attn
