// K2: split-KV GQA decode, persistent, warp-specialised (sm_100a).
//
// Replaces the reference's Triton stage-1 kernels
// (python/sglang/srt/layers/attention/triton_ops/decode_attention.py:44-169
// MHA, :240-401 grouped) and computes the math of the torch_native oracle
// (layers/attention/torch_native_backend.py:112-180).
//
// Layout of one CTA (384 threads, 1 CTA / SM, grid = SM count):
//   warps 8..11 producers: read slot indices (req_to_token row or kv_indices)
//               one chunk ahead, and issue one bulk async copy (TMA engine,
//               SASS UBLKCP) per token row for K and for V into a ring of smem
//               stages; completion is counted in bytes on a "full" mbarrier.
//   warps 0..7  consumers: warp (h, slice) owns kv head h of the CTA's head
//               group and 16 tokens of every chunk.  All `group` q heads that
//               share the kv head ride in the M dimension of one
//               mma.sync.m16n8k16 tile, so every K/V byte is read from HBM and
//               from smem once.  Online softmax in registers (quad shuffles),
//               P stays in registers as the A operand of P.V.
// The chunk sequence of the whole batch is cut into equal per-CTA ranges by
// plan_kernel; a CTA therefore streams its share of the KV bytes with a single
// pipeline fill, whatever the mix of request lengths.  A range that ends
// inside a request leaves a normalised partial (o, lse) in the workspace for
// merge_kernel; a request that fits one range is written straight to `o`.
#pragma once

#include <math_constants.h>

#include <type_traits>

#include "common.cuh"
#include "merge.cuh"

namespace b200attn {

struct DecodeParams {
  const void* q;
  int64_t q_sb, q_sh;            // elements
  const uint8_t* k_pool;
  const uint8_t* v_pool;
  int64_t k_slot_bytes, v_slot_bytes;
  void* o;
  int64_t o_sb, o_sh;            // elements
  float* lse;                    // optional [batch, hq]
  const void* plan;
  const int32_t* indices;
  float* ws_o;
  float* ws_lse;
  int32_t batch, hq;
  int32_t hc, ts, ng, group;
  int32_t row_bytes, row_stride, stage_bytes, stages, chunk_tokens;
  int32_t cp_async;              // 1: producers use 16-byte cp.async (small rows), 0: one bulk copy per row
  int32_t fuse_merge;            // 1: the CTA that writes a request's last partial merges it (no K3 launch)
  int32_t unit_shift;            // log2(row_bytes / 16)
  int32_t swizzle;               // 1: stage rows unpadded, 16-byte unit u of row r at u ^ (r & 7) (cp.async path)
  int32_t xch_off;               // smem offset of the slice-exchange area (0 = none): whole requests merge in the CTA
  float scale_log2;              // sm_scale * k_scale * log2(e)
  float sm_scale_k;              // sm_scale * k_scale
  float logit_cap;
  float v_scale;
#ifdef B200ATTN_DEC_TRACE
  unsigned long long* trace;     // [grid][kTraceSlots] clock stamps (tools/trace_decode.py); nullptr = off
#endif
};

// -DB200ATTN_DEC_TRACE (make trace -> libb200attn_trace.so): per-CTA timeline of one launch.  Slot 0 / 15 hold
// %globaltimer (ns, comparable across SMs) at entry / exit, the others clock64 of the CTA's own SM.
#ifdef B200ATTN_DEC_TRACE
constexpr int kTraceSlots = 16;
enum {
  TR_GT_ENTRY = 0, TR_ENTRY, TR_BARS_READY, TR_P_WALK, TR_P_IDS, TR_P_FIRST_ISSUE, TR_P_LAST_ISSUE, TR_C_Q,
  TR_C_FIRST_FULL, TR_C_FIRST_DONE, TR_C_LAST_DONE, TR_C_EPILOGUE, TR_C_MERGED, TR_C_EXIT, TR_COUNTS, TR_GT_EXIT
};
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define B200_TR(slot) do { if (p.trace) p.trace[blockIdx.x * kTraceSlots + (slot)] = clock64(); } while (0)
#define B200_TR_ONCE(slot, flag) do { if (p.trace && !(flag)) { (flag) = true; p.trace[blockIdx.x * kTraceSlots + (slot)] = clock64(); } } while (0)
#define B200_TR_GT(slot) do { if (p.trace) p.trace[blockIdx.x * kTraceSlots + (slot)] = globaltimer_ns(); } while (0)
#else
#define B200_TR(slot) do { } while (0)
#define B200_TR_ONCE(slot, flag) do { } while (0)
#define B200_TR_GT(slot) do { } while (0)
#endif

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

template <bool HALF>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  if constexpr (HALF) {
    __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
  } else {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
  }
}

__device__ __forceinline__ float q_to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }
__device__ __forceinline__ float q_to_f32(__half x) { return __half2float(x); }
__device__ __forceinline__ void store_q(__nv_bfloat16* p, float x) { *p = __float2bfloat16_rn(x); }
__device__ __forceinline__ void store_q(__half* p, float x) { *p = __float2half_rn(x); }

// S-tile MMA with the Q fragment: q = {a0, a2} (+ {a1, a3} for 16 valid rows).
template <bool HALF, bool ROWS16, int QR>
__device__ __forceinline__ void mma_q(float (&d)[4], const uint32_t (&q)[QR], uint32_t b0,
                                      uint32_t b1) {
  if constexpr (ROWS16) {
    mma16816<HALF>(d, q[0], q[2], q[1], q[3], b0, b1);
  } else {
    mma16816<HALF>(d, q[0], 0u, q[1], 0u, b0, b1);
  }
}

// Walks the pieces (virtual request x local chunk interval) of one range.  Order: the range's LAST
// piece first, then the others in ascending order.  The first and last piece of a range are the ones
// shared with the neighbouring ranges (a request cut by the range boundary); visiting both early
// means every cut request has all its partials -- and its merge -- done long before the kernel ends,
// and each CTA finishes on whole requests whose epilogue is a plain store.
__device__ __forceinline__ RangeDesc load_desc(const RangeDesc* d) {
  RangeDesc r;
  const int4* s = reinterpret_cast<const int4*>(d);
  int4* t = reinterpret_cast<int4*>(&r);
  t[0] = s[0];
  t[1] = s[1];
  t[2] = s[2];
  return r;
}

struct PieceWalk {
  RangeDesc d;
  int ng;
  int i = 0;
  int left;  // chunks of the range not yet handed out
  __device__ __forceinline__ PieceWalk(const RangeDesc& d_, int ng_) : d(d_), ng(ng_) { left = d.g1 - d.g0; }
  // S / row_off: the request's length and the element offset of its index row
  __device__ __forceinline__ bool next(const PlanView& pv, int& b, int& hg, int& lc0, int& lc1,
                                       int& nchunks, int& vreq, int& S, int64_t& row_off) {
    if (i == 0) {  // the range's last virtual request, straight from the descriptor
      i = 1;
      b = d.b;
      hg = d.hg;
      lc0 = d.lc0;
      lc1 = d.lc1;
      nchunks = d.nchunks;
      vreq = d.v_last;
      S = d.seq_len;
      row_off = d.row_off;
      return true;
    }
    const int m = d.v_last - d.v_first + 1;
    while (i < m) {
      const int v = d.v_first + i - 1;
      ++i;
      b = v / ng;
      hg = v - b * ng;
      const int cp0 = pv.chunk_prefix[b];
      const int n = pv.chunk_prefix[b + 1] - cp0;
      const int vstart = ng * cp0 + hg * n;
      const int lo = vstart > d.g0 ? vstart : d.g0;
      const int hi = vstart + n < d.g1 ? vstart + n : d.g1;
      if (lo >= hi) continue;  // empty request
      lc0 = lo - vstart;
      lc1 = hi - vstart;
      nchunks = n;
      vreq = v;
      S = pv.seq_len[b];
      row_off = pv.row_off[b];
      return true;
    }
    return false;
  }
};

template <typename QT, bool FP8, int D, bool ROWS16>
__global__ void __launch_bounds__(kThreads, 1) decode_kernel(const DecodeParams p) {
  constexpr bool HALF = FP8 || std::is_same<QT, __half>::value;
  constexpr int KSTEPS = D / 16;       // k16 steps of Q.K^T
  constexpr int NTILES = D / 8;        // n8 tiles of P.V
  constexpr int QR = ROWS16 ? 4 : 2;   // A-fragment registers per k-step actually loaded
  constexpr int EB = FP8 ? 1 : 2;      // KV element bytes

  extern __shared__ __align__(128) uint8_t smem[];
  const int cta = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar_full = smem_u32(smem);
  const uint32_t bar_empty = bar_full + kMaxStages * 8;
  const uint32_t stage0 = bar_full + kBarrierBytes;
  volatile int32_t* rq = reinterpret_cast<volatile int32_t*>(smem + kRangeRingOff);   // ring of range ids
  volatile int32_t* merge_flag = reinterpret_cast<volatile int32_t*>(smem + kMergeFlagOff);
#ifdef B200ATTN_DEC_TRACE
  bool tr_a = false, tr_b = false, tr_c = false;
  unsigned long long tr_chunks = 0, tr_pieces = 0;
  if (threadIdx.x == 0) {
    B200_TR_GT(TR_GT_ENTRY);
    B200_TR(TR_ENTRY);
    unsigned smid;
    asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
    if (p.trace) p.trace[blockIdx.x * kTraceSlots + TR_P_IDS] = smid;   // slot reused: which SM ran this CTA
  }
#endif
  // Everything that needs no global data first: under programmatic dependent launch this part (CTA launch,
  // barrier init) overlaps the tail of the kernel in front (kv_store / RoPE + store), which has released its
  // dependents at its start.
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(bar_full + s * 8, p.cp_async ? kProducerWarps * 32 : kProducerWarps);
      mbar_init(bar_empty + s * 8, kConsumerWarps);
    }
    fence_barrier_init();
  }
  pdl_wait();  // plan, q and the pool rows of this step are written by kernels in front of this one

  const PlanView pv = plan_view(const_cast<void*>(p.plan), p.batch);
  // the header's first 56 bytes (everything but the ticket) in one batch of loads, issued with the descriptor's:
  // a field read later (csr_mode in the producers, n_empty in the consumers) is then no further round trip
  PlanHeader H;
  {
    const int4* hs = reinterpret_cast<const int4*>(pv.hdr);
    int4* hd = reinterpret_cast<int4*>(&H);
    hd[0] = hs[0];
    hd[1] = hs[1];
    hd[2] = hs[2];
    const int2 h3 = *reinterpret_cast<const int2*>(hs + 3);
    H.total_tokens = h3.x;
    H.n_empty = h3.y;
  }
  // this CTA's static range, fetched together with the header (valid when cta < n_static)
  const RangeDesc d0 = load_desc(pv.range_desc + cta);
  // ... and, for the producers, the slot ids of the range's first chunk (plan_kernel copied them next to the
  // descriptors): issued with the two loads above, so the first copies wait on one round trip, not three.
  int32_t pre_ids[2] = {0, 0};
  if (warp >= kConsumerWarps && cta < kMaxCtas) {
    const int pw_ = warp - kConsumerWarps;
    const int T_ = p.chunk_tokens;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int i = p.cp_async ? (pw_ >> 1) * (T_ >> 1) + lane + 32 * r : 2 * (lane + 32 * r) + (pw_ >> 1);
      const bool mine = p.cp_async ? lane + 32 * r < (T_ >> 1) : i < T_;
      if (mine) pre_ids[r] = __ldg(pv.first_ids + cta * kFirstIds + i);
    }
  }
  // the rate this CTA recorded in earlier launches (updated at exit; fetched here so that the exit is a plain store)
  CtaStat st_prev = {0.f, 0u};
  if (threadIdx.x == 0 && cta < kMaxCtas) st_prev = pv.cta_stat[cta];
  if (cta >= H.grid) return;
  // A plan built for another batch, head geometry or chunk size would send the walk through garbage ranges:
  // do nothing instead (o is left untouched; b200attn_decode documents that plan and call must agree).
  if (H.magic != kPlanMagic || H.batch != p.batch || H.ng != p.ng || H.chunk_tokens != p.chunk_tokens) return;
  const int R = H.n_ranges;
  const int n_dyn = H.n_dynamic, n_grid = H.grid, n_static = H.n_static;

  // Dynamic ranges are claimed through a ticket counter in the plan buffer.  Every launch draws
  // exactly n_dynamic + grid tickets (each CTA stops at its first ticket past the dynamic ranges),
  // so the counter never needs a reset between the layers that share one plan.
  // Without dynamic ranges (the default plan) there is nothing to claim: no atomic.  (148 CTAs incrementing one
  // address took ~1 us, and the K producer warp sat on the result before its first copy: tools/trace_decode.py.)
  auto take_ticket = [&]() -> int {
    if (n_dyn == 0) return R;
    const unsigned long long t = atomicAdd(&pv.hdr->ticket, 1ull);
    const unsigned k = static_cast<unsigned>(t % static_cast<unsigned>(n_dyn + n_grid));
    return k < static_cast<unsigned>(n_dyn) ? n_static + static_cast<int>(k) : R;
  };

  if (threadIdx.x == 0) rq[0] = cta < n_static ? cta : take_ticket();
  __syncthreads();
  if (threadIdx.x == 0) B200_TR(TR_BARS_READY);

  const int T = p.chunk_tokens;
  const uint32_t v_off = T * p.row_stride;  // V rows follow the K rows of a stage
  int b, hg, lc0, lc1, nchunks, vreq, S;
  int64_t row_off;
  int stage = 0;
  uint32_t phase = 0;
  int rn = 0;          // ranges done by this CTA
  int rg = rq[0];      // current range (R = none)

  if (warp >= kConsumerWarps) {
    // ------------------------------ producers ------------------------------
    // The producers (warp group 2) need few registers; they hand theirs to the two consumer warp groups, whose
    // accumulators + fragments + epilogue state did not fit the 168 registers of a 384-thread CTA without spills.
    asm volatile("setmaxnreg.dec.sync.aligned.u32 72;");
    // warp pw: tensor = pw & 1 (K / V), tokens of parity pw >> 1.  Lane l, round r owns chunk
    // token 2 * (l + 32 r) + parity.
    const int pw = warp - kConsumerWarps;
    const int which = pw & 1;
    const int par = pw >> 1;
    const uint8_t* pool = which ? p.v_pool : p.k_pool;
    const int64_t slot_bytes = which ? p.v_slot_bytes : p.k_slot_bytes;
    if (p.cp_async) {
      // Small rows (<= 512 B): a bulk copy is issued per row at ~20 clk / row / SM whatever its
      // size (4.2 TB/s at 256 B, 2.0 TB/s at 128 B, 6.2 at 512 B, profiles/ r01 tp-sim), so rows are
      // moved as 16-byte cp.async units instead: one warp instruction carries 512 B (6.4 TB/s at 512 B,
      // 6.1-6.5 at 256 B, 4.4 at 128 B = the memory system's own random-row ceiling, tools/gather_roofline.cu).
      // From 1 KB up the bulk copy is used (its issue costs the SM nothing).  A variant of this loop that
      // also covered wider rows (several instructions per row, a run-time inner loop) was measured and dropped:
      // no gain at 2 KB (6.62 against 6.68 TB/s), and the inner loop cost the 256- / 512-byte shapes 25 %.
      // Warp (tensor, half)
      // owns tokens [half * T/2, (half + 1) * T/2) of every chunk; lane l holds the slot ids of
      // tokens l and l + 32 of that range (one chunk ahead) and hands them out by shuffle.
      // The issue loop is the producers' whole job and was this path's bound (ncu r01 tp8: ~31 instructions
      // per LDGSTS, producer warps never idle, consumers starved): everything lane-invariant is hoisted, a copy
      // costs one shuffle (slot id), one 32 x 32 -> 64 bit multiply-add (address), the copy, two adds.
      const int ush = p.unit_shift;  // <= 5: this path serves rows of at most 512 bytes (one warp instruction)
      const int umask = (1 << ush) - 1;
      const int Tw = T >> 1;
      const int tbase = par * Tw;
      const int rsh = 5 - ush;                 // log2(rows one warp instruction covers)
      const int rpi = 1 << rsh;
      const int r_in = lane >> ush;            // this lane's row inside an instruction
      const int col = (lane & umask) << 4;     // this lane's 16-byte unit inside the row
      const uint32_t sb32 = static_cast<uint32_t>(slot_bytes);  // < 2^32: checked by the host
      const uint32_t dstep = static_cast<uint32_t>(rpi) * p.row_stride;
      // destination unit of this lane in instruction i: i & 7 decides (rows advance by rpi = 1, 2 or 4 per
      // instruction; every issue run starts on a row that is a multiple of 8 and is a multiple of 8 long,
      // except over 64-byte rows, which are never swizzled)
      uint32_t scol[8];
#pragma unroll
      for (int j = 0; j < 8; ++j)
        scol[j] = p.swizzle ? static_cast<uint32_t>((lane & umask) ^ ((r_in + j * rpi) & 7)) << 4 : static_cast<uint32_t>(col);
      const int cnt0 = (Tw < 32 ? Tw : 32) >> rsh;  // instructions fed by the ids of tokens [0, 32) / [32, 64)
      const int cnt1 = (Tw > 32 ? Tw - 32 : 0) >> rsh;
      while (rg < R) {
        // Claim the range after this one now; the ticket is only looked at when this range's last
        // chunk is issued, so the atomic's round trip (microseconds under load) hides behind the copies.
        int next_rg = R;
        if (pw == 0 && lane == 0) next_rg = take_ticket();
        PieceWalk walk(rn == 0 && rg == cta ? d0 : load_desc(pv.range_desc + rg), p.ng);
        const bool use_pre = rn == 0 && rg == cta && cta < kMaxCtas && !H.csr_mode;
      while (walk.next(pv, b, hg, lc0, lc1, nchunks, vreq, S, row_off)) {
        if (pw == 0 && lane == 0) B200_TR_ONCE(TR_P_WALK, tr_a);
        const int32_t* idx = p.indices + row_off;
        const uint8_t* base = pool + static_cast<int64_t>(hg) * p.row_bytes;
        int32_t nxt[2];
        if (use_pre && walk.i == 1) {   // the static range's first piece: ids came with the descriptor
          nxt[0] = pre_ids[0];
          nxt[1] = pre_ids[1];
        } else {
#pragma unroll
          for (int r = 0; r < 2; ++r) {
            const int i = lane + 32 * r;
            const int t = lc0 * T + tbase + i;
            nxt[r] = (i < Tw && t < S) ? idx[t] : 0;
          }
        }
        for (int lc = lc0; lc < lc1; ++lc) {
          const int32_t cur0 = nxt[0], cur1 = nxt[1];
          if (lc + 1 < lc1) {
#pragma unroll
            for (int r = 0; r < 2; ++r) {
              const int i = lane + 32 * r;
              const int t = (lc + 1) * T + tbase + i;
              nxt[r] = (i < Tw && t < S) ? idx[t] : 0;
            }
          }
          const int nvalid = S - lc * T - tbase;  // valid tokens of this warp's range (may be <= 0)
          mbar_wait(bar_empty + stage * 8, phase ^ 1);
          const uint32_t full = bar_full + stage * 8;
          uint32_t dstp = stage0 + stage * p.stage_bytes + (which ? v_off : 0u) + (tbase + r_in) * p.row_stride;
          const uint8_t* bl = base + col;
          auto issue = [&](auto pred_tag, int32_t cur, int cnt, int tl) {
            constexpr bool PRED = decltype(pred_tag)::value;
            int srcl = r_in;
            for (int i = 0; i < cnt; i += 8) {  // cnt is a multiple of 8, or 4 (64-byte rows)
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                if (j == 4 && cnt == 4) break;
                const uint32_t sl = static_cast<uint32_t>(__shfl_sync(0xffffffffu, cur, srcl));
                const uint8_t* src = bl + static_cast<uint64_t>(sl) * sb32;
                if (!PRED || tl < nvalid) cp_async16(dstp + scol[j], src);
                srcl += rpi;
                dstp += dstep;
                if constexpr (PRED) tl += rpi;
              }
            }
          };
          if (nvalid >= Tw) {  // every chunk but a request's last
            issue(std::false_type{}, cur0, cnt0, 0);
            issue(std::false_type{}, cur1, cnt1, 0);
          } else {
            issue(std::true_type{}, cur0, cnt0, r_in);
            issue(std::true_type{}, cur1, cnt1, 32 + r_in);
          }
          if (pw == 0 && lane == 0 && --walk.left == 0) {
            rq[(rn + 1) & (kRangeRing - 1)] = next_rg;  // published before this range's last arrival
            __threadfence_block();
          }
          cp_async_mbar_arrive_noinc(full);
          if (pw == 0 && lane == 0) { B200_TR_ONCE(TR_P_FIRST_ISSUE, tr_b); B200_TR(TR_P_LAST_ISSUE); }
          if (++stage == p.stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
        named_bar_sync(1, kProducerWarps * 32);  // every producer warp sees the ticket written above
        ++rn;
        rg = rq[rn & (kRangeRing - 1)];
      }
      return;
    }
    while (rg < R) {
      // Claim the range after this one now; the ticket is only looked at when this range's last
      // chunk is issued, so the atomic's round trip (microseconds under load) hides behind the copies.
      int next_rg = R;
      if (pw == 0 && lane == 0) next_rg = take_ticket();
      PieceWalk walk(rn == 0 && rg == cta ? d0 : load_desc(pv.range_desc + rg), p.ng);
      const bool use_pre = rn == 0 && rg == cta && cta < kMaxCtas && !H.csr_mode;
    while (walk.next(pv, b, hg, lc0, lc1, nchunks, vreq, S, row_off)) {
      if (pw == 0 && lane == 0) B200_TR_ONCE(TR_P_WALK, tr_a);
      const int32_t* idx = p.indices + row_off;
      const uint8_t* base = pool + static_cast<int64_t>(hg) * p.row_bytes;
      int32_t nxt[2];
      if (use_pre && walk.i == 1) {   // the static range's first piece: ids came with the descriptor
        nxt[0] = pre_ids[0];
        nxt[1] = pre_ids[1];
      } else {
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          const int i = 2 * (lane + 32 * r) + par;
          const int t = lc0 * T + i;
          nxt[r] = (i < T && t < S) ? idx[t] : 0;
        }
      }
      for (int lc = lc0; lc < lc1; ++lc) {
        int32_t cur[2] = {nxt[0], nxt[1]};
        if (lc + 1 < lc1) {  // index prefetch, one chunk ahead of the copies
#pragma unroll
          for (int r = 0; r < 2; ++r) {
            const int i = 2 * (lane + 32 * r) + par;
            const int t = (lc + 1) * T + i;
            nxt[r] = (i < T && t < S) ? idx[t] : 0;
          }
        }
        const int t0 = lc * T;
        const int nvalid = S - t0 < T ? S - t0 : T;
        const int mine = (nvalid - par + 1) >> 1;  // valid tokens of this warp's parity
        mbar_wait(bar_empty + stage * 8, phase ^ 1);
        const uint32_t full = bar_full + stage * 8;
        if (pw == 0 && lane == 0 && --walk.left == 0) {
          rq[(rn + 1) & (kRangeRing - 1)] = next_rg;  // published before this range's last arrival
          __threadfence_block();
        }
        if (lane == 0) mbar_arrive_expect_tx(full, static_cast<uint32_t>(mine) * p.row_bytes);
        __syncwarp();
        const uint32_t dst = stage0 + stage * p.stage_bytes + (which ? v_off : 0u);
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          const int i = 2 * (lane + 32 * r) + par;
          if (i < nvalid)
            bulk_g2s(dst + i * p.row_stride, base + static_cast<int64_t>(cur[r]) * slot_bytes,
                     p.row_bytes, full);
        }
        if (pw == 0 && lane == 0) { B200_TR_ONCE(TR_P_FIRST_ISSUE, tr_b); B200_TR(TR_P_LAST_ISSUE); }
        if (++stage == p.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
      named_bar_sync(1, kProducerWarps * 32);  // every producer warp sees the ticket written above
      ++rn;
      rg = rq[rn & (kRangeRing - 1)];
    }
    return;
  }

  // -------------------------------- consumers --------------------------------
  asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");   // 128 x 72 + 256 x 216 = 64512 = 384 x 168
  const int h = warp % p.hc;
  const int tsi = warp / p.hc;
  const int gq = lane >> 2, jq = lane & 3;     // MMA fragment coordinates
  const int mat = lane >> 3, mr = lane & 7;    // ldmatrix: matrix id, row
  const bool cap = p.logit_cap > 0.f;
  const float c = cap ? kLog2e : p.scale_log2;
  const float cap_pre = cap ? p.sm_scale_k / p.logit_cap : 0.f;

  // lane-constant parts of the ldmatrix addresses (bytes inside a stage)
  // Two stage layouts.  Padded (bulk-copy path): row stride = row bytes + 16, so the 8 rows of an ldmatrix
  // fall into distinct bank groups.  Swizzled (cp.async path): row stride = row bytes and the 16-byte unit u of
  // row r sits at u ^ (r & 7) -- a quarter warp of LDGSTS then writes one aligned 128-byte line.  (With the
  // padded rows each quarter warp straddled two lines, L1 split the instruction into ~10 wavefronts instead
  // of 4 and fetched 1.5x the sectors from L2, with 2x the tag lookups: ncu r01 tp8.)  Every row a lane
  // addresses has r & 7 == mr, and a head starts on a 128-byte boundary, so the swizzle folds into
  // lane-constant offsets: K unit (mat + 4 k) ^ mr = (mat ^ (mr & 3)) + 4 ((k & 1) ^ (mr >> 2)) + 8 (k >> 1),
  // V unit ((mat >> 1) + 2 n) ^ mr = ((mat >> 1) ^ (mr & 1)) + 2 ((n & 3) ^ (mr >> 1)) + 8 (n >> 2).
  const bool swz = p.swizzle != 0;
  const uint32_t head_off = h * D * EB;
  uint32_t k_lane[2], v_lane;
  {
    const uint32_t krow = (tsi * 16 + mr) * p.row_stride + head_off;
    const uint32_t vrow = (tsi * 16 + (mat & 1) * 8 + mr) * p.row_stride + head_off + v_off;
    if (swz) {
      k_lane[0] = krow + 16 * ((mat ^ (mr & 3)) + 4 * (mr >> 2));
      k_lane[1] = krow + 16 * ((mat ^ (mr & 3)) + 4 * ((mr >> 2) ^ 1));
      v_lane = vrow + 16 * ((mat >> 1) ^ (mr & 1));
    } else {
      k_lane[0] = krow + mat * 16;
      k_lane[1] = k_lane[0] + 64;
      v_lane = vrow + (mat >> 1) * 16;
    }
  }
  const uint32_t v_x = swz ? (mr >> 1) * 32 : 0u;

  if (H.n_empty > 0) {
    // Requests of length 0 have no chunks, so no range reaches them: their rows are defined here
    // (o = 0, lse = -inf), the way the reference's stage 2 leaves an all-masked row.
    for (int i = cta * kConsumerWarps + warp; i < p.batch * p.hq; i += n_grid * kConsumerWarps) {
      const int bb = i / p.hq, hh = i - bb * p.hq;
      if (pv.seq_len[bb] != 0) continue;
      QT* orow = reinterpret_cast<QT*>(p.o) + bb * p.o_sb + hh * p.o_sh;
      for (int d = lane; d < D; d += 32) store_q(orow + d, 0.f);
      if (p.lse && lane == 0) p.lse[static_cast<int64_t>(bb) * p.hq + hh] = -CUDART_INF_F;
    }
  }

  // streaming rate of this CTA (clocks per chunk over its static range, pipeline fill included -- what the plan
  // wants to equalise is when the CTAs END): folded into cta_stat so that the next plan can size the ranges by
  // it.  Stamped per range, not per chunk: on an fp8 pool the consumers are the bound and warp 0 is on every
  // chunk's critical path (thread-0 work inside the chunk loop cost C3 10 %, profiles/r02_baseline).
  const long long clk_first = clock64();
  long long clk_last = clk_first;
  int n_consumed = 0;
  int npiece = 0;     // pieces seen by this CTA: parity picks the flag slot
  while (rg < R) {
  PieceWalk walk(rn == 0 && rg == cta ? d0 : load_desc(pv.range_desc + rg), p.ng);
  while (walk.next(pv, b, hg, lc0, lc1, nchunks, vreq, S, row_off)) {
    const int hq0 = (hg * p.hc + h) * p.group;
    const bool row0 = gq < p.group;
    const bool row1 = ROWS16 && (gq + 8 < p.group);
    const bool full_req = lc0 == 0 && lc1 == nchunks;
    // merge bookkeeping of a request cut by a range boundary, fetched now so that the epilogue does not wait for it
    int v_pieces = 1, v_rfirst = rg;   // a whole request is one piece, in this range
    if (!full_req) {
      v_pieces = pv.vreq_pieces[vreq];
      v_rfirst = pv.vreq_rfirst[vreq];
    }
    int early_ok = 0;   // thread 0: every other piece of this request had arrived when the last chunk began
    const bool try_early = !full_req && p.fuse_merge && v_pieces <= 4;
    ++npiece;
    const int hqv = p.hc * p.group;
    const int64_t prow0 = static_cast<int64_t>(vreq + v_rfirst) * p.ts * hqv;   // first piece's workspace rows
    const int64_t pstep = static_cast<int64_t>(p.ts) * hqv;                      // piece to piece (slice 0 rows)
    const int me = rg - v_rfirst;                                                // this piece's index
    // A few chunks before the end of a cut request's piece, the other pieces' partial rows of this warp's head are
    // pulled into L2: the CTA that merges (normally this one: its neighbour finished the other half long ago, and
    // ~1 GB has gone through L2 since) then reads them in one short round trip instead of a DRAM access queued
    // behind the KV stream (2 us each under load, tools/trace_decode.py: 7 us from the last chunk to the last
    // store on C2).  A row that has not been written yet is harmless to prefetch: L2 is the coherence point.
    const int lc_pf = lc1 - 4 > lc0 ? lc1 - 4 : lc0;

    // ---- Q fragments (A operand), loaded once per piece ----
    uint32_t qf[KSTEPS][QR];
    {
      const QT* q0 = reinterpret_cast<const QT*>(p.q) + b * p.q_sb + (hq0 + gq) * p.q_sh;
      const QT* q1 = q0 + 8 * p.q_sh;
#pragma unroll
      for (int kk = 0; kk < KSTEPS; ++kk) {
        if constexpr (!FP8) {
          const int d0 = kk * 16 + 2 * jq;
          qf[kk][0] = row0 ? *reinterpret_cast<const uint32_t*>(q0 + d0) : 0u;
          qf[kk][1] = row0 ? *reinterpret_cast<const uint32_t*>(q0 + d0 + 8) : 0u;
          if constexpr (ROWS16) {
            qf[kk][2] = row1 ? *reinterpret_cast<const uint32_t*>(q1 + d0) : 0u;
            qf[kk][3] = row1 ? *reinterpret_cast<const uint32_t*>(q1 + d0 + 8) : 0u;
          }
        } else {
          // fp8 K: MMA k-slot (2j, 2j+1, 2j+8, 2j+9) of step kk <-> d = 16kk + 4j + (0..3)
          const int d0 = kk * 16 + 4 * jq;
          float f[4] = {0.f, 0.f, 0.f, 0.f};
          if (row0) {
#pragma unroll
            for (int e = 0; e < 4; ++e) f[e] = fminf(fmaxf(q_to_f32(q0[d0 + e]), -65504.f), 65504.f);
          }
          qf[kk][0] = pack2<true>(f[0], f[1]);
          qf[kk][1] = pack2<true>(f[2], f[3]);
          if constexpr (ROWS16) {
            float f1[4] = {0.f, 0.f, 0.f, 0.f};
            if (row1) {
#pragma unroll
              for (int e = 0; e < 4; ++e) f1[e] = fminf(fmaxf(q_to_f32(q1[d0 + e]), -65504.f), 65504.f);
            }
            qf[kk][2] = pack2<true>(f1[0], f1[1]);
            qf[kk][3] = pack2<true>(f1[2], f1[3]);
          }
        }
      }
    }

    float o[NTILES][QR];
#pragma unroll
    for (int n = 0; n < NTILES; ++n)
#pragma unroll
      for (int e = 0; e < QR; ++e) o[n][e] = 0.f;
    float m0 = -CUDART_INF_F, m1 = -CUDART_INF_F, l0 = 0.f, l1 = 0.f;

#ifdef B200ATTN_DEC_TRACE
    if (threadIdx.x == 0) { B200_TR_ONCE(TR_C_Q, tr_a); ++tr_pieces; }
#endif
    for (int lc = lc0; lc < lc1; ++lc) {
      if (lc == lc_pf && try_early) {
        const int g = lane >> 2, line = lane & 3;
        if (g < p.group && line * 32 < D) {
          const int rowh = h * p.group + g;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            if (i < v_pieces && i != me) {
              const int64_t row = prow0 + i * pstep + rowh;
              asm volatile("prefetch.global.L2 [%0];" ::"l"(p.ws_o + row * D + line * 32));
              if (line == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(p.ws_lse + row));
            }
          }
        }
      }
      if (lc == lc1 - 1 && try_early && threadIdx.x == 0) {   // (warp-uniform tests first: two instructions per chunk)
        // Last-arriver shortcut, decided one chunk early (the load's round trip hides behind the chunk): if
        // every other piece of the request is already in the workspace, this CTA is the one that merges, so
        // its own partial never needs to go through global memory (see the epilogue).
        int arrived;
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(arrived) : "l"(pv.vreq_arrived + vreq) : "memory");
        early_ok = arrived == v_pieces - 1;
      }
      mbar_wait(bar_full + stage * 8, phase);
      if (threadIdx.x == 0) B200_TR_ONCE(TR_C_FIRST_FULL, tr_b);
      const uint32_t sbase = stage0 + stage * p.stage_bytes;

      // ---- S = Q.K^T for 16 tokens (two n8 tiles) ----
      float s[2][4];
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
        const uint32_t ka0 = sbase + k_lane[0] + nt * 8 * p.row_stride;
        const uint32_t ka1 = sbase + k_lane[1] + nt * 8 * p.row_stride;
        if constexpr (!FP8) {
#pragma unroll
          for (int k2 = 0; k2 < KSTEPS / 2; ++k2) {
            uint32_t kf[4];
            ldsm_x4(((k2 & 1) ? ka1 : ka0) + (k2 >> 1) * 128, kf);
            mma_q<HALF, ROWS16, QR>(s[nt], qf[2 * k2], kf[0], kf[1]);
            mma_q<HALF, ROWS16, QR>(s[nt], qf[2 * k2 + 1], kf[2], kf[3]);
          }
        } else {
#pragma unroll
          for (int k4 = 0; k4 < KSTEPS / 4; ++k4) {
            uint32_t kf[4];
            ldsm_x4(((k4 & 1) ? ka1 : ka0) + (k4 >> 1) * 128, kf);
#pragma unroll
            for (int mm = 0; mm < 4; ++mm) {
              const int kk = 4 * k4 + mm;
              const uint32_t b0 = e4m3x2_to_f16x2(kf[mm]);
              const uint32_t b1 = e4m3x2_to_f16x2(kf[mm] >> 16);
              mma_q<true, ROWS16, QR>(s[nt], qf[kk], b0, b1);
            }
          }
        }
      }

      // ---- logit cap, tail mask ----
      if (cap) {
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
          for (int e = 0; e < 4; ++e) s[nt][e] = p.logit_cap * tanhf(s[nt][e] * cap_pre);
      }
      const int tok0 = lc * T + tsi * 16 + 2 * jq;  // token of s[0][0]
      const bool tail = (lc + 1) * T > S;
      if (tail) {
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (tok0 + nt * 8 + (e & 1) >= S) s[nt][e] = -CUDART_INF_F;
      }

      // ---- online softmax (rows gq and, for wide groups, gq + 8) ----
      float mx0 = fmaxf(fmaxf(s[0][0], s[0][1]), fmaxf(s[1][0], s[1][1]));
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
      const float mn0 = fmaxf(m0, mx0);
      const float mr0 = mn0 == -CUDART_INF_F ? 0.f : mn0;
      const float corr0 = fast_exp2((m0 - mr0) * c);
      m0 = mn0;
      const float mc0 = mr0 * c;
      float pr[2][4];
      pr[0][0] = fast_exp2(fmaf(s[0][0], c, -mc0));
      pr[0][1] = fast_exp2(fmaf(s[0][1], c, -mc0));
      pr[1][0] = fast_exp2(fmaf(s[1][0], c, -mc0));
      pr[1][1] = fast_exp2(fmaf(s[1][1], c, -mc0));
      l0 = fmaf(l0, corr0, (pr[0][0] + pr[0][1]) + (pr[1][0] + pr[1][1]));
      float corr1 = 1.f;
      if constexpr (ROWS16) {
        float mx1 = fmaxf(fmaxf(s[0][2], s[0][3]), fmaxf(s[1][2], s[1][3]));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float mn1 = fmaxf(m1, mx1);
        const float mr1 = mn1 == -CUDART_INF_F ? 0.f : mn1;
        corr1 = fast_exp2((m1 - mr1) * c);
        m1 = mn1;
        const float mc1 = mr1 * c;
        pr[0][2] = fast_exp2(fmaf(s[0][2], c, -mc1));
        pr[0][3] = fast_exp2(fmaf(s[0][3], c, -mc1));
        pr[1][2] = fast_exp2(fmaf(s[1][2], c, -mc1));
        pr[1][3] = fast_exp2(fmaf(s[1][3], c, -mc1));
        l1 = fmaf(l1, corr1, (pr[0][2] + pr[0][3]) + (pr[1][2] + pr[1][3]));
      }
      if (__any_sync(0xffffffffu, corr0 != 1.f || corr1 != 1.f)) {
#pragma unroll
        for (int n = 0; n < NTILES; ++n) {
          o[n][0] *= corr0;
          o[n][1] *= corr0;
          if constexpr (ROWS16) {
            o[n][2] *= corr1;
            o[n][3] *= corr1;
          }
        }
      }

      // ---- P (A operand) ----
      const uint32_t pa0 = pack2<HALF>(pr[0][0], pr[0][1]);
      const uint32_t pa2 = pack2<HALF>(pr[1][0], pr[1][1]);
      uint32_t pa1 = 0u, pa3 = 0u;
      if constexpr (ROWS16) {
        pa1 = pack2<HALF>(pr[0][2], pr[0][3]);
        pa3 = pack2<HALF>(pr[1][2], pr[1][3]);
      }
      // V fragments of masked tail tokens are zeroed: the smem rows behind them
      // hold stale bytes of another request (0 * inf would poison this one).
      uint32_t vm0 = 0xffffffffu, vm1 = 0xffffffffu;
      if (tail) {
        vm0 = (tok0 < S ? 0x0000ffffu : 0u) | (tok0 + 1 < S ? 0xffff0000u : 0u);
        vm1 = (tok0 + 8 < S ? 0x0000ffffu : 0u) | (tok0 + 9 < S ? 0xffff0000u : 0u);
      }

      // ---- O += P.V ----  (the masks cost 32 LOP3 per 16 tokens: only the tail chunk pays them)
      auto pv_mma = [&](auto tail_tag) {
        constexpr bool TAIL = decltype(tail_tag)::value;
        const uint32_t m0 = TAIL ? vm0 : 0xffffffffu, m1 = TAIL ? vm1 : 0xffffffffu;
        const uint32_t va = sbase + v_lane;
        if constexpr (!FP8) {
  #pragma unroll
          for (int n2 = 0; n2 < NTILES / 2; ++n2) {
            uint32_t vf[4];
            ldsm_x4_trans(va + (((n2 & 3) * 32) ^ v_x) + (n2 >> 2) * 128, vf);
            if constexpr (ROWS16) {
              mma16816<HALF>(o[2 * n2], pa0, pa1, pa2, pa3, vf[0] & m0, vf[1] & m1);
              mma16816<HALF>(o[2 * n2 + 1], pa0, pa1, pa2, pa3, vf[2] & m0, vf[3] & m1);
            } else {
              float d0[4] = {o[2 * n2][0], o[2 * n2][1], 0.f, 0.f};
              mma16816<HALF>(d0, pa0, 0u, pa2, 0u, vf[0] & m0, vf[1] & m1);
              o[2 * n2][0] = d0[0];
              o[2 * n2][1] = d0[1];
              float d1[4] = {o[2 * n2 + 1][0], o[2 * n2 + 1][1], 0.f, 0.f};
              mma16816<HALF>(d1, pa0, 0u, pa2, 0u, vf[2] & m0, vf[3] & m1);
              o[2 * n2 + 1][0] = d1[0];
              o[2 * n2 + 1][1] = d1[1];
            }
          }
        } else {
          // fp8 V: one x4.trans covers 32 d; a b16 transpose leaves each register
          // with {tok 2j: d, d+1 | tok 2j+1: d, d+1}; PRMT regroups by d parity.
  #pragma unroll
          for (int n4 = 0; n4 < NTILES / 4; ++n4) {
            uint32_t vf[4];
            ldsm_x4_trans(va + (((n4 & 3) * 32) ^ v_x) + (n4 >> 2) * 128, vf);
  #pragma unroll
            for (int gg = 0; gg < 2; ++gg) {
  #pragma unroll
              for (int par = 0; par < 2; ++par) {
                const uint32_t sel = par ? 0x4431u : 0x4420u;
                const uint32_t b0 = e4m3x2_to_f16x2(__byte_perm(vf[2 * gg], 0u, sel)) & m0;
                const uint32_t b1 = e4m3x2_to_f16x2(__byte_perm(vf[2 * gg + 1], 0u, sel)) & m1;
                const int n = n4 * 4 + gg * 2 + par;
                if constexpr (ROWS16) {
                  mma16816<true>(o[n], pa0, pa1, pa2, pa3, b0, b1);
                } else {
                  float d0[4] = {o[n][0], o[n][1], 0.f, 0.f};
                  mma16816<true>(d0, pa0, 0u, pa2, 0u, b0, b1);
                  o[n][0] = d0[0];
                  o[n][1] = d0[1];
                }
              }
            }
          }
        }

      };
      if (tail) pv_mma(std::true_type{}); else pv_mma(std::false_type{});
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_empty + stage * 8);
#ifdef B200ATTN_DEC_TRACE
      if (threadIdx.x == 0) { B200_TR_ONCE(TR_C_FIRST_DONE, tr_c); B200_TR(TR_C_LAST_DONE); ++tr_chunks; }
#endif
      if (++stage == p.stages) {
        stage = 0;
        phase ^= 1;
      }
    }

    // ---- epilogue: normalise, write final o or a partial ----
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    const float inv0 = l0 > 0.f ? p.v_scale / l0 : 0.f;
    const float lse0 = l0 > 0.f ? m0 * c * kLn2 + logf(l0) : -CUDART_INF_F;
    float inv1 = 0.f, lse1 = -CUDART_INF_F;
    if constexpr (ROWS16) {
      l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
      l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
      inv1 = l1 > 0.f ? p.v_scale / l1 : 0.f;
      lse1 = l1 > 0.f ? m1 * c * kLn2 + logf(l1) : -CUDART_INF_F;
    }
    // column (head-dim index) of accumulator element e (0/1) of tile n
    auto col_of = [&](int n, int e) -> int {
      if constexpr (!FP8) return n * 8 + 2 * jq + e;
      return (n >> 1) * 16 + 4 * jq + 2 * e + (n & 1);
    };
    // Workspace rows of this piece: ((vreq + rg) * ts + slice) * hqv + head; with the in-CTA slice merge below
    // only slice 0's rows are used (one partial per piece instead of ts).
    const int64_t piece_row = static_cast<int64_t>(vreq + rg) * p.ts * hqv;
    const bool use_xch = !ROWS16 && p.xch_off != 0;   // implies ts > 1 (make_geometry)

    // Last-arriver shortcut (partial pieces): when every other piece of the request had already arrived one
    // chunk ago, this CTA is the one that merges -- so it merges its own partial straight from registers with
    // the others read from the workspace, instead of write partial -> fence -> atomic -> read everything back
    // (four dependent global round trips of ~1.5 us each under load at the end of a CTA's range:
    // tools/trace_decode.py, 6.6 us from the last chunk to the exit on C2).  Same values, same order, same
    // arithmetic as merge_heads4, so which path ran never shows in the result.
    const bool can_fast = !ROWS16 && !full_req && p.fuse_merge && (use_xch || p.ts == 1);
    volatile int32_t* flags = merge_flag + (npiece & 1) * 2;
    if (can_fast && threadIdx.x == 0) flags[0] = early_ok;
    if (can_fast && !use_xch) named_bar_sync(2, kConsumerWarps * 32);   // (the slice exchange has its own barrier)
    float wk[4], inv_s = 0.f, lse_tot = -CUDART_INF_F;
    // The merge's loads go out in ONE batch -- the other pieces' lse and the rows of the first other piece (the
    // only other piece in all but the rarest plans) -- before anything waits on them: issued one after the other's
    // result (lse -> weights -> row loads per tile) the shortcut still cost 3-4 dependent round trips.
    const int i1 = me == 0 ? 1 : 0;   // the first other piece
    auto merge_lse_load = [&](int rowh, float lse_mine, float (&l)[4]) {
#pragma unroll
      for (int i = 0; i < 4; ++i)
        l[i] = i < v_pieces ? (i == me ? lse_mine : __ldcg(p.ws_lse + prow0 + i * pstep + rowh)) : -CUDART_INF_F;
    };
    auto merge_weights = [&](const float (&l)[4]) {
      const float m = fmaxf(fmaxf(l[0], l[1]), fmaxf(l[2], l[3]));
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        wk[i] = m != -CUDART_INF_F ? __expf(l[i] - m) : 0.f;
        s += wk[i];
      }
      inv_s = s > 0.f ? 1.f / s : 0.f;
      lse_tot = s > 0.f ? m + logf(s) : -CUDART_INF_F;
    };
    // columns col_of(n, 0 / 1) of workspace row w
    auto load_pair = [&](const float* w, int n, float& v0, float& v1) {
      if constexpr (!FP8) {
        const float2 t = __ldcg(reinterpret_cast<const float2*>(w + n * 8 + 2 * jq));
        v0 = t.x;
        v1 = t.y;
      } else {
        v0 = __ldcg(w + col_of(n, 0));
        v1 = __ldcg(w + col_of(n, 1));
      }
    };
    // x1_0 / x1_1: the first other piece's values of this tile, already in registers
    auto merge_tile = [&](int rowh, int n, float x0, float x1, float x1_0, float x1_1, float& y0, float& y1) {
      float a0 = 0.f, a1 = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (i < v_pieces) {
          float v0 = x0, v1 = x1;
          if (i == i1) {
            v0 = x1_0;
            v1 = x1_1;
          } else if (i != me) {
            load_pair(p.ws_o + (prow0 + i * pstep + rowh) * D, n, v0, v1);
          }
          a0 = fmaf(wk[i], v0, a0);
          a1 = fmaf(wk[i], v1, a1);
        }
      }
      y0 = a0 * inv_s;
      y1 = a1 * inv_s;
    };

    bool final_written = false;
    if (p.ts == 1 && full_req) {
      final_written = true;
      QT* o0 = reinterpret_cast<QT*>(p.o) + b * p.o_sb + (hq0 + gq) * p.o_sh;
      if (row0) {
#pragma unroll
        for (int n = 0; n < NTILES; ++n) {
          store_q(o0 + col_of(n, 0), o[n][0] * inv0);
          store_q(o0 + col_of(n, 1), o[n][1] * inv0);
        }
        if (p.lse && jq == 0) p.lse[static_cast<int64_t>(b) * p.hq + hq0 + gq] = lse0;
      }
      if constexpr (ROWS16) {
        if (row1) {
          QT* o1 = o0 + 8 * p.o_sh;
#pragma unroll
          for (int n = 0; n < NTILES; ++n) {
            store_q(o1 + col_of(n, 0), o[n][2] * inv1);
            store_q(o1 + col_of(n, 1), o[n][3] * inv1);
          }
          if (p.lse && jq == 0) p.lse[static_cast<int64_t>(b) * p.hq + hq0 + gq + 8] = lse1;
        }
      }
    } else if (use_xch) {
      // The ts token slices of this piece meet in shared memory (unnormalised O, max, sum) and each warp
      // finishes 1 / ts of its head's columns: a whole request goes straight to `o`; a piece of a request cut
      // by a range boundary leaves ONE normalised partial per head in the workspace instead of ts.  Through
      // the workspace the slice merge cost a write -> fence -> atomic -> read-back chain of global round trips
      // per piece with the consumers idle, and ts x the rows for the cross-CTA merge to read back.
      const int g4 = p.group * 4;
      float2* xall = reinterpret_cast<float2*>(smem + p.xch_off);
      float2* xml = xall + kXchWarps * NTILES * g4;
      const int li = gq * 4 + jq;
      if (row0) {
        float2* xo = xall + warp * (NTILES * g4) + li;
#pragma unroll
        for (int n = 0; n < NTILES; ++n) xo[n * g4] = make_float2(o[n][0], o[n][1]);
        if (jq == 0) xml[warp * 8 + gq] = make_float2(l0 > 0.f ? m0 * c : -CUDART_INF_F, l0);
      }
      named_bar_sync(2, kConsumerWarps * 32);
      const bool fast = can_fast && flags[0] != 0;
      final_written = full_req || fast;
      if (row0) {
        float wgt[kXchWarps];
        float mm = -CUDART_INF_F;
#pragma unroll
        for (int s = 0; s < kXchWarps; ++s)
          if (s < p.ts) mm = fmaxf(mm, xml[(s * p.hc + h) * 8 + gq].x);
        float L = 0.f;
#pragma unroll
        for (int s = 0; s < kXchWarps; ++s) {
          wgt[s] = 0.f;
          if (s < p.ts) {
            const float2 ml = xml[(s * p.hc + h) * 8 + gq];
            wgt[s] = ml.y > 0.f ? fast_exp2(ml.x - mm) : 0.f;
            L = fmaf(wgt[s], ml.y, L);
          }
        }
        const float inv = L > 0.f ? p.v_scale / L : 0.f;
        const float lse_all = L > 0.f ? mm * kLn2 + logf(L) : -CUDART_INF_F;
        const int rowh = h * p.group + gq;
        QT* o0 = reinterpret_cast<QT*>(p.o) + b * p.o_sb + (hq0 + gq) * p.o_sh;
        float* w0 = p.ws_o + (piece_row + rowh) * D;
        // this warp finishes tiles tsi, tsi + ts, ...: at most NTILES / 2 of them (ts >= 2)
        float x1[NTILES / 2][2];
        if (fast) {
          float l[4];
          merge_lse_load(rowh, lse_all, l);
          const float* w1 = p.ws_o + (prow0 + i1 * pstep + rowh) * D;
#pragma unroll
          for (int k = 0; k < NTILES / 2; ++k) {
            const int n = tsi + k * p.ts;
            x1[k][0] = x1[k][1] = 0.f;
            if (n < NTILES) load_pair(w1, n, x1[k][0], x1[k][1]);
          }
          merge_weights(l);
        }
        int kt = 0;
        for (int n = tsi; n < NTILES; n += p.ts, ++kt) {
          float a0 = 0.f, a1 = 0.f;
#pragma unroll
          for (int s = 0; s < kXchWarps; ++s) {
            if (s < p.ts) {
              const float2 v = xall[(s * p.hc + h) * (NTILES * g4) + n * g4 + li];
              a0 = fmaf(wgt[s], v.x, a0);
              a1 = fmaf(wgt[s], v.y, a1);
            }
          }
          if (full_req) {
            store_q(o0 + col_of(n, 0), a0 * inv);
            store_q(o0 + col_of(n, 1), a1 * inv);
          } else if (fast) {
            float y0, y1;
            float f0 = 0.f, f1 = 0.f;
#pragma unroll
            for (int k = 0; k < NTILES / 2; ++k)   // (a run-time index would send x1 to local memory)
              if (k == kt) {
                f0 = x1[k][0];
                f1 = x1[k][1];
              }
            merge_tile(rowh, n, a0 * inv, a1 * inv, f0, f1, y0, y1);
            store_q(o0 + col_of(n, 0), y0);
            store_q(o0 + col_of(n, 1), y1);
          } else {
            w0[col_of(n, 0)] = a0 * inv;
            w0[col_of(n, 1)] = a1 * inv;
          }
        }
        if (jq == 0 && tsi == 0) {
          if (full_req) {
            if (p.lse) p.lse[static_cast<int64_t>(b) * p.hq + hq0 + gq] = lse_all;
          } else if (fast) {
            if (p.lse) p.lse[static_cast<int64_t>(b) * p.hq + hq0 + gq] = lse_tot;
          } else {
            p.ws_lse[piece_row + rowh] = lse_all;
          }
        }
      }
      // the closing barrier (the area is free for the next piece) is the one in front of the arrival counter
      // below when the partial goes through the workspace
      if (final_written || !p.fuse_merge) named_bar_sync(2, kConsumerWarps * 32);
    } else if (can_fast && flags[0] != 0) {
      // ts == 1: this warp owns its head's rows outright
      final_written = true;
      if (row0) {
        const int rowh = h * p.group + gq;
        float l[4], x1[NTILES][2];
        merge_lse_load(rowh, lse0, l);
        const float* w1 = p.ws_o + (prow0 + i1 * pstep + rowh) * D;
#pragma unroll
        for (int n = 0; n < NTILES; ++n) load_pair(w1, n, x1[n][0], x1[n][1]);
        merge_weights(l);
        QT* o0 = reinterpret_cast<QT*>(p.o) + b * p.o_sb + (hq0 + gq) * p.o_sh;
#pragma unroll
        for (int n = 0; n < NTILES; ++n) {
          float y0, y1;
          merge_tile(rowh, n, o[n][0] * inv0, o[n][1] * inv0, x1[n][0], x1[n][1], y0, y1);
          store_q(o0 + col_of(n, 0), y0);
          store_q(o0 + col_of(n, 1), y1);
        }
        if (p.lse && jq == 0) p.lse[static_cast<int64_t>(b) * p.hq + hq0 + gq] = lse_tot;
      }
    } else {
      const int64_t row = piece_row + static_cast<int64_t>(tsi) * hqv + h * p.group + gq;
      if (row0) {
        float* w = p.ws_o + row * D;
#pragma unroll
        for (int n = 0; n < NTILES; ++n) {
          w[col_of(n, 0)] = o[n][0] * inv0;
          w[col_of(n, 1)] = o[n][1] * inv0;
        }
        if (jq == 0) p.ws_lse[row] = lse0;
      }
      if constexpr (ROWS16) {
        if (row1) {
          float* w = p.ws_o + (row + 8) * D;
#pragma unroll
          for (int n = 0; n < NTILES; ++n) {
            w[col_of(n, 0)] = o[n][2] * inv1;
            w[col_of(n, 1)] = o[n][3] * inv1;
          }
          if (jq == 0) p.ws_lse[row + 8] = lse1;
        }
      }
    }
    if (final_written && !full_req && threadIdx.x == 0) {
      pv.vreq_arrived[vreq] = 0;   // shortcut taken: the counter is ready for the next layer's launch
      B200_TR(TR_C_MERGED);
    }
    if (!final_written && p.fuse_merge) {
      // Stage 2 without a second launch: the CTA that lands a virtual request's last piece
      // merges all of them (threadfence + counter, the "last block done" pattern).  The
      // producers keep streaming the next piece's chunks meanwhile.
      __threadfence();
      named_bar_sync(2, kConsumerWarps * 32);
      if (threadIdx.x == 0) {
        const int arrived = atomicAdd(&pv.vreq_arrived[vreq], 1) + 1;
        const int last = arrived == v_pieces;
        if (last) pv.vreq_arrived[vreq] = 0;  // ready for the next layer's launch
        flags[1] = last;
      }
      named_bar_sync(2, kConsumerWarps * 32);
      if (flags[1]) {
        if (threadIdx.x == 0) B200_TR(TR_C_MERGED);
        __threadfence();
        const int nsl = use_xch ? 1 : p.ts;            // partial rows per piece
        const int np = v_pieces * nsl;
        const int64_t step = static_cast<int64_t>(use_xch ? p.ts : 1) * hqv;
        QT* ob = reinterpret_cast<QT*>(p.o) + b * p.o_sb + static_cast<int64_t>(hg) * hqv * p.o_sh;
        float* lb = p.lse ? p.lse + static_cast<int64_t>(b) * p.hq + hg * hqv : nullptr;
        if (np <= 4 && D <= 128) {
          // few partials: four heads per warp at once, every load in flight together (one L2 round trip)
          for (int hl = warp; hl < hqv; hl += 4 * kConsumerWarps) {
            const int nh = (hqv - hl + kConsumerWarps - 1) / kConsumerWarps;
            merge_heads4<QT>(p.ws_o, p.ws_lse, prow0 + hl, kConsumerWarps, nh < 4 ? nh : 4, step, np, D,
                             ob + hl * p.o_sh, kConsumerWarps * p.o_sh, lb ? lb + hl : nullptr, kConsumerWarps, lane);
          }
        } else if (D <= 128) {
          // many partials: one head per lane group (two at D = 64), 16 partials in flight per round
          constexpr int HPW = 128 / D;   // heads per warp and pass
          for (int hl = warp; hl < hqv; hl += HPW * kConsumerWarps) {
            const int nh = (hqv - hl + kConsumerWarps - 1) / kConsumerWarps;
            merge_rows_wide<QT>(p.ws_o, p.ws_lse, prow0 + hl, kConsumerWarps, nh < HPW ? nh : HPW, step, np, D,
                                ob + hl * p.o_sh, kConsumerWarps * p.o_sh, lb ? lb + hl : nullptr, kConsumerWarps, lane);
          }
        } else {
          for (int hl = warp; hl < hqv; hl += kConsumerWarps)
            merge_head<QT>(p.ws_o, p.ws_lse, prow0 + hl, step, np, D, ob + hl * p.o_sh, lb ? lb + hl : nullptr, lane);
        }
      }
    }
    if (threadIdx.x == 0) B200_TR(TR_C_EPILOGUE);
    if (rn == 0) n_consumed += lc1 - lc0;
  }
  if (rn == 0) clk_last = clock64();
  ++rn;
  rg = rq[rn & (kRangeRing - 1)];
  }
  if (threadIdx.x == 0 && n_consumed >= 16 && cta < kMaxCtas) {
    const float r = static_cast<float>(clk_last - clk_first) / static_cast<float>(n_consumed);
    CtaStat st = st_prev;
    const bool ok = st.tag == (kStatTag ^ static_cast<uint32_t>(cta)) && st.clk_per_chunk > 0.f;
    st.clk_per_chunk = ok ? 0.75f * st.clk_per_chunk + 0.25f * r : r;
    st.tag = kStatTag ^ static_cast<uint32_t>(cta);
    pv.cta_stat[cta] = st;
  }
#ifdef B200ATTN_DEC_TRACE
  if (threadIdx.x == 0 && p.trace) {
    B200_TR(TR_C_EXIT);
    p.trace[blockIdx.x * kTraceSlots + TR_COUNTS] = tr_chunks | (tr_pieces << 32);
    B200_TR_GT(TR_GT_EXIT);
  }
#endif
}

template <typename QT, bool FP8>
cudaError_t launch_decode(const DecodeParams& p, int head_dim, bool rows16, int grid, int smem,
                          cudaStream_t stream);

// The dynamic shared-memory limit is a per-function, per-device attribute: set on first use, not per launch.
inline cudaError_t ensure_smem_attr(const void* fn, int smem, int* cache /* [64] */) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = -1;
  if (dev >= 0 && cache[dev] >= smem) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e == cudaSuccess && dev >= 0) cache[dev] = smem;
  return e;
}

#define B200ATTN_DEFINE_LAUNCH(QT, FP8)                                                          \
  template <>                                                                                    \
  cudaError_t launch_decode<QT, FP8>(const DecodeParams& p, int head_dim, bool rows16, int grid, \
                                     int smem, cudaStream_t stream) {                            \
    static int attr_cache[4][64];                                                                \
    void (*fn)(const DecodeParams) = nullptr;                                                    \
    int slot = 0;                                                                                \
    if (head_dim == 64) { fn = rows16 ? decode_kernel<QT, FP8, 64, true> : decode_kernel<QT, FP8, 64, false>; slot = rows16; }        \
    else if (head_dim == 128) { fn = rows16 ? decode_kernel<QT, FP8, 128, true> : decode_kernel<QT, FP8, 128, false>; slot = 2 + rows16; } \
    if (!fn) return cudaErrorInvalidValue;                                                       \
    cudaError_t e = ensure_smem_attr(reinterpret_cast<const void*>(fn), smem, attr_cache[slot]); \
    if (e != cudaSuccess) return e;                                                              \
    cudaLaunchConfig_t cfg = {};                                                                 \
    cfg.gridDim = dim3(grid);                                                                    \
    cfg.blockDim = dim3(kThreads);                                                               \
    cfg.dynamicSmemBytes = smem;                                                                 \
    cfg.stream = stream;                                                                         \
    cudaLaunchAttribute at[1];                                                                   \
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                               \
    at[0].val.programmaticStreamSerializationAllowed = 1;                                        \
    cfg.attrs = at;                                                                              \
    cfg.numAttrs = no_pdl() ? 0 : 1;                                                             \
    return cudaLaunchKernelEx(&cfg, fn, p);                                                      \
  }

}  // namespace b200attn
