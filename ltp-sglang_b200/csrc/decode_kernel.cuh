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
  const int64_t* slot_map;       // optional full -> swa slot translation (SWAKVPool), applied to every id read
  int64_t pool_slots;            // rows of the pools (0 = unknown); checked by the -DB200ATTN_DEBUG build only
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
  int32_t kv_split;              // 1: pipeline slots alternate K rows / V rows of a chunk (stage_bytes = one tensor)
  int32_t subs;                  // 16-token sub-chunks per warp and chunk (chunk_tokens = 16 * ts * subs)
  int32_t e5m2;                  // fp8 pools: 1 = float8_e5m2 bytes, 0 = float8_e4m3fn
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
constexpr int kTraceSlots = 24;
enum {
  TR_GT_ENTRY = 0, TR_ENTRY, TR_BARS_READY, TR_P_WALK, TR_P_IDS, TR_P_FIRST_ISSUE, TR_P_LAST_ISSUE, TR_C_Q,
  TR_C_FIRST_FULL, TR_C_FIRST_DONE, TR_C_LAST_DONE, TR_C_EPILOGUE, TR_C_MERGED, TR_C_EXIT, TR_COUNTS, TR_GT_EXIT,
  // one steady-state chunk (the 65th this CTA consumes) of consumer warp 0, each stamp taken after an instruction
  // that needs the phase's result: wait begins, stage full, S^T done, softmax done (P^T in registers), O^T updated
  TR_K_WAIT, TR_K_FULL, TR_K_QK, TR_K_SOFTMAX, TR_K_PV, TR_K_END, TR_K_WAIT2
};
// clock64 read after `val` is available (the compare stalls on the producing instruction's scoreboard; warps issue in order)
__device__ __forceinline__ unsigned long long clock_after(float val) {
  unsigned long long t;
  asm volatile("{\n\t.reg .pred p;\n\tsetp.eq.f32 p, %1, 0f7fc00001;\n\tmov.u64 %0, %%clock64;\n\t@p add.u64 %0, %0, 1;\n\t}" : "=l"(t) : "f"(val));
  return t;
}
#define B200_TR_DEP(slot, val) do { if (p.trace && tid == 0 && tr_chunks == 64) p.trace[blockIdx.x * kTraceSlots + (slot)] = clock_after(val); } while (0)
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
#define B200_TR_DEP(slot, val) do { } while (0)
#endif

// -DB200ATTN_DEBUG (make debug -> libb200attn_debug.so): a slot id outside the pool traps instead of gathering
// out of bounds; the stage offsets of the consumers are checked against the dynamic smem size at entry.
#ifdef B200ATTN_DEBUG
#define B200_CHECK_SLOT(id) do { if (p.pool_slots > 0 && static_cast<uint64_t>(static_cast<int64_t>(id)) >= static_cast<uint64_t>(p.pool_slots)) __trap(); } while (0)
#else
#define B200_CHECK_SLOT(id) do { } while (0)
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
  constexpr int EB = FP8 ? 1 : 2;      // KV element bytes

  extern __shared__ __align__(128) uint8_t smem[];
  const int cta = blockIdx.x;
  // %tid.x read once into a register the compiler cannot rematerialise: every `tid == 0` test in the chunk
  // loops was otherwise a fresh S2R (tens of clocks each, on the consumers' critical path)
  int tid;
  asm volatile("mov.u32 %0, %%tid.x;" : "=r"(tid));
  const int warp = tid >> 5, lane = tid & 31;
  uint32_t bar_full = smem_u32(smem);
  // kept in a register: left to the compiler the address was rebuilt from %cluster_ctaid (an S2R of tens of
  // clocks) in front of every barrier wait of the consumers' chunk loop
  asm volatile("" : "+r"(bar_full));
  const uint32_t bar_empty = bar_full + kMaxStages * 8;
  const uint32_t stage0 = bar_full + kBarrierBytes;
  volatile int32_t* rq = reinterpret_cast<volatile int32_t*>(smem + kRangeRingOff);   // ring of range ids
  volatile int32_t* merge_flag = reinterpret_cast<volatile int32_t*>(smem + kMergeFlagOff);
#ifdef B200ATTN_DEC_TRACE
  bool tr_a = false, tr_b = false, tr_c = false;
  unsigned long long tr_chunks = 0, tr_pieces = 0, tr_last = 0;
  if (tid == 0) {
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
  if (tid == 0) {
    for (int s = 0; s < p.stages; ++s) {
      // arrivals on a slot's `full`: every producer thread (cp.async path) or warp (bulk path) that fills it --
      // all four producer warps, or the two of one tensor when slots alternate K / V
      mbar_init(bar_full + s * 8, (p.cp_async ? kProducerWarps * 32 : kProducerWarps) >> p.kv_split);
      mbar_init(bar_empty + s * 8, kConsumerWarps);
    }
    fence_barrier_init();
  }
#ifdef B200ATTN_DEBUG
  {
    uint32_t dyn;
    asm volatile("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn));
    const uint32_t need = kBarrierBytes + p.stages * p.stage_bytes;
    if (need > dyn || (p.xch_off != 0 && static_cast<uint32_t>(p.xch_off) < need)) __trap();
  }
#endif
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
    if (p.slot_map) {
#pragma unroll
      for (int r = 0; r < 2; ++r) pre_ids[r] = static_cast<int32_t>(p.slot_map[pre_ids[r]]);
    }
    B200_CHECK_SLOT(pre_ids[0]);
    B200_CHECK_SLOT(pre_ids[1]);
  }
  // the rate this CTA recorded in earlier launches (updated at exit; fetched here so that the exit is a plain store)
  CtaStat st_prev = {0.f, 0u};
  if (tid == 0 && cta < kMaxCtas) st_prev = pv.cta_stat[cta];
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

  if (tid == 0) rq[0] = cta < n_static ? cta : take_ticket();
  __syncthreads();
  if (tid == 0) B200_TR(TR_BARS_READY);

  const int T = p.chunk_tokens;
  const uint32_t v_off = p.kv_split ? 0u : T * p.row_stride;  // V rows follow the K rows of a stage (or own the next slot)
  const int stage_step = 1 + p.kv_split;                      // producers: slots per chunk
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
    if (p.kv_split && which) stage = 1;   // K rows of chunk n in slot 2n, V rows in slot 2n + 1 (mod stages)
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
            if (p.slot_map) nxt[r] = static_cast<int32_t>(p.slot_map[nxt[r]]);
            B200_CHECK_SLOT(nxt[r]);
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
              if (p.slot_map) nxt[r] = static_cast<int32_t>(p.slot_map[nxt[r]]);
            B200_CHECK_SLOT(nxt[r]);
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
          stage += stage_step;
          if (stage >= p.stages) {
            stage -= p.stages;
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
          if (p.slot_map) nxt[r] = static_cast<int32_t>(p.slot_map[nxt[r]]);
            B200_CHECK_SLOT(nxt[r]);
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
            if (p.slot_map) nxt[r] = static_cast<int32_t>(p.slot_map[nxt[r]]);
            B200_CHECK_SLOT(nxt[r]);
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
        stage += stage_step;
        if (stage >= p.stages) {
          stage -= p.stages;
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
  // Transposed formulation: the TOKENS ride in the M dimension of the MMA and the q heads of the group in N.
  //   S^T[16 tok, 8 heads] = K[16 tok, D] . Q^T     A = K tile straight from smem (ldmatrix), B = q in registers
  //   O^T[D, 8 heads]     += V^T[D, 16 tok] . P^T   A = V tile through a transposing ldmatrix, B = P^T
  // Every accumulator row is a real token / head-dim index, so a group of <= 8 q heads wastes nothing (with the
  // heads in M, as in the reference's Triton kernel, half of every m16 tile and of its accumulator registers was
  // padding: 96 HMMA + ~80 zeroing / move instructions per 16 tokens where 16 HMMA do).  P^T becomes the B
  // operand through movmatrix (8x8 b16 transpose, 2 per 16 tokens).  An fp8 pool costs one cvt.f16x2.e4m3x2 per
  // fragment register and nothing else: K bytes arrive in A-fragment order through the 16-bit ldmatrix, V bytes
  // through ldmatrix.m16n16.trans.b8 (tools/probe_frag.cu prints both layouts), with the tokens of a chunk assigned
  // to MMA rows / k slots by a fixed permutation so that no byte shuffle is needed and every 8-row phase of an
  // ldmatrix still reads 8 distinct bank groups:
  //   MMA row R <-> token pi(R) = R ^ ((R & 8) >> 1);   V matrix row i = 4a + 2b + e <-> token (8b + 2a + e) ^ 4b.
  // G in 9..16 (NH = 2) runs a second n8 tile of heads on the same A fragments.
  constexpr int NH = ROWS16 ? 2 : 1;   // n8 tiles of q heads
  constexpr int HPT = 2 * NH;          // heads a thread holds values of: 8 * (k >> 1) + 2 j + (k & 1)
  constexpr int MT = D / 16;           // m16 tiles of O^T = k16 steps of K . Q^T
  const int h = warp % p.hc;
  const int tsi = warp / p.hc;
  const int g = lane >> 2, j = lane & 3;
  const bool cap = p.logit_cap > 0.f;
  const float c = cap ? kLog2e : p.scale_log2;
  const float cap_pre = cap ? p.sm_scale_k / p.logit_cap : 0.f;
  const bool swz = p.swizzle != 0;
  const uint32_t head_off = h * D * EB;

  // Lane-constant parts of the ldmatrix addresses.  Padded rows (bulk-copy path): stride = row bytes + 16.
  // Swizzled rows (cp.async path): unit u of row r at u ^ (r & 7); a head starts on a 128-byte boundary, so the
  // swizzle only touches the low three bits of the unit index: bit 0 folds into the lane constant, bits 1-2 are
  // xor-ed into the (compile-time) unit-pair offset of each instruction.
  uint32_t k_base, k_x, v_base, v_x;
  {
    const int mi = lane >> 3, rr = lane & 7;
    const int R = (mi & 1) * 8 + rr;                         // MMA row this lane addresses (matrices 0/2: rows 0-7)
    const int tokK = FP8 ? (R ^ ((R & 8) >> 1)) : R;
    const int hiK = mi >> 1;                                 // second 16-byte unit of the pair
    const int kx = swz ? (tokK & 7) : 0;
    k_base = (tsi * 16 + tokK) * p.row_stride + head_off + ((hiK ^ (kx & 1)) << 4);
    k_x = (kx & 6) << 4;
    int tokV, hiV;
    if constexpr (FP8) {
      const int i = lane & 15;
      const int kap = 8 * ((i >> 1) & 1) + 2 * (i >> 2) + (i & 1);
      tokV = kap ^ (4 * ((i >> 1) & 1));
      hiV = lane >> 4;
    } else {
      tokV = (mi >> 1) * 8 + rr;
      hiV = mi & 1;
    }
    const int vx = swz ? (tokV & 7) : 0;
    v_base = (tsi * 16 + tokV) * p.row_stride + head_off + v_off + ((hiV ^ (vx & 1)) << 4);
    v_x = (vx & 6) << 4;
  }
  // tokens (inside the warp's 16) behind this thread's S^T elements and behind the k slots of its V fragments
  const int tok_hi = FP8 ? 8 + (g ^ 4) : 8 + g;               // elements 2, 3 (elements 0, 1: token g)
  const int vt0 = 2 * j, vt1 = 2 * j + 1;                     // k slots 2j, 2j + 1
  const int vt2 = FP8 ? 8 + ((2 * j) ^ 4) : 8 + 2 * j;        // k slots 2j + 8, 2j + 9
  const int vt3 = FP8 ? 8 + ((2 * j + 1) ^ 4) : 9 + 2 * j;

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
  // chunk's critical path.
  const long long clk_first = clock64();
  long long clk_last = clk_first;
  int n_consumed = 0;
  int npiece = 0;     // pieces seen by this CTA: parity picks the flag slot
  while (rg < R) {
  PieceWalk walk(rn == 0 && rg == cta ? d0 : load_desc(pv.range_desc + rg), p.ng);
  while (walk.next(pv, b, hg, lc0, lc1, nchunks, vreq, S, row_off)) {
    const int hq0 = (hg * p.hc + h) * p.group;
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
    const int lc_pf = lc1 - 4 > lc0 ? lc1 - 4 : lc0;   // where the other pieces' partial rows are pulled into L2
    // heads this thread holds values of, and whether they exist in the group
    int hl[HPT];
    bool hv[HPT];
#pragma unroll
    for (int k = 0; k < HPT; ++k) {
      hl[k] = 8 * (k >> 1) + 2 * j + (k & 1);
      hv[k] = hl[k] < p.group;
    }

    // ---- Q^T fragments (B operand), loaded once per piece: lane (g, j) holds head g (+ 8) ----
    uint32_t qb[NH][MT][2];
#pragma unroll
    for (int nh = 0; nh < NH; ++nh) {
      const bool qv = 8 * nh + g < p.group;
      const QT* qp = reinterpret_cast<const QT*>(p.q) + b * p.q_sb + (hq0 + 8 * nh + g) * p.q_sh;
#pragma unroll
      for (int ks = 0; ks < MT; ++ks) {
        if constexpr (!FP8) {
          const int dd = ks * 16 + 2 * j;
          qb[nh][ks][0] = qv ? *reinterpret_cast<const uint32_t*>(qp + dd) : 0u;
          qb[nh][ks][1] = qv ? *reinterpret_cast<const uint32_t*>(qp + dd + 8) : 0u;
        } else {
          // fp8 K: k slots (2j, 2j+1 | 2j+8, 2j+9) of step ks <-> d = 16 ks + 4j + (0, 1 | 2, 3)
          const int dd = ks * 16 + 4 * j;
          float f[4] = {0.f, 0.f, 0.f, 0.f};
          if (qv) {
#pragma unroll
            for (int e = 0; e < 4; ++e) f[e] = fminf(fmaxf(q_to_f32(qp[dd + e]), -65504.f), 65504.f);
          }
          qb[nh][ks][0] = pack2<true>(f[0], f[1]);
          qb[nh][ks][1] = pack2<true>(f[2], f[3]);
        }
      }
    }

    float o[NH][MT][4];
#pragma unroll
    for (int nh = 0; nh < NH; ++nh)
#pragma unroll
      for (int t = 0; t < MT; ++t)
#pragma unroll
        for (int e = 0; e < 4; ++e) o[nh][t][e] = 0.f;
    float m[HPT], mc[HPT], l[HPT];   // reference (a running maximum, in logit units), the same times c, row sum
#pragma unroll
    for (int k = 0; k < HPT; ++k) {
      m[k] = -CUDART_INF_F;
      mc[k] = 0.f;
      l[k] = 0.f;
    }

#ifdef B200ATTN_DEC_TRACE
    if (tid == 0) { B200_TR_ONCE(TR_C_Q, tr_a); ++tr_pieces; }
#endif
    for (int lc = lc0; lc < lc1; ++lc) {
      if (lc == lc_pf && try_early) {
        // A few chunks before the end of a cut request's piece, the other pieces' partial rows of this warp's
        // head are pulled into L2: the CTA that merges (normally this one: its neighbour finished the other half
        // long ago, and ~1 GB has gone through L2 since) then reads them in one short round trip instead of a
        // DRAM access queued behind the KV stream.  A row not written yet is harmless to prefetch.
        const int line = lane & 3;
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
      if (lc == lc1 - 1 && try_early && tid == 0) {   // (warp-uniform tests first: two instructions per chunk)
        // Last-arriver shortcut, decided one chunk early (the load's round trip hides behind the chunk): if
        // every other piece of the request is already in the workspace, this CTA is the one that merges, so
        // its own partial never needs to go through global memory (see the epilogue).
        int arrived;
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(arrived) : "l"(pv.vreq_arrived + vreq) : "memory");
        early_ok = arrived == v_pieces - 1;
      }
      B200_TR_DEP(TR_K_WAIT, 0.f);
#ifdef B200ATTN_DEC_TRACE
      if (p.trace && tid == 0 && tr_chunks == 65) p.trace[blockIdx.x * kTraceSlots + TR_K_WAIT2] = clock64();
#endif
      mbar_wait(bar_full + stage * 8, phase);
      if (tid == 0) B200_TR_ONCE(TR_C_FIRST_FULL, tr_b);
      B200_TR_DEP(TR_K_FULL, 0.f);
      const uint32_t sbase = stage0 + stage * p.stage_bytes;

      // A chunk holds SUBS sub-chunks of 16 tokens per warp (fp8 pools with <= 4 token slices: 2, so that a slot
      // is as many BYTES as a 16-bit one and the per-slot costs -- barrier wait, arrive, index loads -- are paid
      // once per 64 KB, and the two sub-chunks' S / softmax / P.V chains interleave in one basic block).
      // Sub-chunk u of warp (h, slice): stage rows (u * ts + slice) * 16 ...
      auto chunk_body = [&](auto subs_tag, auto e5_tag) {
      constexpr int SUBS = decltype(subs_tag)::value;
      constexpr bool E5M2 = decltype(e5_tag)::value;   // (fp8 pools only)
      const uint32_t sub_stride = static_cast<uint32_t>(p.ts) * 16u * p.row_stride;

      // ---- S^T = K . Q^T : 16 tokens x 8 (16) heads per sub-chunk ----
      // (two accumulators per tile, even / odd k steps: the dependent HMMA chain is half as long)
      float s[SUBS][NH][4], sb[SUBS][NH][4];
#pragma unroll
      for (int u = 0; u < SUBS; ++u)
#pragma unroll
        for (int nh = 0; nh < NH; ++nh) {
          s[u][nh][0] = s[u][nh][1] = s[u][nh][2] = s[u][nh][3] = 0.f;
          sb[u][nh][0] = sb[u][nh][1] = sb[u][nh][2] = sb[u][nh][3] = 0.f;
        }
#pragma unroll
      for (int u = 0; u < SUBS; ++u) {
        const uint32_t kb = sbase + k_base + u * sub_stride;
        if constexpr (!FP8) {
#pragma unroll
          for (int ks = 0; ks < MT; ++ks) {
            uint32_t a[4];
            ldsm_x4(kb + ((32u * ks) ^ k_x), a);
#pragma unroll
            for (int nh = 0; nh < NH; ++nh)
              mma16816<HALF>((ks & 1) ? sb[u][nh] : s[u][nh], a[0], a[1], a[2], a[3], qb[nh][ks][0], qb[nh][ks][1]);
          }
        } else {
#pragma unroll
          for (int u2 = 0; u2 < MT / 2; ++u2) {
            uint32_t kf[4], a[4];
            ldsm_x4(kb + ((32u * u2) ^ k_x), kf);
            fp8x4_to_f16x2x2<E5M2>(kf[0], a[0], a[2]);
            fp8x4_to_f16x2x2<E5M2>(kf[1], a[1], a[3]);
#pragma unroll
            for (int nh = 0; nh < NH; ++nh)
              mma16816<true>(s[u][nh], a[0], a[1], a[2], a[3], qb[nh][2 * u2][0], qb[nh][2 * u2][1]);
            fp8x4_to_f16x2x2<E5M2>(kf[2], a[0], a[2]);
            fp8x4_to_f16x2x2<E5M2>(kf[3], a[1], a[3]);
#pragma unroll
            for (int nh = 0; nh < NH; ++nh)
              mma16816<true>(sb[u][nh], a[0], a[1], a[2], a[3], qb[nh][2 * u2 + 1][0], qb[nh][2 * u2 + 1][1]);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < SUBS; ++u)
#pragma unroll
        for (int nh = 0; nh < NH; ++nh)
#pragma unroll
          for (int e = 0; e < 4; ++e) s[u][nh][e] += sb[u][nh][e];

      B200_TR_DEP(TR_K_QK, s[0][0][0]);
      // ---- logit cap, tail mask ----
      if (cap) {
#pragma unroll
        for (int u = 0; u < SUBS; ++u)
#pragma unroll
          for (int nh = 0; nh < NH; ++nh)
#pragma unroll
            for (int e = 0; e < 4; ++e) s[u][nh][e] = p.logit_cap * tanhf(s[u][nh][e] * cap_pre);
      }
      const int t0 = lc * T + tsi * 16;   // token of the warp's row 0 in sub-chunk 0 (sub-chunk u: + u * ts * 16)
      const bool tail = (lc + 1) * T > S;
      if (tail) {
#pragma unroll
        for (int u = 0; u < SUBS; ++u)
#pragma unroll
          for (int nh = 0; nh < NH; ++nh) {
            if (t0 + u * p.ts * 16 + g >= S) s[u][nh][0] = s[u][nh][1] = -CUDART_INF_F;
            if (t0 + u * p.ts * 16 + tok_hi >= S) s[u][nh][2] = s[u][nh][3] = -CUDART_INF_F;
          }
      }

      // ---- online softmax over the token dimension, with a lazily updated reference ----
      // P = exp2(s c - ref) against the reference the piece already has; the true running maximum (three
      // dependent shuffle rounds per head, then the rescale factors -- 400 of the 1250 clocks a warp spent on a
      // chunk, tools/trace_decode.py) is only taken on a piece's first chunk and when some logit rises more than
      // 2^8 above the reference (P <= 256 fits f16 / bf16 with the same relative precision; l and O are fp32).
      // Which path ran does not change the mathematics: o = sum P v / sum P and lse = ref + log sum P for any ref.
      float pr[SUBS][NH][4];
      bool far = lc == lc0;
#pragma unroll
      for (int u = 0; u < SUBS; ++u)
#pragma unroll
        for (int nh = 0; nh < NH; ++nh)
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            pr[u][nh][e] = fmaf(s[u][nh][e], c, -mc[2 * nh + (e & 1)]);
            far |= pr[u][nh][e] > 8.f;
          }
      if (__any_sync(0xffffffffu, far)) {
#pragma unroll
        for (int k = 0; k < HPT; ++k) {
          const int nh = k >> 1, e = k & 1;
          float mx = fmaxf(s[0][nh][e], s[0][nh][e + 2]);
#pragma unroll
          for (int u = 1; u < SUBS; ++u) mx = fmaxf(mx, fmaxf(s[u][nh][e], s[u][nh][e + 2]));
          mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
          mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
          mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
          const float mn = fmaxf(m[k], mx);
          const float mr = mn == -CUDART_INF_F ? 0.f : mn;
          const float corr = fast_exp2((m[k] - mr) * c);
          m[k] = mn;
          mc[k] = mr * c;
          l[k] *= corr;
#pragma unroll
          for (int u = 0; u < SUBS; ++u) {
            pr[u][nh][e] = fmaf(s[u][nh][e], c, -mc[k]);
            pr[u][nh][e + 2] = fmaf(s[u][nh][e + 2], c, -mc[k]);
          }
#pragma unroll
          for (int t = 0; t < MT; ++t) {
            o[nh][t][e] *= corr;
            o[nh][t][e + 2] *= corr;
          }
        }
      }
#pragma unroll
      for (int u = 0; u < SUBS; ++u)
#pragma unroll
        for (int k = 0; k < HPT; ++k) {
          const int nh = k >> 1, e = k & 1;
          pr[u][nh][e] = fast_exp2(pr[u][nh][e]);
          pr[u][nh][e + 2] = fast_exp2(pr[u][nh][e + 2]);
          l[k] += pr[u][nh][e] + pr[u][nh][e + 2];
        }

      // ---- P^T (B operand): pack [token row][heads 2j, 2j+1], transpose 8x8 -> [head g][k slots 2j, 2j+1] ----
      uint32_t pb[SUBS][NH][2];
#pragma unroll
      for (int u = 0; u < SUBS; ++u)
#pragma unroll
        for (int nh = 0; nh < NH; ++nh) {
          pb[u][nh][0] = movmatrix_trans(pack2<HALF>(pr[u][nh][0], pr[u][nh][1]));
          pb[u][nh][1] = movmatrix_trans(pack2<HALF>(pr[u][nh][2], pr[u][nh][3]));
        }
      B200_TR_DEP(TR_K_SOFTMAX, __uint_as_float(pb[0][0][1]));
      // V fragments of masked tail tokens are zeroed: the smem rows behind them hold stale bytes of another
      // request (0 * inf would poison this one).  Only the tail chunk pays the masks.
      uint32_t vm_lo[SUBS], vm_hi[SUBS];
#pragma unroll
      for (int u = 0; u < SUBS; ++u) {
        vm_lo[u] = vm_hi[u] = 0xffffffffu;
        if (tail) {
          const int tu = t0 + u * p.ts * 16;
          vm_lo[u] = (tu + vt0 < S ? 0x0000ffffu : 0u) | (tu + vt1 < S ? 0xffff0000u : 0u);
          vm_hi[u] = (tu + vt2 < S ? 0x0000ffffu : 0u) | (tu + vt3 < S ? 0xffff0000u : 0u);
        }
      }

      // K and V in separate slots: the K slot is done with (its fragments are in registers), the V rows of the
      // chunk are in the next slot
      uint32_t sbase_v = sbase;
      if (p.kv_split) {
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_empty + stage * 8);
        if (++stage == p.stages) {
          stage = 0;
          phase ^= 1;
        }
        mbar_wait(bar_full + stage * 8, phase);
        sbase_v = stage0 + stage * p.stage_bytes;
      }

      // ---- O^T += V^T . P^T ----
      auto pv_mma = [&](auto tail_tag) {
        constexpr bool TAIL = decltype(tail_tag)::value;
#pragma unroll
        for (int u = 0; u < SUBS; ++u) {
          const uint32_t mlo = TAIL ? vm_lo[u] : 0xffffffffu, mhi = TAIL ? vm_hi[u] : 0xffffffffu;
          const uint32_t vb = sbase_v + v_base + u * sub_stride;
          if constexpr (!FP8) {
#pragma unroll
            for (int t = 0; t < MT; ++t) {
              uint32_t a[4];
              ldsm_x4_trans(vb + ((32u * t) ^ v_x), a);
#pragma unroll
              for (int nh = 0; nh < NH; ++nh)
                mma16816<HALF>(o[nh][t], a[0] & mlo, a[1] & mlo, a[2] & mhi, a[3] & mhi, pb[u][nh][0], pb[u][nh][1]);
            }
          } else {
#pragma unroll
            for (int t2 = 0; t2 < MT / 2; ++t2) {
              uint32_t vf[4], a[4];
              ldsm8_x2_trans(vb + ((32u * t2) ^ v_x), vf);
              fp8x4_to_f16x2x2<E5M2>(vf[0], a[0], a[2]);
              fp8x4_to_f16x2x2<E5M2>(vf[1], a[1], a[3]);
#pragma unroll
              for (int nh = 0; nh < NH; ++nh)
                mma16816<true>(o[nh][2 * t2], a[0] & mlo, a[1] & mlo, a[2] & mhi, a[3] & mhi, pb[u][nh][0], pb[u][nh][1]);
              fp8x4_to_f16x2x2<E5M2>(vf[2], a[0], a[2]);
              fp8x4_to_f16x2x2<E5M2>(vf[3], a[1], a[3]);
#pragma unroll
              for (int nh = 0; nh < NH; ++nh)
                mma16816<true>(o[nh][2 * t2 + 1], a[0] & mlo, a[1] & mlo, a[2] & mhi, a[3] & mhi, pb[u][nh][0], pb[u][nh][1]);
            }
          }
        }
      };
      if (tail) pv_mma(std::true_type{}); else pv_mma(std::false_type{});
      B200_TR_DEP(TR_K_PV, o[0][MT - 1][0]);
      };   // chunk_body
      if constexpr (FP8) {
        if (p.e5m2) {
          if (p.subs == 2) chunk_body(std::integral_constant<int, 2>{}, std::true_type{});
          else chunk_body(std::integral_constant<int, 1>{}, std::true_type{});
        } else {
          if (p.subs == 2) chunk_body(std::integral_constant<int, 2>{}, std::false_type{});
          else chunk_body(std::integral_constant<int, 1>{}, std::false_type{});
        }
      } else {
        chunk_body(std::integral_constant<int, 1>{}, std::false_type{});
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_empty + stage * 8);
      B200_TR_DEP(TR_K_END, 0.f);
#ifdef B200ATTN_DEC_TRACE
      // (the last chunk's stamp stays in a register until the piece ends: a global store per chunk cost the traced
      // kernel 600 clocks per chunk on the consumer-bound shapes)
      if (tid == 0) { B200_TR_ONCE(TR_C_FIRST_DONE, tr_c); tr_last = clock64(); ++tr_chunks; }
#endif
      if (++stage == p.stages) {
        stage = 0;
        phase ^= 1;
      }
    }

#ifdef B200ATTN_DEC_TRACE
    if (tid == 0 && p.trace) p.trace[blockIdx.x * kTraceSlots + TR_C_LAST_DONE] = tr_last;
#endif
    // ---- epilogue: normalise, write final o or a partial ----
    // This thread holds, for heads hl[k], the head-dim rows d = 16 t + g + 8 (e >> 1): element (nh, t, e) belongs to
    // head k = 2 nh + (e & 1).  A head's row sum is spread over the 8 token lanes (g).
    float inv[HPT], lse[HPT];
#pragma unroll
    for (int k = 0; k < HPT; ++k) {
      l[k] += __shfl_xor_sync(0xffffffffu, l[k], 4);
      l[k] += __shfl_xor_sync(0xffffffffu, l[k], 8);
      l[k] += __shfl_xor_sync(0xffffffffu, l[k], 16);
      inv[k] = l[k] > 0.f ? p.v_scale / l[k] : 0.f;
      lse[k] = l[k] > 0.f ? m[k] * c * kLn2 + logf(l[k]) : -CUDART_INF_F;
    }
    // Workspace rows of this piece: ((vreq + rg) * ts + slice) * hqv + head; with the in-CTA slice merge below
    // only slice 0's rows are used (one partial per piece instead of ts).
    const int64_t piece_row = static_cast<int64_t>(vreq + rg) * p.ts * hqv;
    const bool use_xch = NH == 1 && p.xch_off != 0;   // implies ts > 1 (make_geometry)

    // Last-arriver shortcut (partial pieces): when every other piece of the request had already arrived one
    // chunk ago, this CTA is the one that merges -- so it merges its own partial straight from registers with
    // the others read from the workspace, instead of write partial -> fence -> atomic -> read everything back.
    // Same values, same order, same arithmetic as merge_heads4, so which path ran never shows in the result.
    const bool can_fast = NH == 1 && !full_req && p.fuse_merge && (use_xch || p.ts == 1);
    volatile int32_t* flags = merge_flag + (npiece & 1) * 2;
    if (can_fast && tid == 0) flags[0] = early_ok;
    if (can_fast && !use_xch) named_bar_sync(2, kConsumerWarps * 32);   // (the slice exchange has its own barrier)
    // The merge's loads go out in ONE batch -- the other pieces' lse and the rows of the first other piece (the
    // only other piece in all but the rarest plans) -- before anything waits on them.
    const int i1 = me == 0 ? 1 : 0;   // the first other piece
    float wk[2][4], inv_s[2] = {0.f, 0.f}, lse_tot[2] = {-CUDART_INF_F, -CUDART_INF_F};
    auto merge_lse_load = [&](int rowh, float lse_mine, float (&lv)[4]) {
#pragma unroll
      for (int i = 0; i < 4; ++i)
        lv[i] = i < v_pieces ? (i == me ? lse_mine : __ldcg(p.ws_lse + prow0 + i * pstep + rowh)) : -CUDART_INF_F;
    };
    auto merge_weights = [&](int k, const float (&lv)[4]) {
      const float mm = fmaxf(fmaxf(lv[0], lv[1]), fmaxf(lv[2], lv[3]));
      float sum = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        wk[k][i] = mm != -CUDART_INF_F ? __expf(lv[i] - mm) : 0.f;
        sum += wk[k][i];
      }
      inv_s[k] = sum > 0.f ? 1.f / sum : 0.f;
      lse_tot[k] = sum > 0.f ? mm + logf(sum) : -CUDART_INF_F;
    };
    // own = this piece's normalised value, x1 = the first other piece's (already in a register)
    auto merge_val = [&](int k, int rowh, int d, float own, float x1) -> float {
      float a = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (i < v_pieces) {
          float v = own;
          if (i == i1) v = x1;
          else if (i != me) v = __ldcg(p.ws_o + (prow0 + i * pstep + rowh) * D + d);
          a = fmaf(wk[k][i], v, a);
        }
      }
      return a * inv_s[k];
    };
    QT* const obase = reinterpret_cast<QT*>(p.o) + b * p.o_sb + hq0 * p.o_sh;
    float* const lbase = p.lse ? p.lse + static_cast<int64_t>(b) * p.hq + hq0 : nullptr;

    bool final_written = false;
    if (p.ts == 1 && full_req) {
      final_written = true;
#pragma unroll
      for (int nh = 0; nh < NH; ++nh)
#pragma unroll
        for (int t = 0; t < MT; ++t)
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int k = 2 * nh + (e & 1);
            if (hv[k]) store_q(obase + hl[k] * p.o_sh + 16 * t + g + 8 * (e >> 1), o[nh][t][e] * inv[k]);
          }
      if (lbase && g == 0) {
#pragma unroll
        for (int k = 0; k < HPT; ++k)
          if (hv[k]) lbase[hl[k]] = lse[k];
      }
    } else if (use_xch) {
      // The ts token slices of this piece meet in shared memory (unnormalised O, max, sum) and each warp
      // finishes the m tiles t = slice, slice + ts, ... of its head: a whole request goes straight to `o`; a piece
      // of a request cut by a range boundary leaves ONE normalised partial per head in the workspace instead of ts.
      float* xall = reinterpret_cast<float*>(smem + p.xch_off);                    // [warp][group][D]
      float2* xml = reinterpret_cast<float2*>(xall + kXchWarps * p.group * D);     // [warp][8] (max * c, sum)
#pragma unroll
      for (int t = 0; t < MT; ++t)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int k = e & 1;
          if (hv[k]) xall[(warp * p.group + hl[k]) * D + 16 * t + g + 8 * (e >> 1)] = o[0][t][e];
        }
      if (g == 0) {
#pragma unroll
        for (int k = 0; k < 2; ++k)
          if (hv[k]) xml[warp * 8 + hl[k]] = make_float2(l[k] > 0.f ? m[k] * c : -CUDART_INF_F, l[k]);
      }
      named_bar_sync(2, kConsumerWarps * 32);
      const bool fast = can_fast && flags[0] != 0;
      final_written = full_req || fast;
      float wgt[kXchWarps][2], invx[2], lse_all[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        float mm = -CUDART_INF_F;
        const int hk = hv[k] ? hl[k] : 0;
#pragma unroll
        for (int sl = 0; sl < kXchWarps; ++sl)
          if (sl < p.ts) mm = fmaxf(mm, xml[(sl * p.hc + h) * 8 + hk].x);
        float L = 0.f;
#pragma unroll
        for (int sl = 0; sl < kXchWarps; ++sl) {
          wgt[sl][k] = 0.f;
          if (sl < p.ts) {
            const float2 ml = xml[(sl * p.hc + h) * 8 + hk];
            wgt[sl][k] = ml.y > 0.f ? fast_exp2(ml.x - mm) : 0.f;
            L = fmaf(wgt[sl][k], ml.y, L);
          }
        }
        invx[k] = L > 0.f ? p.v_scale / L : 0.f;
        lse_all[k] = L > 0.f ? mm * kLn2 + logf(L) : -CUDART_INF_F;
      }
      int rowh[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) rowh[k] = h * p.group + (hv[k] ? hl[k] : 0);
      // this warp finishes tiles tsi, tsi + ts, ...: at most ceil(MT / 2) of them (ts >= 2)
      float x1[(MT + 1) / 2][4];
      if (fast) {
        float lv[2][4];
#pragma unroll
        for (int k = 0; k < 2; ++k) merge_lse_load(rowh[k], lse_all[k], lv[k]);
#pragma unroll
        for (int tt = 0; tt < (MT + 1) / 2; ++tt) {
          const int t = tsi + tt * p.ts;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            x1[tt][e] = 0.f;
            if (t < MT && hv[e & 1])
              x1[tt][e] = __ldcg(p.ws_o + (prow0 + i1 * pstep + rowh[e & 1]) * D + 16 * t + g + 8 * (e >> 1));
          }
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) merge_weights(k, lv[k]);
      }
#pragma unroll
      for (int tt = 0; tt < (MT + 1) / 2; ++tt) {
        const int t = tsi + tt * p.ts;
        if (t < MT) {
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int k = e & 1;
            if (!hv[k]) continue;
            const int d = 16 * t + g + 8 * (e >> 1);
            float a = 0.f;
#pragma unroll
            for (int sl = 0; sl < kXchWarps; ++sl)
              if (sl < p.ts) a = fmaf(wgt[sl][k], xall[((sl * p.hc + h) * p.group + hl[k]) * D + d], a);
            a *= invx[k];
            if (full_req) {
              store_q(obase + hl[k] * p.o_sh + d, a);
            } else if (fast) {
              store_q(obase + hl[k] * p.o_sh + d, merge_val(k, rowh[k], d, a, x1[tt][e]));
            } else {
              p.ws_o[(piece_row + rowh[k]) * D + d] = a;
            }
          }
        }
      }
      if (g == 0 && tsi == 0) {
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          if (!hv[k]) continue;
          if (full_req) {
            if (lbase) lbase[hl[k]] = lse_all[k];
          } else if (fast) {
            if (lbase) lbase[hl[k]] = lse_tot[k];
          } else {
            p.ws_lse[piece_row + rowh[k]] = lse_all[k];
          }
        }
      }
      // the closing barrier (the area is free for the next piece) is the one in front of the arrival counter
      // below when the partial goes through the workspace
      if (final_written || !p.fuse_merge) named_bar_sync(2, kConsumerWarps * 32);
    } else if (can_fast && flags[0] != 0) {
      // ts == 1: this warp owns its head's rows outright
      final_written = true;
      int rowh[2];
      float lv[2][4], x1[MT][4];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        rowh[k] = h * p.group + (hv[k] ? hl[k] : 0);
        merge_lse_load(rowh[k], lse[k], lv[k]);
      }
#pragma unroll
      for (int t = 0; t < MT; ++t)
#pragma unroll
        for (int e = 0; e < 4; ++e)
          x1[t][e] = hv[e & 1] ? __ldcg(p.ws_o + (prow0 + i1 * pstep + rowh[e & 1]) * D + 16 * t + g + 8 * (e >> 1)) : 0.f;
#pragma unroll
      for (int k = 0; k < 2; ++k) merge_weights(k, lv[k]);
#pragma unroll
      for (int t = 0; t < MT; ++t)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int k = e & 1;
          const int d = 16 * t + g + 8 * (e >> 1);
          if (hv[k]) store_q(obase + hl[k] * p.o_sh + d, merge_val(k, rowh[k], d, o[0][t][e] * inv[k], x1[t][e]));
        }
      if (lbase && g == 0) {
#pragma unroll
        for (int k = 0; k < 2; ++k)
          if (hv[k]) lbase[hl[k]] = lse_tot[k];
      }
    } else {
#pragma unroll
      for (int nh = 0; nh < NH; ++nh)
#pragma unroll
        for (int t = 0; t < MT; ++t)
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int k = 2 * nh + (e & 1);
            if (hv[k])
              p.ws_o[(piece_row + static_cast<int64_t>(tsi) * hqv + h * p.group + hl[k]) * D + 16 * t + g + 8 * (e >> 1)] =
                  o[nh][t][e] * inv[k];
          }
      if (g == 0) {
#pragma unroll
        for (int k = 0; k < HPT; ++k)
          if (hv[k]) p.ws_lse[piece_row + static_cast<int64_t>(tsi) * hqv + h * p.group + hl[k]] = lse[k];
      }
    }
    if (final_written && !full_req && tid == 0) {
      pv.vreq_arrived[vreq] = 0;   // shortcut taken: the counter is ready for the next layer's launch
      B200_TR(TR_C_MERGED);
    }
    if (!final_written && p.fuse_merge) {
      // Stage 2 without a second launch: the CTA that lands a virtual request's last piece
      // merges all of them (threadfence + counter, the "last block done" pattern).  The
      // producers keep streaming the next piece's chunks meanwhile.
      __threadfence();
      named_bar_sync(2, kConsumerWarps * 32);
      if (tid == 0) {
        const int arrived = atomicAdd(&pv.vreq_arrived[vreq], 1) + 1;
        const int last = arrived == v_pieces;
        if (last) pv.vreq_arrived[vreq] = 0;  // ready for the next layer's launch
        flags[1] = last;
      }
      named_bar_sync(2, kConsumerWarps * 32);
      if (flags[1]) {
        if (tid == 0) B200_TR(TR_C_MERGED);
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
    if (tid == 0) B200_TR(TR_C_EPILOGUE);
    if (rn == 0) n_consumed += lc1 - lc0;
  }
  if (rn == 0) clk_last = clock64();
  ++rn;
  rg = rq[rn & (kRangeRing - 1)];
  }
  if (tid == 0 && n_consumed >= 16 && cta < kMaxCtas) {
    const float r = static_cast<float>(clk_last - clk_first) / static_cast<float>(n_consumed);
    CtaStat st = st_prev;
    const bool ok = st.tag == (kStatTag ^ static_cast<uint32_t>(cta)) && st.clk_per_chunk > 0.f;
    st.clk_per_chunk = ok ? 0.75f * st.clk_per_chunk + 0.25f * r : r;
    st.tag = kStatTag ^ static_cast<uint32_t>(cta);
    pv.cta_stat[cta] = st;
  }
#ifdef B200ATTN_DEC_TRACE
  if (tid == 0 && p.trace) {
    B200_TR(TR_C_EXIT);
    p.trace[blockIdx.x * kTraceSlots + TR_COUNTS] = tr_chunks | (tr_pieces << 32);
    B200_TR_GT(TR_GT_EXIT);
  }
#endif
}

// One translation unit per (q dtype, fp8 KV, head dim) -- decode_inst.cu -- defines launch_decode_d for its
// triple; launch_decode (api.cu) picks by head dim.  fp8 pools: head dims 64 / 128 / 256; 16-bit: also 80 / 96.
template <typename QT, bool FP8, int D>
cudaError_t launch_decode_d(const DecodeParams& p, bool rows16, int grid, int smem, cudaStream_t stream);

#define B200ATTN_DECLARE_LAUNCH_D(QT, FP8, D) \
  template <>                                 \
  cudaError_t launch_decode_d<QT, FP8, D>(const DecodeParams& p, bool rows16, int grid, int smem, cudaStream_t stream);
#define B200ATTN_DECLARE_LAUNCH_QT(QT)       \
  B200ATTN_DECLARE_LAUNCH_D(QT, false, 64)   \
  B200ATTN_DECLARE_LAUNCH_D(QT, false, 80)   \
  B200ATTN_DECLARE_LAUNCH_D(QT, false, 96)   \
  B200ATTN_DECLARE_LAUNCH_D(QT, false, 128)  \
  B200ATTN_DECLARE_LAUNCH_D(QT, false, 256)  \
  B200ATTN_DECLARE_LAUNCH_D(QT, true, 64)    \
  B200ATTN_DECLARE_LAUNCH_D(QT, true, 128)   \
  B200ATTN_DECLARE_LAUNCH_D(QT, true, 256)
B200ATTN_DECLARE_LAUNCH_QT(__nv_bfloat16)
B200ATTN_DECLARE_LAUNCH_QT(__half)

template <typename QT, bool FP8>
inline cudaError_t launch_decode(const DecodeParams& p, int head_dim, bool rows16, int grid, int smem, cudaStream_t stream) {
  switch (head_dim) {
    case 64: return launch_decode_d<QT, FP8, 64>(p, rows16, grid, smem, stream);
    case 128: return launch_decode_d<QT, FP8, 128>(p, rows16, grid, smem, stream);
    case 256: return launch_decode_d<QT, FP8, 256>(p, rows16, grid, smem, stream);
    default: break;
  }
  if constexpr (!FP8) {
    if (head_dim == 80) return launch_decode_d<QT, FP8, 80>(p, rows16, grid, smem, stream);
    if (head_dim == 96) return launch_decode_d<QT, FP8, 96>(p, rows16, grid, smem, stream);
  }
  return cudaErrorInvalidValue;
}

// The dynamic shared-memory limit is a per-function, per-device attribute: set on first use, not per launch.
inline cudaError_t ensure_smem_attr(const void* fn, int smem, int* cache /* [64] */) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = -1;
  if (dev >= 0 && cache[dev] >= smem) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e == cudaSuccess && dev >= 0) cache[dev] = smem;
  return e;
}

#define B200ATTN_DEFINE_LAUNCH_D(QT, FP8, D)                                                               \
  template <>                                                                                              \
  cudaError_t launch_decode_d<QT, FP8, D>(const DecodeParams& p, bool rows16, int grid, int smem,          \
                                          cudaStream_t stream) {                                           \
    static int attr_cache[2][64];                                                                          \
    void (*fn)(const DecodeParams) = rows16 ? decode_kernel<QT, FP8, D, true> : decode_kernel<QT, FP8, D, false>; \
    cudaError_t e = ensure_smem_attr(reinterpret_cast<const void*>(fn), smem, attr_cache[rows16]);         \
    if (e != cudaSuccess) return e;                                                                        \
    cudaLaunchConfig_t cfg = {};                                                                           \
    cfg.gridDim = dim3(grid);                                                                              \
    cfg.blockDim = dim3(kThreads);                                                                         \
    cfg.dynamicSmemBytes = smem;                                                                           \
    cfg.stream = stream;                                                                                   \
    cudaLaunchAttribute at[1];                                                                             \
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                                         \
    at[0].val.programmaticStreamSerializationAllowed = 1;                                                  \
    cfg.attrs = at;                                                                                        \
    cfg.numAttrs = no_pdl() ? 0 : 1;                                                                       \
    return cudaLaunchKernelEx(&cfg, fn, p);                                                                \
  }

}  // namespace b200attn
