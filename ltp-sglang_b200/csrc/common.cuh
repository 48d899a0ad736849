// Shared host/device definitions of the b200attn library: decode geometry,
// the plan buffer layout, and small PTX wrappers (mbarrier, bulk copy,
// ldmatrix, mma).  sm_100a only.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdlib>

#include "b200attn.h"

namespace b200attn {

// ---------------------------------------------------------------------------
// Decode geometry.  One persistent CTA per SM: 8 consumer warps + 4 producer
// warps.  A CTA serves `hc` kv heads of one "virtual request" (request x head
// group); consumer warp w = (head w % hc, token slice w / hc).  A chunk is
// 16 * ts tokens; each token row of the chunk (hc heads, contiguous in the
// pool) is moved HBM -> smem by one bulk async copy (UBLKCP) into a row padded
// by 16 B so ldmatrix over 8 consecutive tokens is bank-conflict free.
// ---------------------------------------------------------------------------
constexpr int kConsumerWarps = 8;
// Bulk copies take uniform-register operands, so a warp issues its lanes' copies one by one
// (~45 clk each, ncu r01): four producer warps (K/V x even/odd tokens) keep the issue rate
// above what HBM can deliver.  384 threads cost no registers over 288 (same allocation step).
constexpr int kProducerWarps = 4;
constexpr int kThreads = (kConsumerWarps + kProducerWarps) * 32;
constexpr int kRowPad = 16;           // bytes
constexpr int kBarrierBytes = 256;    // mbarrier area at the start of smem
constexpr int kMaxStages = 8;
constexpr int kXchWarps = 8;           // = kConsumerWarps: warps that exchange their slice state
constexpr int kRangeRing = 16;        // range ids in flight between producers and consumers (> kMaxStages)
constexpr int kRangeRingOff = 128;    // byte offsets inside the barrier area
constexpr int kMergeFlagOff = 192;
constexpr int kMaxSmem = 227 * 1024;
constexpr int kMaxRanges = 1024;     // static + dynamic chunk ranges of one plan
constexpr int kMaxHeadGroups = 16;   // num_kv_heads / heads-per-CTA
constexpr int kMaxRowBytes = 2048;    // per token row per tensor staged by one CTA

struct Geometry {
  int hc;            // kv heads per CTA
  int ts;            // token slices (consumer warps per head)
  int ng;            // head groups = num_kv_heads / hc
  int group;         // q heads per kv head
  int chunk_tokens;  // 16 * ts
  int row_bytes;     // hc * head_dim * elem
  int row_stride;    // row_bytes + pad, or row_bytes (swizzled rows)
  int stage_bytes;   // K + V
  int stages;
  int smem_bytes;
  int cp_async;      // rows staged by 16-byte cp.async (<= 512 B) instead of one bulk copy per row
  int swizzle;       // cp.async path: unpadded rows, 16-byte unit u of row r at u ^ (r & 7)
  int xch_off;       // byte offset of the slice-exchange area (whole requests merge their ts token slices
                     // inside the CTA), 0 = none
  int subs;          // 16-token sub-chunks per consumer warp and chunk: 2 for fp8 pools with <= 4 token slices (a slot
                     // then carries the bytes of a 16-bit one and the per-slot costs halve), else 1
  int kv_split;      // 1: a pipeline slot holds the K rows OR the V rows of a chunk (slots alternate K, V, K, ...):
                     // head_dim 256 on a 16-bit pool, where a K + V stage is 129 KB and only one would fit
};

// Tuning aids, read once: B200ATTN_CPASYNC_MAX_ROW (default and maximum 512: wider rows go as bulk copies),
// B200ATTN_NO_SWIZZLE=1 (padded rows on the cp.async path), B200ATTN_NO_XCH=1 (token slices always merged
// through the global workspace).
inline int cp_async_max_row() {
  static const int v = [] {
    const char* e = getenv("B200ATTN_CPASYNC_MAX_ROW");
    const int x = e ? atoi(e) : 512;
    return x > 512 ? 512 : x;  // one warp instruction covers at most one 512-byte row
  }();
  return v;
}
inline bool env_flag(const char* name) {
  const char* e = getenv(name);
  return e && e[0] == '1';
}
inline bool no_swizzle() {
  static const bool v = env_flag("B200ATTN_NO_SWIZZLE");
  return v;
}
inline bool no_pdl() {   // B200ATTN_NO_PDL=1: decode_kernel launched without the programmatic-serialization attribute
  static const bool v = env_flag("B200ATTN_NO_PDL");
  return v;
}
inline bool no_balance() {   // B200ATTN_NO_BALANCE=1: equal static ranges whatever the per-CTA rates say
  static const bool v = env_flag("B200ATTN_NO_BALANCE");
  return v;
}
inline bool no_narrow() {   // B200ATTN_NO_NARROW=1: small batches keep the widest rows (the default geometry)
  static const bool v = env_flag("B200ATTN_NO_NARROW");
  return v;
}
inline bool no_subs() {   // B200ATTN_NO_SUBS=1: fp8 chunks keep one 16-token sub-chunk per warp
  static const bool v = env_flag("B200ATTN_NO_SUBS");
  return v;
}
inline bool no_xch() {
  static const bool v = env_flag("B200ATTN_NO_XCH");
  return v;
}
// Slice exchange: per consumer warp [head_dim / 8 tiles][group * 4 lanes] float2 of unnormalised O, then
// [8 warps][8 rows] float2 (max in log2 units, sum).
inline int xch_bytes(int group, int d) { return kXchWarps * (d / 8) * group * 4 * 8 + kXchWarps * 8 * 8; }

__host__ __device__ inline bool is_fp8(int dtype) { return dtype == B200ATTN_FP8_E4M3 || dtype == B200ATTN_FP8_E5M2; }
__host__ __device__ inline int elem_bytes(int dtype) { return is_fp8(dtype) ? 1 : 2; }

// Returns false when the shape has no kernel.
// `batch` > 0 lets small decode batches trade row width for parallelism: with few (request, head group) pairs
// every request is cut across many of the 148 CTA ranges (C1: 16 requests x 1 head group = ~10 partials each,
// merged by one CTA at the very end of the kernel).  Fewer heads per CTA give more, shorter virtual requests
// (C1 at 2 heads per CTA: 64 of them, 2-3 partials each, merged in registers by the shortcut), as long as a pool
// row slice stays >= 256 bytes (the cp.async path streams those at the bulk-copy rate).  The plan and the
// decode call derive the geometry from the same (shape, batch), so they always agree.
constexpr int kFewVreqs = 64;
inline bool make_geometry(int hq, int hkv, int d, int kv_dtype, Geometry* g, int batch = 0) {
  if (hq <= 0 || hkv <= 0 || hq % hkv != 0) return false;
  if (d != 64 && d != 80 && d != 96 && d != 128 && d != 256) return false;  // instantiated head dims
  int group = hq / hkv;
  if (group > 16) return false;
  int e = elem_bytes(kv_dtype);
  if (e == 1 && d != 64 && d != 128 && d != 256) return false;  // fp8 pools: the instantiated head dims
  int hc = 8;
  while (hc > 1 && (hkv % hc != 0 || hc * d * e > kMaxRowBytes)) hc >>= 1;
  if (hc * d * e > kMaxRowBytes) return false;
  if (hkv / hc > kMaxHeadGroups) return false;
  if (batch > 0 && !no_narrow()) {
    while (hc > 1 && batch * (hkv / hc) < kFewVreqs && (hc / 2) * d * e >= 256 && hkv / (hc / 2) <= kMaxHeadGroups) hc >>= 1;
  }
  g->hc = hc;
  g->ts = kConsumerWarps / hc;
  g->ng = hkv / hc;
  g->group = group;
  g->subs = (e == 1 && 32 * g->ts <= 128 && !no_subs()) ? 2 : 1;   // (chunks hold at most 128 tokens)
  g->chunk_tokens = 16 * g->ts * g->subs;
  g->row_bytes = hc * d * e;
  // (the cp.async issue loop addresses 16-byte units by shifts: power-of-two rows only; 96- / 80-wide heads go as bulk copies)
  g->cp_async = g->row_bytes <= cp_async_max_row() && (g->row_bytes & (g->row_bytes - 1)) == 0;
  g->swizzle = g->cp_async && (d * e) % 128 == 0 && !no_swizzle();  // heads start on 128-byte lines
  g->row_stride = g->swizzle ? g->row_bytes : g->row_bytes + kRowPad;
  g->stage_bytes = 2 * g->chunk_tokens * g->row_stride;
  g->kv_split = 0;
  int stages = (kMaxSmem - kBarrierBytes) / g->stage_bytes;
  if (stages < 2) {   // K and V of a chunk in separate slots
    g->kv_split = 1;
    g->stage_bytes = g->chunk_tokens * g->row_stride;
    stages = (kMaxSmem - kBarrierBytes) / g->stage_bytes;
  }
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) return false;
  g->xch_off = 0;
  int xb = 0;
  if (g->ts > 1 && group <= 8 && !no_xch()) {  // only where it costs no pipeline stage
    xb = xch_bytes(group, d);
    int with = (kMaxSmem - kBarrierBytes - xb) / g->stage_bytes;
    if (with > kMaxStages) with = kMaxStages;
    if (with == stages) g->xch_off = kBarrierBytes + stages * g->stage_bytes;
    else xb = 0;
  }
  g->stages = stages;
  g->smem_bytes = kBarrierBytes + stages * g->stage_bytes + xb;
  return true;
}

// ---------------------------------------------------------------------------
// Plan buffer layout (device memory, written by plan_kernel, read by the
// decode and merge kernels).  All offsets are derived from `batch`.
//
// The chunk sequence of the whole batch (virtual request = request x head
// group, chunk = chunk_tokens tokens) is cut into RANGES: n_static ranges of
// per_static chunks (range c is CTA c's first piece of work) followed by
// n_dynamic ranges of per_dynamic chunks that CTAs claim through the ticket
// counter as they run dry -- the tail of slow SMs is absorbed by fast ones.
// ---------------------------------------------------------------------------
struct PlanHeader {
  int32_t magic;
  int32_t batch;
  int32_t ng;              // head groups
  int32_t chunk_tokens;
  int32_t total_chunks;    // ng * sum_b ceil(S_b / chunk_tokens)
  int32_t n_ranges;        // n_static + n_dynamic
  int32_t n_static;
  int32_t n_dynamic;
  int32_t per_static;      // chunks per static range
  int32_t per_dynamic;     // chunks per dynamic range
  int32_t grid;            // CTAs that take part (the others exit at once)
  int32_t csr_mode;
  int32_t total_tokens;    // sum_b S_b
  int32_t n_empty;         // requests of length 0 (the decode kernel defines their rows)
  unsigned long long ticket;  // dynamic-range tickets; every decode launch consumes n_dynamic + grid
};
static_assert(sizeof(PlanHeader) == 64, "PlanHeader is 64 bytes");
constexpr int32_t kPlanMagic = 0x42323031;  // "B201"

// Per-CTA streaming rate, kept ACROSS plans at a fixed offset of the plan buffer (so a changing batch size does
// not move it): decode_kernel folds every launch's steady-state clocks per chunk into an exponential average,
// and plan_kernel sizes the static ranges in inverse proportion.  Why: one CTA per SM, and the SMs do not stream
// equally fast -- tools/trace_decode.py on C2 / C4: clocks per chunk spread by +-8 % across SMs, the same SMs
// slow on every launch (blockIdx -> SM is stable; rank correlation 0.75-0.95 between launches), so with equal
// ranges the median CTA sat idle for the last 9 % of the kernel.  A buffer that was never zeroed is harmless:
// entries count only with a valid tag, and weights are clamped to [0.85, 1.18].
struct CtaStat {
  float clk_per_chunk;   // exponential average
  uint32_t tag;          // kStatTag ^ cta when valid
};
constexpr uint32_t kStatTag = 0x5ca1ab1eu;
constexpr int kMaxCtas = kMaxRanges / 2;
constexpr int kFirstIds = 128;   // >= the largest chunk (16 * 8 tokens): slot ids of a static range's first chunk

// What a CTA needs to start streaming range r without walking the plan's tables: the range's bounds and the
// piece it visits first (PieceWalk: the range's LAST virtual request).  Read with one load at kernel entry,
// next to the header, so the producers' first copy waits on desc -> slot ids -> rows instead of
// header -> range tables -> chunk prefix -> seq_len / row_off -> slot ids -> rows (each a dependent global
// round trip of 0.5-0.7 us: tools/trace_decode.py, 3.3 us from CTA entry to the first copy).
struct __align__(16) RangeDesc {
  int32_t v_first, v_last;   // first / last virtual request of the range
  int32_t g0, g1;            // chunk interval [g0, g1) of the range
  int32_t b, hg;             // the first piece visited: request, head group
  int32_t lc0, lc1;          // its chunk interval inside the request
  int32_t nchunks, seq_len;  // of request b
  int64_t row_off;           // element offset of request b's index row
};
static_assert(sizeof(RangeDesc) == 48, "RangeDesc is 48 bytes");

struct PlanView {
  PlanHeader* hdr;
  int32_t* seq_len;       // [batch]
  int32_t* chunk_prefix;  // [batch + 1] exclusive prefix of ceil(S_b / T)
  int64_t* row_off;       // [batch] element offset of request b's index row
  CtaStat* cta_stat;      // [kMaxCtas] fixed offset: survives re-planning with another batch size
  int32_t* range_start;   // [kMaxRanges + 1] first chunk of range r; entry n_ranges = total_chunks
  int32_t* range_vreq;    // [kMaxRanges] first virtual request of range r
  int32_t* range_vlast;   // [kMaxRanges] last virtual request of range r
  int32_t* vreq_rfirst;   // [batch * kMaxHeadGroups] first range that touches virtual request v
  int32_t* vreq_pieces;   // [batch * kMaxHeadGroups] ranges that touch v (its partials to merge)
  int32_t* vreq_arrived;  // [batch * kMaxHeadGroups] pieces written so far (self-resetting)
  RangeDesc* range_desc;  // [kMaxRanges]
  int32_t* first_ids;     // [kMaxCtas][kFirstIds] slot ids of the first chunk static range c streams (req_to_token mode)
};

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

__host__ __device__ inline size_t plan_bytes_for(int batch) {
  size_t n = sizeof(PlanHeader);
  n += (size_t)kMaxCtas * sizeof(CtaStat);     // cta_stat
  n += align_up((size_t)(kMaxRanges + 1) * 4, 16);  // range_start
  n += align_up((size_t)batch * 4, 16);        // seq_len
  n += align_up((size_t)(batch + 1) * 4, 16);  // chunk_prefix
  n += align_up((size_t)batch * 8, 16);        // row_off
  n += 2 * (size_t)kMaxRanges * 4;             // range_vreq, range_vlast
  n += 3 * align_up((size_t)batch * kMaxHeadGroups * 4, 16);
  n += (size_t)kMaxRanges * sizeof(RangeDesc);
  n += (size_t)kMaxCtas * kFirstIds * 4;
  return n;
}

__host__ __device__ inline PlanView plan_view(void* base, int batch) {
  PlanView v;
  char* p = reinterpret_cast<char*>(base);
  v.hdr = reinterpret_cast<PlanHeader*>(p);
  p += sizeof(PlanHeader);
  v.cta_stat = reinterpret_cast<CtaStat*>(p);
  p += (size_t)kMaxCtas * sizeof(CtaStat);
  v.range_start = reinterpret_cast<int32_t*>(p);
  p += align_up((size_t)(kMaxRanges + 1) * 4, 16);
  v.seq_len = reinterpret_cast<int32_t*>(p);
  p += align_up((size_t)batch * 4, 16);
  v.chunk_prefix = reinterpret_cast<int32_t*>(p);
  p += align_up((size_t)(batch + 1) * 4, 16);
  v.row_off = reinterpret_cast<int64_t*>(p);
  p += align_up((size_t)batch * 8, 16);
  v.range_vreq = reinterpret_cast<int32_t*>(p);
  p += (size_t)kMaxRanges * 4;
  v.range_vlast = reinterpret_cast<int32_t*>(p);
  p += (size_t)kMaxRanges * 4;
  const size_t tbl = align_up((size_t)batch * kMaxHeadGroups * 4, 16);
  v.vreq_rfirst = reinterpret_cast<int32_t*>(p);
  p += tbl;
  v.vreq_pieces = reinterpret_cast<int32_t*>(p);
  p += tbl;
  v.vreq_arrived = reinterpret_cast<int32_t*>(p);
  p += tbl;
  v.range_desc = reinterpret_cast<RangeDesc*>(p);
  p += (size_t)kMaxRanges * sizeof(RangeDesc);
  v.first_ids = reinterpret_cast<int32_t*>(p);
  return v;
}

// Workspace: partial outputs.  Piece (virtual request v, range r) has id v + r
// (unique and dense because both only grow along the chunk sequence); each
// piece holds ts slices of [hq_per_vreq, dv] fp32 plus [hq_per_vreq] lse.
inline size_t workspace_floats(int batch, int hq, int dv, const Geometry& g, size_t* lse_off) {
  // (a small batch may run with more head groups than the default geometry: at most 2 * kFewVreqs virtual requests)
  size_t pieces = (size_t)batch * g.ng + kMaxRanges;
  if (pieces < (size_t)2 * kFewVreqs + kMaxRanges) pieces = (size_t)2 * kFewVreqs + kMaxRanges;
  size_t rows = pieces * g.ts * (size_t)(hq / g.ng);
  if (lse_off) *lse_off = rows * dv;
  return rows * dv + rows;
}

int sm_count();        // SMs of the current device (148 when there is no device)
void count_launch();   // bumps the process-wide launch counter

#ifdef __CUDACC__
// ---------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// 1-D bulk async copy global -> shared (TMA engine, SASS UBLKCP); completion is
// signalled on an mbarrier as a transaction byte count.  16 B granularity.
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes,
                                         uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}

// 16-byte per-thread async copy global -> shared (SASS LDGSTS), L1 bypass.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}

// Same, with zero fill: when !valid nothing is read and the 16 destination bytes become zero.
__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void* src, bool valid) {
  const uint32_t n = valid ? 16u : 0u;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}

__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }

template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// Arrive on `bar` once all cp.async issued so far by this thread have landed.  .noinc: the
// arrival is one of the barrier's expected count (init count = arriving threads).
__device__ __forceinline__ void cp_async_mbar_arrive_noinc(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// Named barrier among `threads` threads (a subset of the CTA's warps).
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}

__device__ __forceinline__ void ldsm_x4_trans(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}

// D(16x8, f32) += A(16x16) * B(16x8).  HALF selects f16 vs bf16 operands.
template <bool HALF>
__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                         uint32_t a3, uint32_t b0, uint32_t b1) {
  if constexpr (HALF) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
        "{%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
        "{%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
}

// Two e4m3 bytes (low 16 bits of x) -> f16x2, exact.
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2(uint32_t x) {
  uint32_t r;
  asm("{\n\t.reg .b16 t;\n\tcvt.u16.u32 t, %1;\n\tcvt.rn.f16x2.e4m3x2 %0, t;\n\t}"
      : "=r"(r)
      : "r"(x));
  return r;
}

// Four e4m3 bytes -> two f16x2 (bytes 0, 1 -> lo; bytes 2, 3 -> hi), exact.
__device__ __forceinline__ void e4m3x4_to_f16x2x2(uint32_t x, uint32_t& lo, uint32_t& hi) {
  asm("{\n\t.reg .b16 a, b;\n\tmov.b32 {a, b}, %2;\n\tcvt.rn.f16x2.e4m3x2 %0, a;\n\tcvt.rn.f16x2.e4m3x2 %1, b;\n\t}"
      : "=r"(lo), "=r"(hi)
      : "r"(x));
}

// Four fp8 bytes -> two f16x2, exact.  e5m2 is the upper byte of an f16: a byte permute places it (PRMT).
template <bool E5M2>
__device__ __forceinline__ void fp8x4_to_f16x2x2(uint32_t x, uint32_t& lo, uint32_t& hi) {
  if constexpr (E5M2) {
    lo = __byte_perm(x, 0u, 0x1404);   // bytes (0, b0, 0, b1)
    hi = __byte_perm(x, 0u, 0x3424);   // bytes (0, b2, 0, b3)
  } else {
    e4m3x4_to_f16x2x2(x, lo, hi);
  }
}

// 8x8 transpose of 16-bit elements across the warp: lane (g, j) gives row g, columns 2j, 2j+1 and receives
// column g, rows 2j, 2j+1 (tools/probe_frag.cu).
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t x) {
  uint32_t y;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(y) : "r"(x));
  return y;
}

// Two 16 x 16 BYTE matrices, transposed (sm_100a; SASS LDSM.8.MT1616): lanes 0-15 / 16-31 give the row addresses
// of matrix 0 / 1; lane (g, j) receives r[0] = rows 4j..4j+3 of column g, r[1] = the same rows of column g + 8
// (matrix 0), r[2], r[3] likewise from matrix 1 (tools/probe_frag.cu).
__device__ __forceinline__ void ldsm8_x2_trans(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m16n16.x2.trans.shared.b8 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}

// Programmatic dependent launch (sm_90+).  launch_dependents: the next kernel on the stream, if it was launched
// with the programmatic-serialization attribute, may start its CTAs now.  wait: block until the kernels this
// one depends on have completed and their writes are visible (a no-op for a normally launched kernel).
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
constexpr int kPlanThreads = 1024;

__device__ __forceinline__ int64_t load_idx(const void* p, int i, int idx64) {
  return idx64 ? reinterpret_cast<const int64_t*>(p)[i]
               : static_cast<int64_t>(reinterpret_cast<const int32_t*>(p)[i]);
}

// Block-wide inclusive scan of one value per thread (1024 threads).
template <typename T>
__device__ __forceinline__ T block_scan_incl(T v, T* warp_sums) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    T t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  if (lane == 31) warp_sums[warp] = v;
  __syncthreads();
  if (warp == 0) {
    T w = warp_sums[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      T t = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += t;
    }
    warp_sums[lane] = w;
  }
  __syncthreads();
  if (warp > 0) v += warp_sums[warp - 1];
  __syncthreads();
  return v;
}
#endif  // __CUDACC__

}  // namespace b200attn
