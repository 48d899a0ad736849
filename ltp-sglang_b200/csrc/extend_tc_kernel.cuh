// K5 (tensor-core generation 5): extend / prefill attention on tcgen05 + TMEM (sm_100a).
//
// Same contract as extend_kernel.cuh (the reference's Triton extend kernel,
// python/sglang/srt/layers/attention/triton_ops/extend_attention.py:41-303), head_dim 128.
//
// One CTA = (request, kv head, q tile of 256 MMA rows) = two 128-row sub-tiles that share every
// K/V tile.  A row is a (token, q head of the kv head's group) pair, so GQA packs G heads into M.
//   warps 0-7 / 8-15  softmax warps of sub-tile 0 / 1, two threads per row (column halves; warp w serves TMEM
//                     lane quarter w & 3).  General kernel: the thread's 64 S columns are read ONCE from TMEM
//                     (two tcgen05.ld in flight), masked, the row max is exchanged with the other half's warp
//                     (pairwise named barrier), P goes back as packed 16-bit pairs into the columns S occupied
//                     (tcgen05.st) -- the first 32 tokens' P is published early so that half of P.V runs under
//                     the remaining exps -- and O is rescaled in TMEM only when the max grew by more than 2^8.
//                     (Small-q instantiation, ALT: the older two-pass body on 32-column chunks.)
//   warps 16-19       loaders: gather K / V rows of the paged pool (prefix through the index list,
//                     new tokens through out_cache_loc) with 16-byte cp.async straight into the
//                     128-byte-swizzled K-major layout tcgen05 reads (row r, chunk c -> c ^ (r & 7)),
//                     ring of kSlots 32 KB slots, full / empty mbarriers; also the Q tile.
//   warp 20           issues every MMA (warp-uniform, elect inside): S_i = Q_i . K^T   (SS, both K-major)
//                                                           O_i += P_i . V     (TS: P from TMEM,
//                                                           V MN-major = the same row layout)
//                     in the order PV0(n) S0(n+1) PV1(n) S1(n+1), so the softmax of one sub-tile
//                     overlaps the MMAs of the other; tcgen05.commit signals the mbarriers.
//   warps 21-23       idle: they exist so that every SM sub-partition holds 6 x 80 launch registers, which
//                     setmaxnreg then moves to the softmax warps (96) from loaders (64), MMA (32), idle (24).
// TMEM (512 columns): S0 [0,128) S1 [128,256) O0 [256,384) O1 [384,512); P_i aliases S_i [0,32) + [64,96).
// The descriptor encodings are pinned by tools/probe_umma.cu.
#pragma once

#include <math_constants.h>

#include <type_traits>

#include "extend_kernel.cuh"

namespace b200attn {

constexpr int kTcSoftmaxWarps = 16;   // 2 sub-tiles x 2 column halves x 4 lane quarters
constexpr int kTcLoaderWarp0 = 16;    // warps 16-19 load, warp 20 issues the MMAs
constexpr int kTcMmaWarp = 20;
constexpr int kTcThreads = 21 * 32;   // 16 softmax + 4 loader + 1 MMA warps, 80 registers each
constexpr int kTcBM = 128;            // rows per sub-tile (TMEM lanes)
constexpr int kTcBN = 128;            // kv tokens per tile
constexpr int kTcD = 128;            // widest head_dim; the shared / tensor memory geometry is laid out for it
constexpr int kTcSlots = 4;           // K / V ring slots
constexpr int kTcSlab = kTcBN * 128;  // bytes of one 64-element-wide slab of 128 rows
constexpr int kTcSlotBytes = 2 * kTcSlab;
constexpr int kTcQBytes = 2 * kTcSlotBytes;  // two sub-tiles
#ifndef B200ATTN_TC_LAG
#define B200ATTN_TC_LAG 2
#endif
constexpr int kTcLag = B200ATTN_TC_LAG;  // cp.async groups a loader thread keeps in flight (< kTcSlots)
constexpr int kTcBarBytes = 9216;    // mbarriers, TMEM base, row max / row sum exchange, 4 slot-id buffers
constexpr int kTcSmemBytes = kTcQBytes + kTcSlots * kTcSlotBytes + kTcBarBytes + 1024;  // + alignment slack
constexpr float kTcRescaleLog2 = 8.f;  // lazy rescale threshold (log2 units)
#ifdef B200ATTN_TC_TWO_PASS
constexpr bool kTcSinglePass = false;   // A/B build: the general kernel keeps the two-pass softmax body
#else
constexpr bool kTcSinglePass = true;
#endif
#ifdef B200ATTN_TC_ALT_TWO_PASS
constexpr bool kTcSinglePassAlt = false;   // A/B build: the small-q (ALT) instantiation keeps the two-pass body
#else
constexpr bool kTcSinglePassAlt = true;
#endif
#ifdef B200ATTN_TC_NO_EARLY_P
constexpr bool kTcEarlyP = false;
#else
constexpr bool kTcEarlyP = kTcSinglePass;   // (the two-pass body publishes P once per tile)
#endif
#ifdef B200ATTN_TC_LOADER_WIDENS
constexpr bool kTcWidenWarps = false;   // A/B build: the loaders widen fp8 stages themselves
#else
constexpr bool kTcWidenWarps = true;
#endif
#ifndef B200ATTN_TC_REGS
#define B200ATTN_TC_REGS 96, 64, 32
#endif
constexpr int kTcRegs[3] = {B200ATTN_TC_REGS};  // softmax / loader / MMA warps after setmaxnreg
constexpr int kTcSoftmaxRegs = kTcRegs[0], kTcLoaderRegs = kTcRegs[1], kTcMmaRegs = kTcRegs[2];
// setmaxnreg moves registers inside what the launch allocated: 80 per warp (the compiler's cap for a 6-warp
// sub-partition).  The general kernel is launched with three idle warps (21-23) so that EVERY sub-partition holds
// 6 x 80 = 480: 4 softmax + 1 loader + the MMA warp or an idle one (24 registers).
constexpr int kTcLaunchThreads = 24 * 32;
constexpr int kTcWideThreads = 16 * 32;   // WIDE: 8 softmax + 4 loader + 1 MMA + 3 idle warps, 128 registers at launch
static_assert(4 * kTcSoftmaxRegs + kTcLoaderRegs + kTcMmaRegs <= 480, "registers of sub-partition 0");

// ---- tcgen05 wrappers ----
// The MMA / commit wrappers are executed by a WHOLE warp (uniform control flow) and elect one lane
// inside the asm: operands then live in uniform registers.  Issued from a lane == 0 branch instead,
// every MMA costs ~125 clk of R2UR / waterfall code on the issuing thread (tools/probe_umma.cu).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = static_cast<uint64_t>((addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= 1ull << 46;  // descriptor version (sm_100)
  d |= 2ull << 61;  // SWIZZLE_128B
  return d;
}
// kind::f16 instruction descriptor: fp32 accumulate, A/B format fmt (0 = f16, 1 = bf16)
__host__ __device__ constexpr uint32_t umma_idesc(int m, int n, int fmt, int b_mn_major) {
  return (1u << 4) | (static_cast<uint32_t>(fmt) << 7) | (static_cast<uint32_t>(fmt) << 10) |
         (static_cast<uint32_t>(b_mn_major) << 16) | (static_cast<uint32_t>(n >> 3) << 17) |
         (static_cast<uint32_t>(m >> 4) << 24);
}
__device__ __forceinline__ void umma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p, e;\n\telect.sync _|e, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
      "l"(a), "l"(b), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void umma_ts(uint32_t d, uint32_t a_tmem, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p, e;\n\telect.sync _|e, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "@e tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d),
      "r"(a_tmem), "l"(b), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred e;\n\telect.sync _|e, 0xffffffff;\n\t"
      "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(bar)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

#define B200_R16(r, o) r[o + 0], r[o + 1], r[o + 2], r[o + 3], r[o + 4], r[o + 5], r[o + 6], r[o + 7], r[o + 8], \
    r[o + 9], r[o + 10], r[o + 11], r[o + 12], r[o + 13], r[o + 14], r[o + 15]

// 32 lanes x 32 consecutive columns -> 32 registers per thread (thread = lane); no wait.
template <int OFF, int LEN>
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&r)[LEN]) {
  uint32_t* u = reinterpret_cast<uint32_t*>(&r[OFF]);
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
template <int OFF, int LEN>
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const float (&r)[LEN]) {
  const uint32_t* u = reinterpret_cast<const uint32_t*>(&r[OFF]);
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%32], "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,"
      "%29,%30,%31};" ::"r"(u[0]),
      "r"(u[1]), "r"(u[2]), "r"(u[3]), "r"(u[4]), "r"(u[5]), "r"(u[6]), "r"(u[7]), "r"(u[8]), "r"(u[9]), "r"(u[10]),
      "r"(u[11]), "r"(u[12]), "r"(u[13]), "r"(u[14]), "r"(u[15]), "r"(u[16]), "r"(u[17]), "r"(u[18]), "r"(u[19]),
      "r"(u[20]), "r"(u[21]), "r"(u[22]), "r"(u[23]), "r"(u[24]), "r"(u[25]), "r"(u[26]), "r"(u[27]), "r"(u[28]),
      "r"(u[29]), "r"(u[30]), "r"(u[31]), "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&r)[16]) {
  uint32_t* u = reinterpret_cast<uint32_t*>(&r[0]);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]), "=r"(u[8]),
        "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&u)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%16], {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15};" ::"r"(
          u[0]),
      "r"(u[1]), "r"(u[2]), "r"(u[3]), "r"(u[4]), "r"(u[5]), "r"(u[6]), "r"(u[7]), "r"(u[8]), "r"(u[9]), "r"(u[10]),
      "r"(u[11]), "r"(u[12]), "r"(u[13]), "r"(u[14]), "r"(u[15]), "r"(taddr)
      : "memory");
}

// mbarrier wait with a watchdog: a protocol bug traps instead of hanging the GPU.
// (Measured and dropped, profiles/r02_extend2: a suspend-time hint on try_wait -- the softmax warps' s_full spin is
//  45 % of the executed instructions, yet parking them changed nothing -- and a pure test_wait poll.)
__device__ __forceinline__ void mbar_wait_wd(uint32_t bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 20)) __trap();
  }
}

template <int REGS>
__device__ __forceinline__ void setmaxnreg_inc() {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(REGS));
}
template <int REGS>
__device__ __forceinline__ void setmaxnreg_dec() {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(REGS));
}

// -DB200ATTN_TC_TRACE: CTA (0,0,0) writes clock64() timestamps of its pipeline events into the lse
// buffer (tools/trace_tc.py decodes them); a tuning build, never the shipped library.
#ifdef B200ATTN_TC_TRACE
#define TC_TR(kind, n) do { if (tr && (n) < 256) tr[(kind) * 256 + (n)] = clock64(); } while (0)
#else
#define TC_TR(kind, n) do { } while (0)
#endif

// FP8: the pool holds float8_e4m3fn bytes.  The loaders cp.async the raw rows (128 B per token and head)
// into the upper slab of the ring slot and, once landed, widen them in place to f16 (exact) in the swizzled
// layout; both MMAs then run in f16 with q converted (and clamped) once, k_scale folded into the softmax
// scale and v_scale into the final normalisation -- the decode kernel's fp8 semantics.
// ALT: compiled-in support for the small-q mode (launched when no CTA of the grid can have more than 128
// rows); the general kernel carries none of its bookkeeping in the softmax loop's registers.
// CAP: logit cap (cap * tanh(x / cap)) compiled in.  As a run-time flag the compiler if-converted it: the
// softmax loop evaluated tanh for every element and selected afterwards -- ~17 instructions and two extra
// MUFU ops per element on a loop that is issue-bound -- although no Llama-class model sets a cap.
// D: head_dim 128 or 64.  64 keeps the 128-wide geometry (ring slots, O columns) and uses its first slab /
// first 64 O columns: rows are one 128-byte swizzle atom wide, Q.K^T takes 4 k-steps, P.V has N = 64.
// WIDE (general kernel only): ONE thread per row -- 8 softmax warps (warp w: sub-tile w / 4, TMEM lane quarter w & 3)
// on 208 registers each read the whole 128-column S row in one burst of four loads, so there is no row-max exchange
// between warps, half as many warps have to arrive before P is published and only two softmax warps share a
// sub-partition's MUFU / issue slots.  16 warps are launched (8 softmax, 4 loaders, MMA, 3 idle) with 128 registers
// each; setmaxnreg turns that into 208 / 64 / 32 / 24 (2 x 208 + 64 + 32 = 512 on the fullest sub-partition).
template <typename QT, int D, bool FP8, bool ALT, bool CAP, bool WIDE = false>
__global__ void __launch_bounds__(WIDE ? kTcWideThreads : kTcLaunchThreads, 1) extend_tc_kernel(const ExtendParams p) {
  static_assert(!(WIDE && ALT), "the small-q mode keeps two threads per row");
  constexpr int kTcSoftmaxWarps = WIDE ? 8 : b200attn::kTcSoftmaxWarps;   // (shadow the two-threads-per-row constants)
  constexpr int kTcLoaderWarp0 = kTcSoftmaxWarps;
  constexpr int kTcMmaWarp = kTcLoaderWarp0 + 4;
  constexpr int kTcSoftmaxRegs = WIDE ? 208 : b200attn::kTcSoftmaxRegs;
  // cp.async groups in flight per loader thread: the small-q mode over a 16-bit pool streams (HBM-bound) and wants three
  // 32 KB stages in flight per SM (verify shape 407 -> 389 us, prefix 16384: 393 -> 360 us = 5.96 TB/s); the fp8 and the
  // tensor-bound instantiations lose with three (P2f 797 -> 709 TFLOP/s: a stage is published one iteration later)
  constexpr int kTcLag = (ALT && !FP8) ? 3 : b200attn::kTcLag;
  // fp8 pools, small-q mode: warps 21-23 widen a landed stage to f16 in place, so the loaders only issue copies --
  // with the widening in the loader loop (LDS -> cvt -> STS between two stages' cp.async) the loaders bound every
  // fp8 shape (the verify shape ran at 2.3 TB/s against 5.2 on a 16-bit pool)
  constexpr bool WIDEN = FP8 && ALT && kTcWidenWarps;   // (the general kernel is tensor-bound: there the extra hop costs more than it frees, P2f 790 -> 741 TFLOP/s)
  constexpr int kTcLoaderRegs = WIDEN ? 48 : b200attn::kTcLoaderRegs;
  static_assert(D == 128 || (D == 64 && !FP8), "head_dim 128, or 64 over a 16-bit pool");
  constexpr bool NARROW = FP8 || D == 64;  // a pool row (one head) is 128 bytes: 8 chunks, 8 rows per loader thread
  constexpr int OH = D / 2;                // O columns per thread (two threads per row)
  constexpr bool OUT_HALF = std::is_same<QT, __half>::value;
  constexpr bool HALF = OUT_HALF || FP8;  // MMA operand type
  constexpr uint32_t IDESC_QK = umma_idesc(kTcBM, kTcBN, HALF ? 0 : 1, 0);
  constexpr uint32_t IDESC_PV = umma_idesc(kTcBM, D, HALF ? 0 : 1, 1);

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;  // swizzle atoms are 1024-byte aligned
  const uint32_t q_base = smem0;
  const uint32_t kv_base = smem0 + kTcQBytes;
  const uint32_t bars = kv_base + kTcSlots * kTcSlotBytes;
  const uint32_t bar_kv_full = bars;                      // [kTcSlots]
  const uint32_t bar_kv_empty = bars + 8 * kTcSlots;      // [kTcSlots]
  const uint32_t bar_q_full = bars + 16 * kTcSlots;
  const uint32_t bar_s_full = bar_q_full + 8;             // [2]
  const uint32_t bar_p_full = bar_s_full + 16;            // [2]
  const uint32_t bar_o_full = bar_p_full + 16;
  const uint32_t bar_p_half = bar_o_full + 8;             // [2] general kernel: P of every thread's first 32 tokens is in TMEM
  const uint32_t bar_raw_full = bar_p_half + 16;          // [kTcSlots] fp8: a stage's raw bytes have landed (loaders -> widening warps)
  const uint32_t tmem_slot = bar_raw_full + 8 * kTcSlots;
  // row max [parity][sub][half][row] and row sum [sub][half][row] exchanged between the two threads of a row
  float* xch = reinterpret_cast<float*>(smem_raw + (bars + 1024 - smem_u32(smem_raw)));
  float* lxch = xch + 2 * 2 * 2 * kTcBM;
  int32_t* sid = reinterpret_cast<int32_t*>(lxch + 2 * 2 * kTcBM);  // [4][kTcBN] slot ids of tile n at n & 3
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y;
#ifdef B200ATTN_TC_TRACE
  long long* tr = (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (((threadIdx.x & (WIDE ? 127 : 255)) == 0 && warp < kTcSoftmaxWarps) || threadIdx.x == kTcLoaderWarp0 * 32 || threadIdx.x == kTcMmaWarp * 32))
                      ? reinterpret_cast<long long*>(p.lse) : nullptr;
#endif

#ifdef B200ATTN_TC_TRACE
  // every CTA: globaltimer at entry / exit and its SM id, at byte offset 32 KB of the lse buffer
  long long* cta_tr = reinterpret_cast<long long*>(p.lse) + 4096 +
                      4 * ((blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x);
  if (threadIdx.x == 0) {
    unsigned long long gt;
    unsigned smid;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(gt));
    asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
    cta_tr[0] = static_cast<long long>(gt);
    cta_tr[2] = smid;
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) reinterpret_cast<long long*>(p.lse)[12 * 256 + 255] = clock64();
  }
#endif
  // ---- tile geometry (CTA-uniform) ----
  // all of the request's scalars in one round trip (the loads are independent; the early return below would
  // otherwise serialise them into three)
  const int q0 = p.qo_indptr[b];
  const int q1 = p.qo_indptr[b + 1];
  const int kv0 = p.kv_indptr[b];
  const int kv1 = p.kv_indptr[b + 1];
  int64_t req_row = 0, pl_full = 0;
  if (!p.kv_indices) {
    req_row = p.idx64 ? static_cast<const int64_t*>(p.req_pool_indices)[b]
                      : static_cast<const int32_t*>(p.req_pool_indices)[b];
    pl_full = p.idx64 ? static_cast<const int64_t*>(p.prefix_lens)[b] : static_cast<const int32_t*>(p.prefix_lens)[b];
  }
  const int Le = q1 - q0;
  const int tok_per_cta = (2 * kTcBM) >> p.gp_shift;
  const int ntq = (Le + tok_per_cta - 1) / tok_per_cta;
  const int t = ntq - 1 - static_cast<int>(blockIdx.x);  // longest tiles first
  if (t < 0) return;
  const int tok0 = t * tok_per_cta;
  const int Lp = kv1 - kv0;
  const bool use_mask = p.custom_mask != nullptr;
  const int tile_tok_end = tok0 + tok_per_cta < Le ? tok0 + tok_per_cta : Le;
  const int ext_end = (p.causal && !use_mask) ? tile_tok_end : Le;
  const int Ltot = Lp + ext_end;
  const int N = (Ltot + kTcBN - 1) / kTcBN;
  const int rows_valid = (tile_tok_end - tok0) << p.gp_shift;
  const int nsub = rows_valid > kTcBM ? 2 : 1;
  // Small-q mode ("KV ping-pong"): a tile of at most 128 rows (decode-like batches, TARGET_VERIFY) would
  // leave sub-tile 1 idle and serialise S -> softmax -> PV on one accumulator.  Instead both sub-tiles
  // take the SAME 128 q rows and alternate over the KV tiles (sub-tile v owns tiles n = v mod 2), each
  // with its own running max / sum / O; the two partial states are merged in the epilogue (an in-CTA
  // 2-way split-KV).  K / V stages are then consumed in the order K0 K1 V0 K2 V1 K3 ...
  const bool alt = ALT && nsub == 1 && N >= 2;
  const int32_t* pidx = p.kv_indices ? p.kv_indices + kv0 : p.req_to_token + req_row * p.r2t_stride + (pl_full - Lp);

  if (threadIdx.x == 0) {
    for (int s = 0; s < kTcSlots; ++s) {
      mbar_init(bar_kv_full + 8 * s, WIDEN ? 96 : 128);   // fp8: the three widening warps publish a stage
      mbar_init(bar_kv_empty + 8 * s, 1);
      mbar_init(bar_raw_full + 8 * s, 128);
    }
    mbar_init(bar_q_full, 128);
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_s_full + 8 * i, 1);
      mbar_init(bar_p_full + 8 * i, WIDE ? 128 : 256);
      mbar_init(bar_p_half + 8 * i, WIDE ? 128 : 256);
    }
    mbar_init(bar_o_full, 1);
    fence_barrier_init();
  }
  if (warp == kTcMmaWarp) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(tmem_slot));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot_ptr, 0);  // provably warp-uniform
  TC_TR(12, 250);  // roles start (each tracing thread overwrites: the last one wins, they are close)

  if (warp < kTcSoftmaxWarps) {
    // =============================== softmax warps ===============================
    // (registers: the launch gives every warp 80 -- one sub-partition holds 6 of the 21 warps -- the loaders and
    //  the MMA warp hand back what they do not need and the softmax warps take it: 4 x 96 + 64 + 32 = 480)
    setmaxnreg_inc<kTcSoftmaxRegs>();
    if constexpr (WIDE) {
    // ---------- one thread per row ----------
    const int sub = warp >> 2;
    if (sub < nsub) {
      const int r = (warp & 3) * 32 + lane;
      const int R = sub * kTcBM + r;
      const int itok = tok0 + (R >> p.gp_shift);
      const int hd = R & ((1 << p.gp_shift) - 1);
      const bool rv = hd < p.group && itok < Le;
      const int i_min_sub = tok0 + ((sub * kTcBM) >> p.gp_shift);
      const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
      const uint32_t tS = tmem + lane_off + sub * kTcBN;                 // this row's 128 S columns
      const uint32_t tO = tmem + lane_off + 2 * kTcBN + sub * kTcD;      // and its D O columns
      constexpr bool cap = CAP;
      const float c = cap ? kLog2e : p.scale_log2;
      const float cap_pre = cap ? p.sm_scale / p.logit_cap : 0.f;
      const int64_t mask_row_len = static_cast<int64_t>(Lp) + Le;
      const uint8_t* mrow = nullptr;
      if (use_mask) {
        const int64_t ms = p.mask64 ? static_cast<const int64_t*>(p.mask_indptr)[b]
                                    : static_cast<const int32_t*>(p.mask_indptr)[b];
        mrow = p.custom_mask + ms + (rv ? itok : 0) * mask_row_len;
      }
      float m = -CUDART_INF_F, l = 0.f;
      for (int n = 0; n < N; ++n) {
        mbar_wait_wd(bar_s_full + 8 * sub, n & 1);
        tc_fence_after();
        TC_TR(sub * 3, n);
        const int j0 = n * kTcBN, j_end = j0 + kTcBN;
        const bool need_mask = j_end > Ltot || (p.window > 0 && j0 < Lp) ||
                               (use_mask && (!p.skip_prefix_mask || j_end > Lp)) ||
                               (!use_mask && p.causal && j_end - 1 - Lp > i_min_sub);
        auto row_body = [&](auto masked_tag) {
          constexpr bool MASKED = decltype(masked_tag)::value;
          float s[128];
          tmem_ld32<0>(tS, s);
          tmem_ld32<32>(tS + 32, s);
          tmem_ld32<64>(tS + 64, s);
          tmem_ld32<96>(tS + 96, s);
          tmem_wait_ld();
          if constexpr (cap) {
#pragma unroll
            for (int j = 0; j < 128; ++j) s[j] = p.logit_cap * tanhf(s[j] * cap_pre);
          }
          if constexpr (MASKED) {
            // one visible range [lo, lo + span) per row (window front, causal / tile end back) plus the custom-mask
            // bytes from mask_from on: see the two-threads-per-row body below
            const int hi = !rv ? 0 : ((p.causal && !use_mask) ? min(Ltot, Lp + itok + 1) : Ltot);
            const int lo = p.window > 0 ? min(max(itok - p.window, 0), Lp) : 0;
            const unsigned span = hi > lo ? static_cast<unsigned>(hi - lo) : 0u;
            const int mask_from = !use_mask ? 0x7fffffff : (p.skip_prefix_mask ? Lp : 0);
#pragma unroll
            for (int jj = 0; jj < 128; ++jj)
              s[jj] = static_cast<unsigned>(j0 + jj - lo) < span ? s[jj] : -CUDART_INF_F;
            if (j0 + 128 > mask_from) {
#pragma unroll
              for (int jj = 0; jj < 128; ++jj) {
                const int j = j0 + jj;
                if (j >= mask_from && j < hi && mrow[j] == 0) s[jj] = -CUDART_INF_F;
              }
            }
          }
          float mx0 = -CUDART_INF_F, mx1 = -CUDART_INF_F, mx2 = -CUDART_INF_F, mx3 = -CUDART_INF_F;
#pragma unroll
          for (int j = 0; j < 128; j += 4) {
            mx0 = fmaxf(mx0, s[j]);
            mx1 = fmaxf(mx1, s[j + 1]);
            mx2 = fmaxf(mx2, s[j + 2]);
            mx3 = fmaxf(mx3, s[j + 3]);
          }
          TC_TR(sub * 3 + 1, n);
          const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
          // lazy rescale: keep the stale max while the new one is within 2^8 of it (p stays <= 256)
          const bool grow = mx > m && (m == -CUDART_INF_F || (mx - m) * c > kTcRescaleLog2);
          if (__any_sync(0xffffffffu, grow)) {
            float corr = 1.f;
            if (grow) {
              corr = fast_exp2((m - mx) * c);  // m = -inf -> 0
              m = mx;
              l *= corr;
            }
            if (n > 0) {
              // this sub-tile's previous P.V has completed: its commit precedes the one that signalled s_full
#pragma unroll
              for (int cc = 0; cc < D; cc += 16) {   // 16 columns at a time: s[128] is live
                float o[16];
                tmem_ld16(tO + cc, o);
                tmem_wait_ld();
                uint32_t ou[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) ou[j] = __float_as_uint(o[j] * corr);
                tmem_st16(tO + cc, ou);
              }
            }
          }
          const float mc = (m == -CUDART_INF_F ? 0.f : m) * c;
          float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
          for (int cc = 0; cc < 128; cc += 32) {
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              const float p0 = fast_exp2(fmaf(s[cc + j], c, -mc));
              const float p1 = fast_exp2(fmaf(s[cc + j + 1], c, -mc));
              sum0 += p0;
              sum1 += p1;
              pk[j >> 1] = pack2<HALF>(p0, p1);
            }
            // packed P of tokens cc .. cc + 31 -> columns cc / 2 .. of the row's own S tile (already consumed)
            tmem_st16(tS + (cc >> 1), pk);
            if (cc == 32) {
              // the first 64 tokens' P is in TMEM: the MMA warp runs their four P.V k-steps under the other 64 exps
              tmem_wait_st();
              tc_fence_before();
              mbar_arrive(bar_p_half + 8 * sub);
            }
          }
          l += sum0 + sum1;
          tmem_wait_st();
          tc_fence_before();
          mbar_arrive(bar_p_full + 8 * sub);
          TC_TR(sub * 3 + 2, n);
        };
        if (need_mask) row_body(std::true_type{});
        else row_body(std::false_type{});
      }
      // ---- epilogue: the row's D columns of O / l -> global ----
      const float mlog2 = (m == -CUDART_INF_F ? 0.f : m) * c;
      mbar_wait_wd(bar_o_full, 0);
      tc_fence_after();
      TC_TR(13, 250);  // epilogue: O complete
      const float inv = l > 0.f ? p.v_scale / l : 0.f;
      QT* orow = reinterpret_cast<QT*>(p.o) + static_cast<int64_t>(q0 + (rv ? itok : 0)) * p.o_st +
                 static_cast<int64_t>(h * p.group + (rv ? hd : 0)) * p.o_sh;
#pragma unroll
      for (int cc = 0; cc < D; cc += 32) {
        float o[32];
        tmem_ld32<0>(tO + cc, o);
        tmem_wait_ld();
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
      if (rv && p.lse)
        p.lse[static_cast<int64_t>(q0 + itok) * p.hq + h * p.group + hd] =
            l > 0.f ? mlog2 * kLn2 + logf(l) : -CUDART_INF_F;
#endif
      TC_TR(13, 251);  // epilogue: outputs stored
      tc_fence_before();
    }
    } else {
    // Two threads per row: warp w serves sub-tile w / 8, column half (w / 4) & 1, TMEM lane quarter w & 3
    // (a warp reaches only the 32 lanes of quarter warp_id % 4).  The halves exchange their row max
    // through shared memory (one named barrier per tile) and their row sum once at the end.
    const int sub = warp >> 3;
    if (sub < nsub || alt) {
      const int half = (warp >> 2) & 1;
      const int r = (warp & 3) * 32 + lane;
      const int R = alt ? r : sub * kTcBM + r;
      const int itok = tok0 + (R >> p.gp_shift);
      const int hd = R & ((1 << p.gp_shift) - 1);
      const bool rv = hd < p.group && itok < Le;
      const int i_min_sub = alt ? tok0 : tok0 + ((sub * kTcBM) >> p.gp_shift);
      const int dn = alt ? 2 : 1;
      // a warp whose 32 rows all lie past the tile's last valid row (small-q batches: 8 draft tokens x 4 heads
      // fill one warp of four) only keeps the barriers moving: its P rows are never written (whatever bits sit
      // there feed O rows nobody reads -- rows are independent) and it touches neither TMEM nor the MUFU
      const bool dead = ALT && (alt ? 0 : sub * kTcBM) + (warp & 3) * 32 >= rows_valid;
      const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
      const uint32_t tS = tmem + lane_off + sub * kTcBN + half * 64;        // this thread's S columns
      const uint32_t tO = tmem + lane_off + 2 * kTcBN + sub * kTcD + half * OH;
      constexpr bool cap = CAP;
      const float c = cap ? kLog2e : p.scale_log2;
      const float cap_pre = cap ? p.sm_scale / p.logit_cap : 0.f;
      const int64_t mask_row_len = static_cast<int64_t>(Lp) + Le;
      const uint8_t* mrow = nullptr;
      if (use_mask) {
        const int64_t ms = p.mask64 ? static_cast<const int64_t*>(p.mask_indptr)[b]
                                    : static_cast<const int32_t*>(p.mask_indptr)[b];
        mrow = p.custom_mask + ms + (rv ? itok : 0) * mask_row_len;
      }
      float* xmine = xch + (sub * 2 + half) * kTcBM + r;
      float* xother = xch + (sub * 2 + (half ^ 1)) * kTcBM + r;
      float m = -CUDART_INF_F, l = 0.f;

      int it = 0;  // tiles this sub-tile has processed
      for (int n = alt ? sub : 0; n < N; n += dn, ++it) {
        mbar_wait_wd(bar_s_full + 8 * sub, it & 1);
        tc_fence_after();
        TC_TR(sub * 3, n);
        if (dead) {
          // a sync, not an arrive: the warp must not run ahead into the epilogue's use of the barrier.  (Single-pass
          // body: the row max is exchanged on the pair's own barrier, and both warps of a pair are dead together.)
          if constexpr (kTcSinglePass && kTcSinglePassAlt) {
            named_bar_sync(6 + sub * 4 + (warp & 3), 64);
            mbar_arrive(bar_p_half + 8 * sub);
          } else {
            named_bar_sync(1 + sub, 256);
          }
          mbar_arrive(bar_p_full + 8 * sub);
          continue;
        }
        const int j0 = n * kTcBN, j_end = j0 + kTcBN;
        const bool need_mask = j_end > Ltot || (p.window > 0 && j0 < Lp) ||
                               (use_mask && (!p.skip_prefix_mask || j_end > Lp)) ||
                               (!use_mask && p.causal && j_end - 1 - Lp > i_min_sub);
        // The tile body exists twice, with and without the mask code: left to the compiler, the range test is
        // if-converted into the unmasked path too (3 instructions per element on an issue-bound loop).
        auto tile_body = [&](auto masked_tag) {
        constexpr bool MASKED = decltype(masked_tag)::value;
        // 32 of this thread's 64 S columns -> registers, logit cap and masks applied.  The row is read
        // twice (max pass, exp pass): TMEM loads are cheap, 64 live values per thread are not.
        auto load_chunk = [&](int cc, float (&s)[32]) {
          tmem_ld32<0>(tS + cc, s);
          tmem_wait_ld();
          if constexpr (cap) {
#pragma unroll
            for (int j = 0; j < 32; ++j) s[j] = p.logit_cap * tanhf(s[j] * cap_pre);
          }
          if constexpr (MASKED) {
            // Visibility of key j for this row, custom-mask bytes aside, is one range [lo, lo + span): the
            // sliding window cuts the front of the prefix part (i <= j + W, prefix keys only), causality /
            // the tile's end the back.  The mask bytes (speculative verify) are read only from mask_from on.
            // (Derived here, on masked tiles only, to keep them out of the main loop's live registers.)
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
        TC_TR(sub * 3 + 1, n);
        float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
        // row max over both halves
        const int par = (it & 1) * (2 * 2 * kTcBM);
        xmine[par] = mx;
        named_bar_sync(1 + sub, 256);
        mx = fmaxf(mx, xother[par]);
        // lazy rescale: keep the stale max while the new one is within 2^8 of it (p stays <= 256).
        // Both threads of a row see the same m and mx, so they take the same decision.
        const bool grow = mx > m && (m == -CUDART_INF_F || (mx - m) * c > kTcRescaleLog2);
        if (__any_sync(0xffffffffu, grow)) {
          float corr = 1.f;
          if (grow) {
            corr = fast_exp2((m - mx) * c);  // m = -inf -> 0
            m = mx;
            l *= corr;
          }
          if (it > 0) {
            // this sub-tile's previous PV has completed: its commit precedes the one that signalled s_full
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
          // packed P of these 32 tokens goes into the first 32 columns of this thread's OWN S half
          // (columns it has already consumed): no thread ever overwrites S data another still needs
          tmem_st16(tS + (cc >> 1), pk);
        }
        l += sum0 + sum1;
        tmem_wait_st();
        tc_fence_before();
        mbar_arrive(bar_p_full + 8 * sub);
        TC_TR(sub * 3 + 2, n);
        };
        // General kernel: the thread's 64 S columns are read ONCE (two loads in flight, one wait) and stay in
        // registers across the row-max exchange -- the two-pass body above pays four load round trips per tile on
        // a chain (S -> softmax -> P.V -> next S) that is latency-bound; 104 registers make room for it.
        auto tile_body1 = [&](auto masked_tag) {
          constexpr bool MASKED = decltype(masked_tag)::value;
          float s[64];
#ifdef B200ATTN_TC_TRACE
          // per-warp skew: lane 0 of every softmax warp of CTA 0 stamps s_full seen / P published (kinds 32 + 2 warp, + 1)
          long long* trw = (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0 && n < 256)
                               ? reinterpret_cast<long long*>(p.lse) + (32 + 2 * warp) * 256 + n : nullptr;
          if (trw) trw[0] = clock64();
#endif
          tmem_ld32<0>(tS, s);
          tmem_ld32<32>(tS + 32, s);
          tmem_wait_ld();
          if constexpr (cap) {
#pragma unroll
            for (int j = 0; j < 64; ++j) s[j] = p.logit_cap * tanhf(s[j] * cap_pre);
          }
          if constexpr (MASKED) {
            const int hi = !rv ? 0 : ((p.causal && !use_mask) ? min(Ltot, Lp + itok + 1) : Ltot);
            const int lo = p.window > 0 ? min(max(itok - p.window, 0), Lp) : 0;
            const unsigned span = hi > lo ? static_cast<unsigned>(hi - lo) : 0u;
            const int mask_from = !use_mask ? 0x7fffffff : (p.skip_prefix_mask ? Lp : 0);
            const int jb = j0 + half * 64;
#pragma unroll
            for (int jj = 0; jj < 64; ++jj)
              s[jj] = static_cast<unsigned>(jb + jj - lo) < span ? s[jj] : -CUDART_INF_F;
            if (jb + 64 > mask_from) {
#pragma unroll
              for (int jj = 0; jj < 64; ++jj) {
                const int j = jb + jj;
                if (j >= mask_from && j < hi && mrow[j] == 0) s[jj] = -CUDART_INF_F;
              }
            }
          }
          float mx0 = -CUDART_INF_F, mx1 = -CUDART_INF_F, mx2 = -CUDART_INF_F, mx3 = -CUDART_INF_F;
#pragma unroll
          for (int j = 0; j < 64; j += 4) {
            mx0 = fmaxf(mx0, s[j]);
            mx1 = fmaxf(mx1, s[j + 1]);
            mx2 = fmaxf(mx2, s[j + 2]);
            mx3 = fmaxf(mx3, s[j + 3]);
          }
          TC_TR(sub * 3 + 1, n);
          float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
          const int par = (it & 1) * (2 * 2 * kTcBM);
          xmine[par] = mx;
          // only the two warps that share these 32 rows meet (barriers 6..13): a barrier over all eight warps
          // of the sub-tile waits for the slowest of them every tile
#ifdef B200ATTN_TC_WIDE_XCH
          named_bar_sync(1 + sub, 256);
#else
          named_bar_sync(6 + sub * 4 + (warp & 3), 64);
#endif
          mx = fmaxf(mx, xother[par]);
          if (sub == 0) TC_TR(14, n);
          const bool grow = mx > m && (m == -CUDART_INF_F || (mx - m) * c > kTcRescaleLog2);
          if (__any_sync(0xffffffffu, grow)) {
            float corr = 1.f;
            if (grow) {
              corr = fast_exp2((m - mx) * c);  // m = -inf -> 0
              m = mx;
              l *= corr;
            }
            if (it > 0) {
#pragma unroll
              for (int cc = 0; cc < OH; cc += 16) {   // 16 columns at a time: s[64] is live
                float o[16];
                tmem_ld16(tO + cc, o);
                tmem_wait_ld();
                uint32_t ou[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) ou[j] = __float_as_uint(o[j] * corr);
                tmem_st16(tO + cc, ou);
              }
            }
          }
          const float mc = (m == -CUDART_INF_F ? 0.f : m) * c;
          float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
          for (int cc = 0; cc < 64; cc += 32) {
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              const float p0 = fast_exp2(fmaf(s[cc + j], c, -mc));
              const float p1 = fast_exp2(fmaf(s[cc + j + 1], c, -mc));
              sum0 += p0;
              sum1 += p1;
              pk[j >> 1] = pack2<HALF>(p0, p1);
            }
            tmem_st16(tS + (cc >> 1), pk);
            if (kTcEarlyP && cc == 0) {
              // publish the first 32 tokens' P now: the MMA warp runs their half of P.V under the other 32 exps
              tmem_wait_st();
              tc_fence_before();
              mbar_arrive(bar_p_half + 8 * sub);
            }
          }
          l += sum0 + sum1;
          if (sub == 0) TC_TR(15, n);
          tmem_wait_st();
          tc_fence_before();
          mbar_arrive(bar_p_full + 8 * sub);
#ifdef B200ATTN_TC_TRACE
          if (trw) trw[256] = clock64();
#endif
          TC_TR(sub * 3 + 2, n);
        };
        if constexpr ((ALT && !kTcSinglePassAlt) || !kTcSinglePass) {
          if (need_mask) tile_body(std::true_type{});
          else tile_body(std::false_type{});
        } else {
          if (need_mask) tile_body1(std::true_type{});
          else tile_body1(std::false_type{});
        }
      }

      // ---- epilogue: row sum over both halves, then this thread's 64 columns of O / l -> global ----
      lxch[(sub * 2 + half) * kTcBM + r] = l;
      named_bar_sync(1 + sub, 256);
      l += lxch[(sub * 2 + (half ^ 1)) * kTcBM + r];
      // small-q mode: sub-tile 1 hands its state (max in log2 units, sum) to sub-tile 0, which merges
      float a0 = 1.f, a1 = 0.f, mlog2 = (m == -CUDART_INF_F ? 0.f : m) * c;
      if (alt) {
        if (sub == 1 && half == 0) {
          xch[2 * kTcBM + r] = m == -CUDART_INF_F ? -CUDART_INF_F : m * c;
          xch[3 * kTcBM + r] = l;
        }
        named_bar_sync(5, 512);
        if (sub == 0) {
          const float m0c = m == -CUDART_INF_F ? -CUDART_INF_F : m * c;
          const float m1c = xch[2 * kTcBM + r], l1 = xch[3 * kTcBM + r];
          const float mm = fmaxf(m0c, m1c);
          a0 = (l > 0.f) ? fast_exp2(m0c - mm) : 0.f;
          a1 = (l1 > 0.f) ? fast_exp2(m1c - mm) : 0.f;
          l = a0 * l + a1 * l1;
          mlog2 = mm == -CUDART_INF_F ? 0.f : mm;
        }
      }
      if ((!alt || sub == 0) && !dead) {
      mbar_wait_wd(bar_o_full, 0);
      tc_fence_after();
      TC_TR(13, 250);  // epilogue: O complete
      const float inv = l > 0.f ? p.v_scale / l : 0.f;
      QT* orow = reinterpret_cast<QT*>(p.o) + static_cast<int64_t>(q0 + (rv ? itok : 0)) * p.o_st +
                 static_cast<int64_t>(h * p.group + (rv ? hd : 0)) * p.o_sh + half * OH;
#pragma unroll
      for (int cc = 0; cc < OH; cc += 32) {
        float o[32];
        tmem_ld32<0>(tO + cc, o);
        tmem_wait_ld();
        if (alt) {
          // O = a0 O0 + a1 O1 in two 16-column passes (register budget); a factor of 0 drops its term
#pragma unroll
          for (int j = 0; j < 32; ++j) o[j] = a0 > 0.f ? o[j] * a0 : 0.f;
#pragma unroll
          for (int c16 = 0; c16 < 32; c16 += 16) {  // tcgen05.ld is warp-collective: no per-row branch around it
            float o1[16];
            tmem_ld16(tO + kTcD + cc + c16, o1);
            tmem_wait_ld();
#pragma unroll
            for (int j = 0; j < 16; ++j) o[c16 + j] = a1 > 0.f ? fmaf(o1[j], a1, o[c16 + j]) : o[c16 + j];
          }
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
        p.lse[static_cast<int64_t>(q0 + itok) * p.hq + h * p.group + hd] =
            l > 0.f ? mlog2 * kLn2 + logf(l) : -CUDART_INF_F;
#endif
      }
      TC_TR(13, 251);  // epilogue: outputs stored
      tc_fence_before();
    }
    }  // !WIDE
  } else {
  if (warp < kTcMmaWarp) {
    // =============================== loaders (4 warps) ===============================
    setmaxnreg_dec<kTcLoaderRegs>();
    const int tid = threadIdx.x - kTcLoaderWarp0 * 32;  // 0..127
    const int chunk = tid & 15, rbase = tid >> 4;       // 256-byte rows: row = rbase + 8 * i
    const uint32_t sw_col = static_cast<uint32_t>(chunk >> 3) * kTcSlab;
    const int chunk8 = tid & 7, rb8 = tid >> 3;         // 128-byte rows (fp8, head_dim 64): row = rb8 + 16 * i
    // ---- Q: rows of both sub-tiles ----
    {
      const int gmask = (1 << p.gp_shift) - 1;
      const int nrows = nsub * kTcBM;
      constexpr int QSTEP = D == 64 ? 16 : 8;
      const int qchunk = D == 64 ? chunk8 : chunk, qrb = D == 64 ? rb8 : rbase;
      for (int i = 0; i < nrows / QSTEP; ++i) {
        const int R = qrb + QSTEP * i;
        const int itok = tok0 + (R >> p.gp_shift), hd = R & gmask;
        const bool ok = hd < p.group && itok < Le;
        const QT* src = reinterpret_cast<const QT*>(p.q) + (ok ? static_cast<int64_t>(q0 + itok) * p.q_st +
                                                                     static_cast<int64_t>(h * p.group + hd) * p.q_sh
                                                               : 0) + qchunk * 8;
        const int rr = R & (kTcBM - 1);
        const uint32_t dst = q_base + (R >> 7) * kTcSlotBytes + (D == 64 ? 0u : sw_col) + rr * 128 +
                             (((qchunk & 7) ^ (rr & 7)) << 4);
        if constexpr (FP8 && !OUT_HALF) {  // bf16 q -> f16, clamped to the f16 range
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
      cp_async_commit();  // the oldest group: complete at the first publish below, which also releases Q
      TC_TR(12, 251);  // Q issued
    }
    bool q_pub = false;
    // ---- K / V ring ----
    const uint8_t* k_head = p.k_pool + static_cast<int64_t>(h) * D * (FP8 ? 1 : 2) + (NARROW ? chunk8 : chunk) * 16;
    const uint8_t* v_head = p.v_pool + static_cast<int64_t>(h) * D * (FP8 ? 1 : 2) + (NARROW ? chunk8 : chunk) * 16;
    // stage st is complete for this thread: (fp8: widen it in place) and hand it to the MMA warp
    auto publish = [&](int st) {
      if constexpr (WIDEN) {
        // this thread's raw bytes of stage st (and, the first time, its part of Q) are in shared memory
        if (!q_pub) {
          fence_proxy_async();
          mbar_arrive(bar_q_full);
          q_pub = true;
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
      if (!q_pub) {
        mbar_arrive(bar_q_full);
        q_pub = true;
      }
      mbar_arrive(bar_kv_full + 8 * (st % kTcSlots));
    };
    // Slot ids of a tile are read once (one per thread, coalesced) into a double-buffered shared
    // array; a 128-thread named barrier at the tile's K stage publishes them to both loader warps.
    // Slot ids run ahead of the copies: tiles n + 1 and n + 2 sit in shared memory, tile n + 3 is in flight
    // in a register (`pending`), so the index load's latency never joins the K / V copy's on the critical path.
    auto load_id = [&](int n) {
      const int j = n * kTcBN + tid;
      int32_t sl = -1;
      if (j < Ltot) sl = j < Lp ? pidx[j] : static_cast<int32_t>(p.extend_loc[q0 + j - Lp]);
      return sl;
    };
    // Layout of a slot-id buffer: the ids of the rows ONE loader thread copies are adjacent, so a stage reads
    // them with vector loads into registers BEFORE its first cp.async.  (Read one by one between the copies
    // -- each asm volatile cp.async orders the next shared load behind it -- the issue loop ran LDS ->
    // address -> LDGSTS serially, ~90 clk per copy, 1400 clk per stage: the whole kernel was bound by it.)
    const int sid_pos = NARROW ? (tid & 15) * 8 + (tid >> 4) : (tid & 7) * 16 + (tid >> 3);
    sid[sid_pos] = load_id(0);
    sid[kTcBN + sid_pos] = load_id(1);
    int32_t pending = load_id(2);
    TC_TR(12, 252);  // first slot ids stored
    const int S = 2 * N;
    for (int st = 0; st < S; ++st) {
      // publish stage st - kTcLag first: its copies have landed once at most kTcLag - 1 newer groups are
      // pending.  Doing this before the empty-slot wait keeps publication independent of the ring.
      TC_TR(13, st);
      if (st >= kTcLag) {
        cp_async_wait<kTcLag - 1>();
        publish(st - kTcLag);
        TC_TR(11, st - kTcLag);
      }
      const int slot = st % kTcSlots;
      const uint32_t ph = static_cast<uint32_t>(st / kTcSlots) & 1u;
      mbar_wait_wd(bar_kv_empty + 8 * slot, ph ^ 1u);
      TC_TR(10, st);
      // stage -> (tensor, tile): K0 V0 K1 V1 ..., or in small-q mode K0 K1 V0 K2 V1 ... K(N-1) V(N-2) V(N-1)
      bool is_v = st & 1;
      int tile = st >> 1;
      if (alt && st > 0) {
        is_v = st == S - 1 || !(st & 1);
        tile = st == S - 1 ? N - 1 : ((st & 1) ? (st + 1) >> 1 : (st >> 1) - 1);
      }
      if (!is_v) named_bar_sync(3, 128);  // slot ids of this tile are visible
      const uint8_t* base = is_v ? v_head : k_head;
      const uint32_t sb = static_cast<uint32_t>(is_v ? p.v_slot_bytes : p.k_slot_bytes);  // < 2^32 (host check): one 32 x 32 -> 64 mad per row address
      if constexpr (NARROW) {
        const int4* idv = reinterpret_cast<const int4*>(sid + (tile & 3) * kTcBN + rb8 * 8);
        const int4 i0 = idv[0], i1 = idv[1];
        const int32_t ids[8] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w};
        // fp8: raw bytes, linear, into the upper slab (widened in place at publish time);
        // head_dim 64: the row is the whole 128-byte swizzle atom of slab 0
        const uint32_t dst0 = kv_base + slot * kTcSlotBytes + (FP8 ? kTcSlab + chunk8 * 16 : 0);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int32_t sl = ids[i];
          const bool ok = sl >= 0;
          const int row = rb8 + 16 * i;
          cp_async16_zfill(dst0 + row * 128 + (FP8 ? 0 : ((chunk8 ^ (row & 7)) << 4)), base + static_cast<uint64_t>(static_cast<uint32_t>(ok ? sl : 0)) * sb, ok);
        }
      } else {
        const int4* idv = reinterpret_cast<const int4*>(sid + (tile & 3) * kTcBN + rbase * 16);
        const int4 i0 = idv[0], i1 = idv[1], i2 = idv[2], i3 = idv[3];
        const int32_t ids[16] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w,
                                 i2.x, i2.y, i2.z, i2.w, i3.x, i3.y, i3.z, i3.w};
        const uint32_t dst0 = kv_base + slot * kTcSlotBytes + sw_col;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int row = rbase + 8 * i;
          const int32_t sl = ids[i];
          const bool ok = sl >= 0;
          cp_async16_zfill(dst0 + row * 128 + (((chunk & 7) ^ (row & 7)) << 4), base + static_cast<uint64_t>(static_cast<uint32_t>(ok ? sl : 0)) * sb, ok);
        }
      }
      cp_async_commit();
      TC_TR(12, st);
      // publish the ids of tile + 2: their buffer last served tile - 2, whose V stage every loader thread had
      // issued before this thread could pass the barrier of K(tile); then start the load for tile + 3
      if (is_v && tile + 2 < N) {
        sid[((tile + 2) & 3) * kTcBN + sid_pos] = pending;
        pending = load_id(tile + 3);
      }
    }
    // drain: the last min(S, kTcLag) stages
    cp_async_wait<0>();
    for (int st = (S > kTcLag ? S - kTcLag : 0); st < S; ++st) publish(st);
    TC_TR(13, 252);  // loader done
  } else if (warp == kTcMmaWarp) {
    // =============================== MMA issuer (warp 20) ===============================
    setmaxnreg_dec<kTcMmaRegs>();
    {
#ifdef B200ATTN_TC_TRACE
      tr = (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0) ? reinterpret_cast<long long*>(p.lse) : nullptr;
#endif
      // descriptors differ only in the 14-bit start-address field: build one base per operand kind and
      // add the (compile-time) offsets, in 16-byte units
      const uint64_t dq = umma_desc_sw128(q_base, 16, 1024);
      const uint64_t dk = umma_desc_sw128(kv_base, 16, 1024);
      const uint64_t dv = umma_desc_sw128(kv_base, kTcSlab, 1024);
      auto issue_qk = [&](int sub, int slot) {
        const uint64_t a0 = dq + static_cast<uint32_t>((alt ? 0 : sub) * (kTcSlotBytes >> 4));
        const uint64_t b0 = dk + static_cast<uint32_t>(slot * (kTcSlotBytes >> 4));
        const uint32_t d = tmem + sub * kTcBN;
#pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint32_t off = ((k >> 2) * kTcSlab + (k & 3) * 32) >> 4;
          umma_ss(d, a0 + off, b0 + off, IDESC_QK, k > 0);
        }
      };
      auto issue_pv = [&](int sub, int slot, bool acc) {
        const uint64_t v0 = dv + static_cast<uint32_t>(slot * (kTcSlotBytes >> 4));
        const uint32_t d = tmem + 2 * kTcBN + sub * kTcD, a = tmem + sub * kTcBN;
        umma_ts(d, a, v0, IDESC_PV, acc ? 1u : 0u);
#pragma unroll
        for (int k = 1; k < kTcBN / 16; ++k)  // P: columns [0, 32) hold tokens 0-63, [64, 96) tokens 64-127
          umma_ts(d, a + k * 8 + (k >= 4 ? 32 : 0), v0 + k * (2048 >> 4), IDESC_PV, 1u);
      };
      // P.V of the k-steps whose 16 tokens lie in every thread's first (which = 0) / second 32-token chunk:
      // half h of a row holds tokens 64 h .., so chunk c is k-steps 4 h + 2 c, 4 h + 2 c + 1
      auto issue_pv_part = [&](int sub, int slot, int which, bool acc) {
        const uint64_t v0 = dv + static_cast<uint32_t>(slot * (kTcSlotBytes >> 4));
        const uint32_t d = tmem + 2 * kTcBN + sub * kTcD, a = tmem + sub * kTcBN;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          // WIDE: a row's packed P is contiguous (token t -> column t / 2), its first half is k-steps 0..3
          const int k = WIDE ? which * 4 + i : (i >> 1) * 4 + which * 2 + (i & 1);
          umma_ts(d, a + k * 8 + ((!WIDE && k >= 4) ? 32 : 0), v0 + k * (2048 >> 4), IDESC_PV, (acc || i > 0) ? 1u : 0u);
        }
      };
      auto issue_pv_tile = [&](int sub, int slot, int n) {
        if constexpr (kTcEarlyP && !ALT) {
          mbar_wait_wd(bar_p_half + 8 * sub, n & 1);
          tc_fence_after();
          issue_pv_part(sub, slot, 0, n > 0);
          mbar_wait_wd(bar_p_full + 8 * sub, n & 1);
          tc_fence_after();
          if (sub == 0) TC_TR(7, n);
          else TC_TR(9, n);
          issue_pv_part(sub, slot, 1, true);
        } else {
          mbar_wait_wd(bar_p_full + 8 * sub, n & 1);
          tc_fence_after();
          if (sub == 0) TC_TR(7, n);
          else TC_TR(9, n);
          issue_pv(sub, slot, n > 0);
        }
      };
      auto full_wait = [&](int st) {
        mbar_wait_wd(bar_kv_full + 8 * (st % kTcSlots), static_cast<uint32_t>(st / kTcSlots) & 1u);
      };
      mbar_wait_wd(bar_q_full, 0);
      full_wait(0);
      tc_fence_after();
      issue_qk(0, 0);
      umma_commit(bar_s_full);
      if (alt) {
        // one chain: S0(0) S1(1) | PV_v(j) S_v(j+2) with v = j & 1 -- every stage feeds one sub-tile
        umma_commit(bar_kv_empty);
        full_wait(1);
        tc_fence_after();
        issue_qk(1, 1);
        umma_commit(bar_s_full + 8);
        umma_commit(bar_kv_empty + 8);
        for (int j = 0; j < N; ++j) {
          const int v = j & 1;
          const int vst = j < N - 1 ? 2 * j + 2 : 2 * N - 1;
          full_wait(vst);
          TC_TR(6, j);
          mbar_wait_wd(bar_p_full + 8 * v, (j >> 1) & 1);
          tc_fence_after();
          TC_TR(7, j);
          issue_pv(v, vst % kTcSlots, j >= 2);
          umma_commit(bar_kv_empty + 8 * (vst % kTcSlots));
          if (j + 2 < N) {
            const int kst = 2 * j + 3;
            full_wait(kst);
            tc_fence_after();
            TC_TR(8, j);
            issue_qk(v, kst % kTcSlots);
            umma_commit(bar_s_full + 8 * v);
            umma_commit(bar_kv_empty + 8 * (kst % kTcSlots));
          }
        }
      } else {
      if (nsub == 2) {
        issue_qk(1, 0);
        umma_commit(bar_s_full + 8);
      }
      umma_commit(bar_kv_empty);  // K_0's slot
      for (int n = 0; n < N; ++n) {
        const int vst = 2 * n + 1, kst = 2 * n + 2;
        const int vslot = vst % kTcSlots, kslot = kst % kTcSlots;
        const bool more = n + 1 < N;
        full_wait(vst);
        TC_TR(6, n);
        issue_pv_tile(0, vslot, n);
        if (more) {
          full_wait(kst);
          tc_fence_after();
          TC_TR(8, n);
          issue_qk(0, kslot);
          umma_commit(bar_s_full);
        }
        if (nsub == 2) issue_pv_tile(1, vslot, n);
        umma_commit(bar_kv_empty + 8 * vslot);
        if (more) {
          if (nsub == 2) {
            issue_qk(1, kslot);
            umma_commit(bar_s_full + 8);
          }
          umma_commit(bar_kv_empty + 8 * kslot);
        }
      }
      }
      umma_commit(bar_o_full);
      TC_TR(13, 253);  // last MMA committed
      tc_fence_before();
    }
  } else if constexpr (WIDEN) {
    // =============================== fp8 widening warps (21-23) ===============================
    // A stage holds 128 raw rows of 128 fp8 bytes in its upper slab; row r widens to 256 bytes = row r of both
    // slabs, in the swizzled layout.  Thread (row group rb, chunk c) turns 16 bytes into two 16-byte f16 chunks.
    // The outputs of a row's chunks 4-7 land on that row's own raw bytes and nowhere else, and the row's eight
    // chunks sit in eight adjacent lanes: a __syncwarp between the loads and the stores is all the ordering needed.
    setmaxnreg_dec<48>();
    const int wt = threadIdx.x - (kTcMmaWarp + 1) * 32;  // 0..95
    const int chunk8 = wt & 7, rb = wt >> 3;             // rows rb + 12 i
    const int S = 2 * N;
    for (int st = 0; st < S; ++st) {
      const int slot = st % kTcSlots;
      mbar_wait_wd(bar_raw_full + 8 * slot, static_cast<uint32_t>(st / kTcSlots) & 1u);
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
  } else {
    setmaxnreg_dec<24>();   // warps 21-23 only lend their registers
  }
  }
  __syncthreads();
#ifdef B200ATTN_TC_TRACE
  if (threadIdx.x == 0) {
    unsigned long long gt;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(gt));
    cta_tr[1] = static_cast<long long>(gt);
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) reinterpret_cast<long long*>(p.lse)[13 * 256 + 255] = clock64();
  }
#endif
  if (warp == kTcMmaWarp) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
  }
}

}  // namespace b200attn
