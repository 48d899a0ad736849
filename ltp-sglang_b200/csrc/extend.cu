// C-ABI entry of the extend (prefill with cached prefix) attention path: argument checks and the
// launch of extend_kernel (extend_kernel.cuh).
#include <cstdlib>

#include "extend_tc_smallq.cuh"

using namespace b200attn;

namespace {

template <typename QT, bool FP8>
cudaError_t launch_extend(const ExtendParams& p, int d, dim3 grid, cudaStream_t stream) {
  void (*fn)(const ExtendParams) = nullptr;
  const bool cap = p.logit_cap > 0.f;
  if (d == 64) fn = cap ? extend_kernel<QT, 64, FP8, true> : extend_kernel<QT, 64, FP8, false>;
  else if (d == 128) fn = cap ? extend_kernel<QT, 128, FP8, true> : extend_kernel<QT, 128, FP8, false>;
  else if (d == 256) fn = cap ? extend_kernel<QT, 256, FP8, true> : extend_kernel<QT, 256, FP8, false>;   // Gemma-sized heads
  if constexpr (!FP8) {   // head dims 80 / 96 (16-bit pools, as in the decode kernel)
    if (d == 80) fn = cap ? extend_kernel<QT, 80, false, true> : extend_kernel<QT, 80, false, false>;
    else if (d == 96) fn = cap ? extend_kernel<QT, 96, false, true> : extend_kernel<QT, 96, false, false>;
  }
  if (!fn) return cudaErrorInvalidValue;
  const int smem = extend_smem_bytes(d);
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  fn<<<grid, kExtThreads, smem, stream>>>(p);
  return cudaGetLastError();
}

// The general kernel runs one thread per row (WIDE, 16 warps); B200ATTN_TC_LAYOUT=two picks the older
// two-threads-per-row instantiation (24 warps) for A/B runs and the tests.  The small-q mode always has two.
static bool tc_wide() {
  static const bool wide = [] {
    const char* e = std::getenv("B200ATTN_TC_LAYOUT");
    return !(e && e[0] == 't');
  }();
  return wide;
}

template <typename QT, bool FP8>
cudaError_t launch_extend_tc(const ExtendParams& p, int d, dim3 grid, bool small_q, cudaStream_t stream) {
  void (*fn)(const ExtendParams);
  const bool wide = !small_q && !FP8 && tc_wide();   // (fp8 pools: measured a shade slower, P2f 775 against 789 TFLOP/s)
#define B200_TC_PICK(DD, F8, CP)                                                                      \
  (small_q ? extend_tc_kernel<QT, DD, F8, true, CP, false>                                            \
           : (wide ? extend_tc_kernel<QT, DD, F8, false, CP, true> : extend_tc_kernel<QT, DD, F8, false, CP, false>))
  if constexpr (!FP8) {
    if (d == 64) {  // head_dim 64: 16-bit pools without a logit cap (the caller checked)
      fn = B200_TC_PICK(64, false, false);
    } else if (p.logit_cap > 0.f) {
      fn = B200_TC_PICK(128, false, true);
    } else {
      fn = B200_TC_PICK(128, false, false);
    }
  } else if (p.logit_cap > 0.f) {
    fn = B200_TC_PICK(128, true, true);
  } else {
    fn = B200_TC_PICK(128, true, false);
  }
#undef B200_TC_PICK
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes);
  if (e != cudaSuccess) return e;
  fn<<<grid, wide ? kTcWideThreads : kTcLaunchThreads, kTcSmemBytes, stream>>>(p);
  return cudaGetLastError();
}

// Small-q batches (no request has more than 128 (token, head) rows): the persistent kernel of
// extend_tc_smallq.cuh, one CTA per SM looping over (request, kv head) items.
template <typename QT, bool FP8>
cudaError_t launch_extend_tc_smallq(const ExtendParams& p, int d, int items, cudaStream_t stream) {
  void (*fn)(const ExtendParams);
  if constexpr (!FP8) {
    if (d == 64) fn = extend_tc_smallq_kernel<QT, 64, false, false>;
    else fn = p.logit_cap > 0.f ? extend_tc_smallq_kernel<QT, 128, false, true> : extend_tc_smallq_kernel<QT, 128, false, false>;
  } else {
    fn = p.logit_cap > 0.f ? extend_tc_smallq_kernel<QT, 128, true, true> : extend_tc_smallq_kernel<QT, 128, true, false>;
  }
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, kSqSmemBytes);
  if (e != cudaSuccess) return e;
  const int sms = sm_count();
  fn<<<items < sms ? items : sms, FP8 ? kTcLaunchThreads : kTcThreads, kSqSmemBytes, stream>>>(p);
  return cudaGetLastError();
}

// Small-q batches have two kernels (measured, tools/time_smallq.py, profiles/r01_smallq_persistent_vs_oneshot.txt):
// the persistent one wins wherever per-item overhead matters -- prefixes up to ~1k tokens (1.05-1.27x), every
// fp8 pool (1.02-1.55x: the in-place widening hides behind the next item's loads) -- the one-CTA-per-item kernel
// on long 16-bit prefixes (1.07-1.17x at 4k-16k tokens: its loader loop is leaner; at 80 registers per thread the
// persistent loader's extra cursor state costs ~500 instructions per stage).  The host does not know the prefix
// lengths (device arrays) unless the caller passes avg_kv_len_hint (sglang has seq_lens_sum on the host): the
// default goes by pool type and, when given, by the hint; B200ATTN_SMALLQ=persistent|oneshot forces one
// (a test / tuning aid; read once).
int smallq_mode() {
  static const int v = [] {
    const char* e = getenv("B200ATTN_SMALLQ");
    return !e ? 0 : e[0] == 'p' ? 1 : e[0] == 'o' ? 2 : 0;
  }();
  return v;
}

// B200ATTN_EXTEND_IMPL=mma forces the mma.sync kernel (a test / tuning aid; read once).
bool extend_force_mma() {
  static const bool v = [] {
    const char* e = getenv("B200ATTN_EXTEND_IMPL");
    return e && e[0] == 'm';
  }();
  return v;
}

struct ExtendPlanParams {
  const void* extend_seq_lens;
  const void* prefix_lens;
  int32_t idx64, batch, uniform, window;
  int32_t* qo_indptr;
  int32_t* kv_indptr;
  int64_t* mask_indptr;
};

__global__ void __launch_bounds__(kPlanThreads) extend_plan_kernel(const ExtendPlanParams p) {
  __shared__ int warp_sums[32];
  __shared__ long long warp_sums64[32];
  __shared__ int carry_q, carry_kv;
  __shared__ long long carry_m;
  const int tid = threadIdx.x;
  if (tid == 0) {
    carry_q = carry_kv = 0;
    carry_m = 0;
    p.qo_indptr[0] = 0;
    p.kv_indptr[0] = 0;
    if (p.mask_indptr) p.mask_indptr[0] = 0;
  }
  __syncthreads();
  for (int base = 0; base < p.batch; base += kPlanThreads) {
    const int b = base + tid;
    int le = 0, lp = 0;
    long long pl = 0;
    if (b < p.batch) {
      le = p.uniform > 0 ? p.uniform : static_cast<int>(load_idx(p.extend_seq_lens, b, p.idx64));
      pl = load_idx(p.prefix_lens, b, p.idx64);
      if (le < 0) le = 0;
      if (pl < 0) pl = 0;
      lp = static_cast<int>(pl);
      if (p.window > 0 && lp > p.window + 1) lp = p.window + 1;
    }
    const int q_incl = block_scan_incl(le, warp_sums);
    const int kv_incl = block_scan_incl(lp, warp_sums);
    long long m_incl = 0;
    if (p.mask_indptr) m_incl = block_scan_incl(static_cast<long long>(le) * (pl + le), warp_sums64);
    if (b < p.batch) {
      p.qo_indptr[b + 1] = carry_q + q_incl;
      p.kv_indptr[b + 1] = carry_kv + kv_incl;
      if (p.mask_indptr) p.mask_indptr[b + 1] = carry_m + m_incl;
    }
    __syncthreads();
    if (tid == kPlanThreads - 1) {
      carry_q += q_incl;
      carry_kv += kv_incl;
      carry_m += m_incl;
    }
    __syncthreads();
  }
}

}  // namespace

extern "C" __attribute__((visibility("default"))) int b200attn_extend_plan(const b200attn_extend_plan_args* a,
                                                                           b200attn_stream_t stream_) {
  if (!a || a->batch < 0 || !a->qo_indptr || !a->kv_indptr || !a->prefix_lens) return B200ATTN_E_INVALID;
  if (!a->extend_seq_lens && a->uniform_extend_len <= 0) return B200ATTN_E_INVALID;
  if (a->index_dtype != B200ATTN_I32 && a->index_dtype != B200ATTN_I64) return B200ATTN_E_INVALID;
  ExtendPlanParams p;
  p.extend_seq_lens = a->extend_seq_lens;
  p.prefix_lens = a->prefix_lens;
  p.idx64 = a->index_dtype == B200ATTN_I64;
  p.batch = a->batch;
  p.uniform = a->uniform_extend_len > 0 ? a->uniform_extend_len : 0;
  p.window = a->sliding_window > 0 ? a->sliding_window : 0;
  p.qo_indptr = a->qo_indptr;
  p.kv_indptr = a->kv_indptr;
  p.mask_indptr = a->mask_indptr;
  extend_plan_kernel<<<1, kPlanThreads, 0, static_cast<cudaStream_t>(stream_)>>>(p);
  if (cudaGetLastError() != cudaSuccess) return B200ATTN_E_CUDA;
  count_launch();
  return B200ATTN_OK;
}

extern "C" __attribute__((visibility("default"))) int b200attn_extend(const b200attn_extend_args* a,
                                                                      b200attn_stream_t stream_) {
  if (!a) return B200ATTN_E_INVALID;
  if (a->batch < 0 || a->total_tokens < 0 || a->max_extend_len < 0) return B200ATTN_E_INVALID;
  if (a->batch == 0 || a->total_tokens == 0 || a->max_extend_len == 0) return B200ATTN_OK;
  if (!a->q || !a->o || !a->k_pool || !a->v_pool || !a->qo_indptr || !a->kv_indptr || !a->extend_loc)
    return B200ATTN_E_INVALID;
  if (!a->kv_indices && (!a->req_to_token || !a->req_pool_indices || !a->prefix_lens)) return B200ATTN_E_INVALID;
  if (a->custom_mask && !a->mask_indptr) return B200ATTN_E_INVALID;
  if ((a->index_dtype != B200ATTN_I32 && a->index_dtype != B200ATTN_I64) ||
      (a->mask_indptr_dtype != B200ATTN_I32 && a->mask_indptr_dtype != B200ATTN_I64))
    return B200ATTN_E_INVALID;
  if (a->q_dtype != B200ATTN_BF16 && a->q_dtype != B200ATTN_F16) return B200ATTN_E_UNSUPPORTED;
  const bool fp8 = is_fp8(a->kv_dtype);
  if (!fp8 && a->kv_dtype != a->q_dtype) return B200ATTN_E_UNSUPPORTED;
  const int esz = fp8 ? 1 : 2;
  const float k_scale = fp8 ? a->k_scale : 1.f, v_scale = fp8 ? a->v_scale : 1.f;
  if (fp8 && (!(k_scale > 0.f) || !(v_scale > 0.f))) return B200ATTN_E_INVALID;
  if (a->v_head_dim != a->head_dim) return B200ATTN_E_UNSUPPORTED;
  if (a->head_dim != 64 && a->head_dim != 128 && a->head_dim != 256 &&
      !(!fp8 && (a->head_dim == 80 || a->head_dim == 96)))   // the decode kernel's head dims
    return B200ATTN_E_UNSUPPORTED;
  if (a->num_q_heads <= 0 || a->num_kv_heads <= 0 || a->num_q_heads % a->num_kv_heads != 0)
    return B200ATTN_E_INVALID;
  const int group = a->num_q_heads / a->num_kv_heads;
  if (group > 16) return B200ATTN_E_UNSUPPORTED;
  if (a->num_kv_heads > 65535 || a->batch > 65535) return B200ATTN_E_UNSUPPORTED;
  if (!(a->sm_scale > 0.f) || a->logit_cap < 0.f) return B200ATTN_E_INVALID;
  if ((reinterpret_cast<uintptr_t>(a->k_pool) | reinterpret_cast<uintptr_t>(a->v_pool)) & 15) return B200ATTN_E_ALIGN;
  if ((a->k_pool_stride_slot * esz) % 16 != 0 || (a->v_pool_stride_slot * esz) % 16 != 0) return B200ATTN_E_ALIGN;
  // the loaders form a row address as base + slot (u32) x slot bytes (u32)
  if (a->k_pool_stride_slot <= 0 || a->v_pool_stride_slot <= 0 || a->k_pool_stride_slot * esz > 0xffffffffll ||
      a->v_pool_stride_slot * esz > 0xffffffffll)
    return B200ATTN_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(a->q) & 3) || a->q_stride_token % 2 != 0 || a->q_stride_head % 2 != 0)
    return B200ATTN_E_ALIGN;
  if ((reinterpret_cast<uintptr_t>(a->o) & 3) || a->o_stride_token % 2 != 0 || a->o_stride_head % 2 != 0)
    return B200ATTN_E_ALIGN;

  ExtendParams p;
  p.q = a->q;
  p.q_st = a->q_stride_token;
  p.q_sh = a->q_stride_head;
  p.o = a->o;
  p.o_st = a->o_stride_token;
  p.o_sh = a->o_stride_head;
  p.lse = a->lse;
  p.k_pool = static_cast<const uint8_t*>(a->k_pool);
  p.v_pool = static_cast<const uint8_t*>(a->v_pool);
  p.k_slot_bytes = a->k_pool_stride_slot * esz;
  p.v_slot_bytes = a->v_pool_stride_slot * esz;
  p.qo_indptr = a->qo_indptr;
  p.kv_indptr = a->kv_indptr;
  p.kv_indices = a->kv_indices;
  p.req_to_token = a->req_to_token;
  p.r2t_stride = a->req_to_token_stride;
  p.req_pool_indices = a->req_pool_indices;
  p.prefix_lens = a->prefix_lens;
  p.extend_loc = a->extend_loc;
  p.custom_mask = a->custom_mask;
  p.mask_indptr = a->mask_indptr;
  p.idx64 = a->index_dtype == B200ATTN_I64;
  p.mask64 = a->mask_indptr_dtype == B200ATTN_I64;
  p.hq = a->num_q_heads;
  p.group = group;
  p.gp_shift = 0;
  while ((1 << p.gp_shift) < group) ++p.gp_shift;
  p.causal = a->causal != 0;
  p.window = a->sliding_window > 0 ? a->sliding_window : 0;
  p.skip_prefix_mask = a->skip_prefix_custom_mask != 0;
  p.sm_scale = a->sm_scale * k_scale;  // fp8: K is read unscaled, k_scale rides in the softmax scale
  p.scale_log2 = p.sm_scale * kLog2e;
  p.logit_cap = a->logit_cap;
  p.v_scale = v_scale;
  p.e5m2 = a->kv_dtype == B200ATTN_FP8_E5M2;
  p.batch = a->batch;

  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  // head_dim 128 (and 64 over a 16-bit pool without logit cap) with 16-byte aligned q / o rows runs on tcgen05
  // (extend_tc_kernel.cuh); everything else on the mma.sync kernel.
  const bool tc_ok = (a->head_dim == kTcD || (a->head_dim == 64 && !fp8 && !(a->logit_cap > 0.f))) &&
                     !extend_force_mma() &&
                     !((reinterpret_cast<uintptr_t>(a->q) | reinterpret_cast<uintptr_t>(a->o)) & 15) &&
                     a->q_stride_token % 8 == 0 && a->q_stride_head % 8 == 0 && a->o_stride_token % 8 == 0 &&
                     a->o_stride_head % 8 == 0;
  cudaError_t err;
  if (tc_ok) {
    const int tok = (2 * kTcBM) >> p.gp_shift;
    dim3 grid((a->max_extend_len + tok - 1) / tok, a->num_kv_heads, a->batch);
    // every request has at most 128 (token, head) rows -- decode-like and TARGET_VERIFY batches: the small-q
    // kernel splits each CTA's KV range over its two accumulators instead of leaving one idle
    const bool small_q = (static_cast<int64_t>(a->max_extend_len) << p.gp_shift) <= kTcBM;
    // auto: fp8 pools (with the widening warps the persistent kernel is at or ahead of the one-shot one at every
    // prefix length: 389 / 397 us at 4096, 391 / 377 at 16384, 371 / 498 at 512), and 16-bit pools whose mean KV
    // length the host gave as at most 768 tokens (the one-shot kernel of round 2's second pass wins from ~1k on:
    // 277 / 282 / 345 us against 277 / 305 / 385 at 1024 / 1536 / 2048; 360 against 324 at 512)
    const bool sq_persistent = smallq_mode() == 1 ||
                               (smallq_mode() == 0 && (fp8 || (a->avg_kv_len_hint > 0 && a->avg_kv_len_hint <= 768)));
    if (small_q && sq_persistent) {
      const int items = a->batch * a->num_kv_heads;
      if (fp8)
        err = a->q_dtype == B200ATTN_F16 ? launch_extend_tc_smallq<__half, true>(p, a->head_dim, items, stream)
                                         : launch_extend_tc_smallq<__nv_bfloat16, true>(p, a->head_dim, items, stream);
      else
        err = a->q_dtype == B200ATTN_F16 ? launch_extend_tc_smallq<__half, false>(p, a->head_dim, items, stream)
                                         : launch_extend_tc_smallq<__nv_bfloat16, false>(p, a->head_dim, items, stream);
    } else if (fp8)
      err = a->q_dtype == B200ATTN_F16 ? launch_extend_tc<__half, true>(p, a->head_dim, grid, small_q, stream)
                                       : launch_extend_tc<__nv_bfloat16, true>(p, a->head_dim, grid, small_q, stream);
    else
      err = a->q_dtype == B200ATTN_F16 ? launch_extend_tc<__half, false>(p, a->head_dim, grid, small_q, stream)
                                       : launch_extend_tc<__nv_bfloat16, false>(p, a->head_dim, grid, small_q, stream);
  } else {
    const int bm_tok = kExtRows >> p.gp_shift;
    dim3 grid((a->max_extend_len + bm_tok - 1) / bm_tok, a->num_kv_heads, a->batch);
    if (fp8)
      err = a->q_dtype == B200ATTN_F16 ? launch_extend<__half, true>(p, a->head_dim, grid, stream)
                                       : launch_extend<__nv_bfloat16, true>(p, a->head_dim, grid, stream);
    else
      err = a->q_dtype == B200ATTN_F16 ? launch_extend<__half, false>(p, a->head_dim, grid, stream)
                                       : launch_extend<__nv_bfloat16, false>(p, a->head_dim, grid, stream);
  }
  if (err != cudaSuccess) return B200ATTN_E_CUDA;
  count_launch();
  return B200ATTN_OK;
}
