// K3: LSE merge of the split-KV partials (stage 2), and the stand-alone 2-way
// state merge.
//
// Reference: _fwd_kernel_stage2,
// python/sglang/srt/layers/attention/triton_ops/decode_attention.py:491-552
// (online merge: m' = max(m, lse_i), acc = acc e^{m-m'} + e^{lse_i-m'} o_i);
// merge_state_v2, sgl-kernel/csrc/attention/merge_attn_states.cu:31-106.
#include <math_constants.h>

#include "merge.cuh"

namespace b200attn {

// One warp per (request, q head).  Pieces of virtual request v live at workspace rows
// ((v + r) * ts + slice) * hqv + head_in_vreq for every range r that touches the request
// (plan: vreq_rfirst / vreq_pieces).  Stand-alone form of the merge decode_kernel runs itself when
// fuse_merge is set; kept for phases = 2 (timing / testing each stage alone).
template <typename QT>
__global__ void __launch_bounds__(128) merge_kernel(const MergeParams p) {
  const PlanView pv = plan_view(const_cast<void*>(p.plan), p.batch);
  const int warp_global = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (warp_global >= p.batch * p.hq) return;
  const int b = warp_global / p.hq;
  const int hq = warp_global - b * p.hq;
  const int hqv = p.hq / p.ng;
  const int hg = hq / hqv;
  const int hl = hq - hg * hqv;
  const int v = b * p.ng + hg;
  QT* out = reinterpret_cast<QT*>(p.o) + b * p.o_sb + hq * p.o_sh;
  float* lse_out = p.lse ? p.lse + static_cast<int64_t>(b) * p.hq + hq : nullptr;
  const int pieces = pv.vreq_pieces[v];
  if (pieces == 0) {  // empty request: defined output, no pieces were written
    for (int d = lane; d < p.dv; d += 32) out[d] = QT(0.f);
    if (lse_out && lane == 0) *lse_out = -CUDART_INF_F;
    return;
  }
  if (p.ts == 1 && pieces == 1) return;  // decode_kernel wrote o directly
  const int64_t row0 = static_cast<int64_t>(v + pv.vreq_rfirst[v]) * p.ts * hqv + hl;
  merge_head<QT>(p.ws_o, p.ws_lse, row0, hqv, pieces * p.ts, p.dv, out, lse_out, lane);
}

template __global__ void merge_kernel<__nv_bfloat16>(const MergeParams);
template __global__ void merge_kernel<__half>(const MergeParams);

cudaError_t launch_merge(const MergeParams& p, bool half, cudaStream_t stream) {
  const int warps = p.batch * p.hq;
  const int blocks = (warps + 3) / 4;
  if (half) merge_kernel<__half><<<blocks, 128, 0, stream>>>(p);
  else merge_kernel<__nv_bfloat16><<<blocks, 128, 0, stream>>>(p);
  return cudaGetLastError();
}

// 2-way merge; one thread per 128-bit pack (8 half / bf16 elements, 4 floats: the reference dispatches
// float, half and bf16, merge_attn_states.cu:111-122).  +inf lse counts as -inf (merge_attn_states.cu:68-69).
__device__ __forceinline__ float ms_to_f32(float x) { return x; }
__device__ __forceinline__ float ms_to_f32(__half x) { return __half2float(x); }
__device__ __forceinline__ float ms_to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }
template <typename QT>
__device__ __forceinline__ QT ms_from_f32(float x) { return QT(x); }

template <typename QT>
__global__ void __launch_bounds__(128) merge_states_kernel(const QT* __restrict__ oa,
                                                          const float* __restrict__ la,
                                                          const QT* __restrict__ ob,
                                                          const float* __restrict__ lb,
                                                          QT* __restrict__ out,
                                                          float* __restrict__ lout, int tokens,
                                                          int heads, int dim) {
  constexpr int PACK = 16 / sizeof(QT);
  const int packs = dim / PACK;
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= static_cast<int64_t>(tokens) * heads * packs) return;
  const int64_t th = i / packs;
  const int pk = static_cast<int>(i - th * packs);
  float a = la[th], b = lb[th];
  if (isinf(a)) a = -CUDART_INF_F;
  if (isinf(b)) b = -CUDART_INF_F;
  const float mx = fmaxf(a, b);
  const float ea = expf(a - mx), eb = expf(b - mx);
  const float se = ea + eb;
  const float wa = ea / se, wb = eb / se;
  const int64_t off = th * dim + pk * PACK;
  QT xa[PACK], xb[PACK], y[PACK];
  *reinterpret_cast<uint4*>(xa) = *reinterpret_cast<const uint4*>(oa + off);
  *reinterpret_cast<uint4*>(xb) = *reinterpret_cast<const uint4*>(ob + off);
#pragma unroll
  for (int e = 0; e < PACK; ++e) y[e] = ms_from_f32<QT>(ms_to_f32(xa[e]) * wa + ms_to_f32(xb[e]) * wb);
  *reinterpret_cast<uint4*>(out + off) = *reinterpret_cast<const uint4*>(y);
  if (lout && pk == 0) lout[th] = logf(se) + mx;
}

template <typename QT>
static void launch_merge_states(const b200attn_merge_states_args* a, cudaStream_t s) {
  const int64_t n = static_cast<int64_t>(a->tokens) * a->heads * (a->head_dim / (16 / sizeof(QT)));
  const unsigned blocks = static_cast<unsigned>((n + 127) / 128);
  merge_states_kernel<QT><<<blocks, 128, 0, s>>>(static_cast<const QT*>(a->o_a), a->lse_a, static_cast<const QT*>(a->o_b),
                                                 a->lse_b, static_cast<QT*>(a->o_out), a->lse_out, a->tokens, a->heads,
                                                 a->head_dim);
}

}  // namespace b200attn

using namespace b200attn;

extern "C" __attribute__((visibility("default"))) int b200attn_merge_states(const b200attn_merge_states_args* a, b200attn_stream_t s_) {
  if (!a || !a->o_a || !a->o_b || !a->lse_a || !a->lse_b || !a->o_out) return B200ATTN_E_INVALID;
  if (a->tokens < 0 || a->heads <= 0 || a->head_dim <= 0) return B200ATTN_E_INVALID;
  if (a->dtype != B200ATTN_BF16 && a->dtype != B200ATTN_F16 && a->dtype != B200ATTN_F32) return B200ATTN_E_UNSUPPORTED;
  if (a->head_dim % (a->dtype == B200ATTN_F32 ? 4 : 8) != 0) return B200ATTN_E_UNSUPPORTED;  // 128-bit packs (:185-202)
  if ((reinterpret_cast<uintptr_t>(a->o_a) | reinterpret_cast<uintptr_t>(a->o_b) |
       reinterpret_cast<uintptr_t>(a->o_out)) & 15)
    return B200ATTN_E_ALIGN;
  if (a->tokens == 0) return B200ATTN_OK;
  cudaStream_t s = static_cast<cudaStream_t>(s_);
  if (a->dtype == B200ATTN_BF16) launch_merge_states<__nv_bfloat16>(a, s);
  else if (a->dtype == B200ATTN_F16) launch_merge_states<__half>(a, s);
  else launch_merge_states<float>(a, s);
  count_launch();
  return cudaGetLastError() == cudaSuccess ? B200ATTN_OK : B200ATTN_E_CUDA;
}
