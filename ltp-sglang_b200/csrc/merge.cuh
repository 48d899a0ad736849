#pragma once
#include <math_constants.h>

#include "common.cuh"

namespace b200attn {

struct MergeParams {
  const void* plan;
  const float* ws_o;
  const float* ws_lse;
  void* o;
  int64_t o_sb, o_sh;
  float* lse;
  int32_t batch, hq, ng, ts, dv;
};

cudaError_t launch_merge(const MergeParams& p, bool half, cudaStream_t stream);

#ifdef __CUDACC__
// N-way LSE merge of one (request, q head) by one warp: partial i lives at workspace row
// row0 + i * row_step (normalised fp32 o[dv] and lse = ln sum exp).  Same recurrence as the
// reference's stage 2 (triton_ops/decode_attention.py:529-552), evaluated as max / weights / sum.
// Reads go through L2 (ld.global.cg): the rows were written by other SMs moments ago.
template <typename QT>
__device__ __forceinline__ void merge_rows(const float* __restrict__ ws_o, const float* __restrict__ ws_lse,
                                           int64_t row0, int row_step, int np, int dv, QT* out, float* lse_out,
                                           int lane) {
  // Online form, four partials per round so their loads (lse and o rows alike) fly together: the
  // merge sits on a CTA's critical path, and each dependent round trip to L2 costs ~1 us under load.
  const int d = lane * 4;  // dv <= 128: one float4 per lane
  const bool act = d < dv;
  float m = -CUDART_INF_F, s = 0.f;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int i0 = 0; i0 < np; i0 += 4) {
    float l[4];
    float4 x[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const bool ok = i0 + k < np;
      const int64_t row = row0 + static_cast<int64_t>(ok ? i0 + k : i0) * row_step;
      l[k] = ok ? __ldcg(ws_lse + row) : -CUDART_INF_F;
      x[k] = (ok && act) ? __ldcg(reinterpret_cast<const float4*>(ws_o + row * dv + d)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float mn = fmaxf(fmaxf(fmaxf(l[0], l[1]), fmaxf(l[2], l[3])), m);
    if (mn == -CUDART_INF_F) continue;  // nothing but empty partials so far
    const float sc = __expf(m - mn);
    s *= sc;
    acc.x *= sc;
    acc.y *= sc;
    acc.z *= sc;
    acc.w *= sc;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float w = __expf(l[k] - mn);  // 0 for an empty partial (lse = -inf)
      s += w;
      acc.x = fmaf(w, x[k].x, acc.x);
      acc.y = fmaf(w, x[k].y, acc.y);
      acc.z = fmaf(w, x[k].z, acc.z);
      acc.w = fmaf(w, x[k].w, acc.w);
    }
    m = mn;
  }
  const float inv = s > 0.f ? 1.f / s : 0.f;
  if (act) {
    out[d] = QT(acc.x * inv);
    out[d + 1] = QT(acc.y * inv);
    out[d + 2] = QT(acc.z * inv);
    out[d + 3] = QT(acc.w * inv);
  }
  if (lse_out && lane == 0) *lse_out = s > 0.f ? m + logf(s) : -CUDART_INF_F;
}
#endif

}  // namespace b200attn
