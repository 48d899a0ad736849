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
//
// The merge sits on the critical path of the CTA that lands a request's last partial -- at the very end of
// the kernel when a request spans many CTAs -- and every dependent round trip to L2 costs 0.5-1 us
// (tools/trace_decode.py: C1's 16 merging CTAs ran 12 us longer than the others with 4 loads in flight).
// So R partials are loaded per round, all loads of a round issued before the first use; the caller picks R
// against the number of partials (merge_heads).  dv <= 128 (one float4 per lane) or 256 (two).
template <typename QT, int R, int V>
__device__ __forceinline__ void merge_rows(const float* __restrict__ ws_o, const float* __restrict__ ws_lse,
                                           int64_t row0, int64_t row_step, int np, int dv, QT* out, float* lse_out,
                                           int lane) {
  float m = -CUDART_INF_F, s = 0.f;
  float4 acc[V];
  bool act[V];
#pragma unroll
  for (int v = 0; v < V; ++v) {
    acc[v] = make_float4(0.f, 0.f, 0.f, 0.f);
    act[v] = (lane + 32 * v) * 4 < dv;
  }
  for (int i0 = 0; i0 < np; i0 += R) {
    float l[R];
    float4 x[R][V];
#pragma unroll
    for (int k = 0; k < R; ++k) {
      const bool ok = i0 + k < np;
      const int64_t row = row0 + static_cast<int64_t>(ok ? i0 + k : i0) * row_step;
      l[k] = ok ? __ldcg(ws_lse + row) : -CUDART_INF_F;
#pragma unroll
      for (int v = 0; v < V; ++v)
        x[k][v] = (ok && act[v]) ? __ldcg(reinterpret_cast<const float4*>(ws_o + row * dv + (lane + 32 * v) * 4))
                                 : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float mn = m;
#pragma unroll
    for (int k = 0; k < R; ++k) mn = fmaxf(mn, l[k]);
    if (mn == -CUDART_INF_F) continue;  // nothing but empty partials so far
    const float sc = __expf(m - mn);
    s *= sc;
#pragma unroll
    for (int v = 0; v < V; ++v) {
      acc[v].x *= sc;
      acc[v].y *= sc;
      acc[v].z *= sc;
      acc[v].w *= sc;
    }
#pragma unroll
    for (int k = 0; k < R; ++k) {
      const float w = __expf(l[k] - mn);  // 0 for an empty partial (lse = -inf)
      s += w;
#pragma unroll
      for (int v = 0; v < V; ++v) {
        acc[v].x = fmaf(w, x[k][v].x, acc[v].x);
        acc[v].y = fmaf(w, x[k][v].y, acc[v].y);
        acc[v].z = fmaf(w, x[k][v].z, acc[v].z);
        acc[v].w = fmaf(w, x[k][v].w, acc[v].w);
      }
    }
    m = mn;
  }
  const float inv = s > 0.f ? 1.f / s : 0.f;
#pragma unroll
  for (int v = 0; v < V; ++v) {
    if (act[v]) {
      const int d = (lane + 32 * v) * 4;
      out[d] = QT(acc[v].x * inv);
      out[d + 1] = QT(acc[v].y * inv);
      out[d + 2] = QT(acc[v].z * inv);
      out[d + 3] = QT(acc[v].w * inv);
    }
  }
  if (lse_out && lane == 0) *lse_out = s > 0.f ? m + logf(s) : -CUDART_INF_F;
}

// Up to four heads of one request, each with at most four partials, merged by one warp with every load
// issued up front (one L2 round trip in all).  Head j of the call is workspace row row0 + j * head_step
// (+ i * row_step for partial i) and output row out + j * out_step; dv <= 128.
template <typename QT>
__device__ __forceinline__ void merge_heads4(const float* __restrict__ ws_o, const float* __restrict__ ws_lse,
                                             int64_t row0, int64_t head_step, int nheads, int64_t row_step, int np,
                                             int dv, QT* out, int64_t out_step, float* lse_out, int64_t lse_step,
                                             int lane) {
  const int d = lane * 4;
  const bool act = d < dv;
  float l[4][4];
  float4 x[4][4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const bool ok = j < nheads && k < np;
      const int64_t row = row0 + (ok ? j * head_step + k * row_step : 0);
      l[j][k] = ok ? __ldcg(ws_lse + row) : -CUDART_INF_F;
      x[j][k] = (ok && act) ? __ldcg(reinterpret_cast<const float4*>(ws_o + row * dv + d)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (j >= nheads) break;
    const float m = fmaxf(fmaxf(l[j][0], l[j][1]), fmaxf(l[j][2], l[j][3]));
    float s = 0.f;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (m != -CUDART_INF_F) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float w = __expf(l[j][k] - m);
        s += w;
        acc.x = fmaf(w, x[j][k].x, acc.x);
        acc.y = fmaf(w, x[j][k].y, acc.y);
        acc.z = fmaf(w, x[j][k].z, acc.z);
        acc.w = fmaf(w, x[j][k].w, acc.w);
      }
    }
    const float inv = s > 0.f ? 1.f / s : 0.f;
    if (act) {
      QT* o = out + j * out_step + d;
      o[0] = QT(acc.x * inv);
      o[1] = QT(acc.y * inv);
      o[2] = QT(acc.z * inv);
      o[3] = QT(acc.w * inv);
    }
    if (lse_out && lane == 0) lse_out[j * lse_step] = s > 0.f ? m + logf(s) : -CUDART_INF_F;
  }
}

// Many partials per head (a request cut across many CTA ranges: small batches such as C1, where every request
// spans ~10 of the 148 ranges).  The lanes of a warp split into 32 / (dv / 4) groups, one head per group (two
// heads at dv = 64), and up to R = 16 partials are in flight per round, so a request of <= 16 pieces costs each
// warp hq / (8 * groups) round trips instead of 2 per head (C1: 2 instead of 8; the merge of the last request
// ran 12 us past the end of the stream).  Head j of the call: workspace row row0 + j * head_step, output
// out + j * out_step; heads j >= nheads are skipped.  dv = 64 or 128.
template <typename QT>
__device__ __forceinline__ void merge_rows_wide(const float* __restrict__ ws_o, const float* __restrict__ ws_lse,
                                                int64_t row0, int64_t head_step, int nheads, int64_t row_step, int np,
                                                int dv, QT* out, int64_t out_step, float* lse_out, int64_t lse_step,
                                                int lane) {
  constexpr int R = 16;
  const int lanes_per_head = dv >> 2;            // 16 or 32
  const int j = lane / lanes_per_head;           // this lane's head of the call
  const int d = (lane - j * lanes_per_head) * 4;
  if (j >= nheads) return;
  row0 += j * head_step;
  float m = -CUDART_INF_F, s = 0.f;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int i0 = 0; i0 < np; i0 += R) {
    float l[R];
    float4 x[R];
#pragma unroll
    for (int k = 0; k < R; ++k) {
      const bool ok = i0 + k < np;
      const int64_t row = row0 + static_cast<int64_t>(ok ? i0 + k : i0) * row_step;
      l[k] = ok ? __ldcg(ws_lse + row) : -CUDART_INF_F;
      x[k] = ok ? __ldcg(reinterpret_cast<const float4*>(ws_o + row * dv + d)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float mn = m;
#pragma unroll
    for (int k = 0; k < R; ++k) mn = fmaxf(mn, l[k]);
    if (mn == -CUDART_INF_F) continue;
    const float sc = __expf(m - mn);
    s *= sc;
    acc.x *= sc;
    acc.y *= sc;
    acc.z *= sc;
    acc.w *= sc;
#pragma unroll
    for (int k = 0; k < R; ++k) {
      const float w = __expf(l[k] - mn);
      s += w;
      acc.x = fmaf(w, x[k].x, acc.x);
      acc.y = fmaf(w, x[k].y, acc.y);
      acc.z = fmaf(w, x[k].z, acc.z);
      acc.w = fmaf(w, x[k].w, acc.w);
    }
    m = mn;
  }
  const float inv = s > 0.f ? 1.f / s : 0.f;
  QT* o = out + j * out_step + d;
  o[0] = QT(acc.x * inv);
  o[1] = QT(acc.y * inv);
  o[2] = QT(acc.z * inv);
  o[3] = QT(acc.w * inv);
  if (lse_out && d == 0) lse_out[j * lse_step] = s > 0.f ? m + logf(s) : -CUDART_INF_F;
}

// One (request, head) row merged by this warp; picks the unroll against the number of partials.
template <typename QT>
__device__ __forceinline__ void merge_head(const float* __restrict__ ws_o, const float* __restrict__ ws_lse,
                                           int64_t row0, int64_t row_step, int np, int dv, QT* out, float* lse_out,
                                           int lane) {
  if (dv > 128) {
    merge_rows<QT, 4, 2>(ws_o, ws_lse, row0, row_step, np, dv, out, lse_out, lane);
  } else if (np <= 4) {
    merge_rows<QT, 4, 1>(ws_o, ws_lse, row0, row_step, np, dv, out, lse_out, lane);
  } else {
    merge_rows<QT, 16, 1>(ws_o, ws_lse, row0, row_step, np, dv, out, lse_out, lane);   // rounds of 16, as merge_rows_wide
  }
}
#endif

}  // namespace b200attn
