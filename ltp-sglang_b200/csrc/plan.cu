// K0: plan.  One launch per forward (all layers share it): scans seq_lens into
// kv_indptr (bit-exact with torch.cumsum, triton_backend.py:172), optionally
// flattens req_to_token rows into kv_indices (bit-exact with
// create_flashinfer_kv_indices_triton, layers/attention/utils.py:10-46), and
// cuts the chunk sequence of the whole batch into equal per-CTA ranges for the
// persistent decode kernel.  This replaces the reference's per-request split
// heuristic (get_num_kv_splits_triton, triton_backend.py:875-924): any split
// plan gives the same math, so the plan is built for 148 SMs instead.
// Everything runs on the device: no host sync, CUDA-graph safe.
#include <cstdlib>

#include "common.cuh"

namespace b200attn {

struct PlanParams {
  const int32_t* req_to_token;
  int64_t r2t_stride;
  const void* req_pool_indices;
  const void* seq_lens;
  const int32_t* csr_indptr;
  int32_t* kv_indptr;
  void* plan;
  int32_t idx64;
  int32_t batch;
  int32_t ng;
  int32_t chunk_tokens;
  int32_t max_ctas;
  int32_t min_chunks;
  int32_t dyn_pct;      // share of the chunks handed out dynamically, percent
  int32_t dyn_chunks;   // chunks per dynamic range
  int32_t window;       // sliding window W (0 = none): the last min(S, W + 1) tokens
  int32_t balance;      // 1: size the static ranges by the per-CTA streaming rates decode_kernel has recorded
};

__global__ void __launch_bounds__(kPlanThreads) plan_kernel(PlanParams p) {
  __shared__ int warp_sums[32];
  __shared__ int carry_tok, carry_chunk, n_empty;
  PlanView pv = plan_view(p.plan, p.batch);
  const int tid = threadIdx.x;
  if (tid == 0) {
    carry_tok = 0;
    carry_chunk = 0;
    n_empty = 0;
    pv.chunk_prefix[0] = 0;
    if (p.kv_indptr) p.kv_indptr[0] = 0;
  }
  __syncthreads();

  const int T = p.chunk_tokens;
  for (int base = 0; base < p.batch; base += kPlanThreads) {
    const int b = base + tid;
    int s = 0;
    int64_t row = 0;
    if (b < p.batch) {
      if (p.csr_indptr) {
        const int lo = p.csr_indptr[b];
        s = p.csr_indptr[b + 1] - lo;
        row = lo;
      } else {
        int64_t sl = load_idx(p.seq_lens, b, p.idx64);
        s = sl < 0 ? 0 : static_cast<int>(sl);
        row = load_idx(p.req_pool_indices, b, p.idx64) * p.r2t_stride;
        if (p.window > 0 && s > p.window + 1) {  // triton_backend.py:936-946: lens = min(S, W + 1), start = S - lens
          row += s - (p.window + 1);
          s = p.window + 1;
        }
      }
      if (s < 0) s = 0;
    }
    const int nch = (s + T - 1) / T;
    const int tok_incl = block_scan_incl(s, warp_sums);
    const int ch_incl = block_scan_incl(nch, warp_sums);
    if (b < p.batch) {
      if (s == 0) atomicAdd(&n_empty, 1);
      pv.seq_len[b] = s;
      pv.row_off[b] = row;
      pv.chunk_prefix[b + 1] = carry_chunk + ch_incl;
      if (p.kv_indptr) p.kv_indptr[b + 1] = carry_tok + tok_incl;
    }
    __syncthreads();
    if (tid == kPlanThreads - 1) {
      carry_tok += tok_incl;
      carry_chunk += ch_incl;
    }
    __syncthreads();
  }

  // chunk_prefix[] was written by this block: make it visible to all threads.
  __threadfence_block();
  __syncthreads();

  const int total = carry_chunk * p.ng;
  __shared__ PlanHeader sh;
  if (tid == 0) {
    PlanHeader h;
    h.magic = kPlanMagic;
    h.batch = p.batch;
    h.ng = p.ng;
    h.chunk_tokens = T;
    h.total_chunks = total;
    h.grid = p.max_ctas;
    h.csr_mode = p.csr_indptr != nullptr;
    h.total_tokens = carry_tok;
    h.n_empty = n_empty;
    h.ticket = 0ull;
    const int G = p.max_ctas;
    const int minc = p.min_chunks < 1 ? 1 : p.min_chunks;
    int per_s, n_static, per_d = 1, n_dyn = 0;
    if (total <= G * minc || p.dyn_pct <= 0) {
      // small batch (or dynamic ranges disabled): equal static ranges only
      per_s = (total + G - 1) / G;
      if (per_s < minc) per_s = minc;
      n_static = (total + per_s - 1) / per_s;
    } else {
      per_s = static_cast<int>(static_cast<int64_t>(total) * (100 - p.dyn_pct) / (100ll * G));
      if (per_s < minc) per_s = minc;
      int64_t covered = static_cast<int64_t>(per_s) * G;
      if (covered >= total) {
        per_s = (total + G - 1) / G;
        n_static = (total + per_s - 1) / per_s;
      } else {
        n_static = G;
        const int rem = total - static_cast<int>(covered);
        per_d = p.dyn_chunks > minc ? p.dyn_chunks : minc;
        const int room = kMaxRanges - G;
        const int need = (rem + room - 1) / room;
        if (per_d < need) per_d = need;
        n_dyn = (rem + per_d - 1) / per_d;
      }
    }
    h.per_static = per_s;
    h.per_dynamic = per_d;
    h.n_static = n_static;
    h.n_dynamic = n_dyn;
    h.n_ranges = n_static + n_dyn;
    *pv.hdr = h;
    sh = h;
  }
  __syncthreads();
  const PlanHeader h = sh;

  // Range table.  Static range c is CTA c's: equal shares, or -- when every CTA has a share worth balancing and
  // decode_kernel has left its streaming rates in cta_stat -- shares in inverse proportion to the rate, so that
  // slow and fast SMs finish together.  Integer weights (4096 = average) keep the split exact and reproducible.
  {
    const int G = h.grid;
    const bool weighted = p.balance && h.n_static == G && h.per_static >= 16;
    int rate = 0, valid = 0;
    if (weighted && tid < G) {
      const CtaStat st = pv.cta_stat[tid];
      if (st.tag == (kStatTag ^ static_cast<uint32_t>(tid)) && st.clk_per_chunk > 16.f && st.clk_per_chunk < 1.0e6f) {
        rate = static_cast<int>(st.clk_per_chunk);
        valid = 1;
      }
    }
    const int rate_sum = block_scan_incl(rate, warp_sums);
    const int valid_cnt = block_scan_incl(valid, warp_sums);
    __shared__ int mean_rate_sh;
    if (tid == kPlanThreads - 1) mean_rate_sh = valid_cnt > 0 ? rate_sum / valid_cnt : 0;
    __syncthreads();
    int wgt = 0;
    if (tid < h.n_static) {
      wgt = 4096;
      if (valid && mean_rate_sh > 0) {
        const long long x = static_cast<long long>(mean_rate_sh) * 4096 / rate;
        wgt = x < 3482 ? 3482 : (x > 4833 ? 4833 : static_cast<int>(x));   // [0.85, 1.18]
      }
    }
    const int cum = block_scan_incl(wgt, warp_sums);
    __shared__ int wsum_sh;
    if (tid == kPlanThreads - 1) wsum_sh = cum;
    __syncthreads();
    const int static_chunks = h.n_dynamic > 0 ? h.n_static * h.per_static : h.total_chunks;
    if (tid < h.n_static) {
      int g;
      if (weighted && mean_rate_sh > 0)
        g = static_cast<int>(static_cast<long long>(static_chunks) * (cum - wgt) / wsum_sh);
      else
        g = tid * h.per_static;
      pv.range_start[tid] = g < h.total_chunks ? g : h.total_chunks;
    }
    for (int r = h.n_static + tid; r <= h.n_ranges; r += kPlanThreads) {
      const long long g = static_cast<long long>(static_chunks) + static_cast<long long>(r - h.n_static) * h.per_dynamic;
      pv.range_start[r] = g < h.total_chunks ? static_cast<int>(g) : h.total_chunks;
    }
    if (h.n_dynamic == 0 && tid == 0) pv.range_start[h.n_ranges] = h.total_chunks;
  }
  __threadfence_block();
  __syncthreads();
  auto range_begin = [&](int r) -> int { return pv.range_start[r]; };
  auto range_of = [&](int g) -> int {   // largest r with range_start[r] <= g (empty ranges share a start: take the last)
    int lo = 0, hi = h.n_ranges;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (pv.range_start[mid] <= g) lo = mid; else hi = mid;
    }
    return lo;
  };

  // First and last virtual request (b * ng + hg) of every range: the ones whose chunk interval
  // contains the range's first / last chunk.  Virtual request (b, hg) starts at ng * prefix[b] + hg * n_b.
  for (int r = tid; r < 2 * kMaxRanges; r += kPlanThreads) {
    const int rr = r >> 1, last = r & 1;
    int vq = p.batch * p.ng;
    if (rr < h.n_ranges) {
      const int64_t g0 = last ? range_begin(rr + 1) - 1 : range_begin(rr);
      int lo = 0, hi = p.batch;  // largest b with ng * prefix[b] <= g0
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (static_cast<int64_t>(pv.chunk_prefix[mid]) * p.ng <= g0) lo = mid; else hi = mid;
      }
      const int nb = pv.chunk_prefix[lo + 1] - pv.chunk_prefix[lo];
      const int hg = nb > 0 ? static_cast<int>((g0 - static_cast<int64_t>(pv.chunk_prefix[lo]) * p.ng) / nb) : 0;
      vq = lo * p.ng + hg;
    }
    (last ? pv.range_vlast : pv.range_vreq)[rr] = vq;
  }
  // Range descriptors: bounds + the first piece the decode kernel's PieceWalk visits (the range's last
  // virtual request, whose chunk interval holds the range's last chunk and is therefore never empty).
  __syncthreads();   // range_vreq / range_vlast above were written by other threads of this block
  for (int r = tid; r < h.n_ranges; r += kPlanThreads) {
    RangeDesc d;
    d.v_first = pv.range_vreq[r];
    d.v_last = pv.range_vlast[r];
    d.g0 = range_begin(r);
    d.g1 = range_begin(r + 1);
    d.b = d.v_last / p.ng;
    d.hg = d.v_last - d.b * p.ng;
    const int cp0 = pv.chunk_prefix[d.b];
    d.nchunks = pv.chunk_prefix[d.b + 1] - cp0;
    const int vstart = p.ng * cp0 + d.hg * d.nchunks;
    d.lc0 = (vstart > d.g0 ? vstart : d.g0) - vstart;
    d.lc1 = (vstart + d.nchunks < d.g1 ? vstart + d.nchunks : d.g1) - vstart;
    d.seq_len = pv.seq_len[d.b];
    d.row_off = pv.row_off[d.b];
    pv.range_desc[r] = d;
  }
  // Slot ids of the first chunk each static range streams (its first piece's chunk lc0), copied next to the
  // descriptors: the decode producers' first copies then wait on ONE plan-buffer load instead of
  // descriptor -> req_to_token row -> pool rows (a dependent DRAM round trip of ~1 us at every launch of every
  // layer: tools/trace_decode.py).  req_to_token mode only (a CSR index list is not known to the plan).
  if (!p.csr_indptr) {
    __syncthreads();   // range_desc written above
    const int ns = h.n_static < kMaxCtas ? h.n_static : kMaxCtas;
    for (int i = tid; i < ns * T; i += kPlanThreads) {
      const int r = i / T, j = i - r * T;
      const RangeDesc& d = pv.range_desc[r];
      const int t = d.lc0 * T + j;
      pv.first_ids[r * kFirstIds + j] = t < d.seq_len ? p.req_to_token[d.row_off + t] : 0;
    }
  }
  // Per virtual request: the ranges that touch it (= its partial pieces) and the arrival counter.
  for (int v = tid; v < p.batch * p.ng; v += kPlanThreads) {
    const int b = v / p.ng, hg = v - b * p.ng;
    const int cp0 = pv.chunk_prefix[b];
    const int n = pv.chunk_prefix[b + 1] - cp0;
    int rf = 0, np = 0;
    if (n > 0) {
      const int vstart = p.ng * cp0 + hg * n;
      rf = range_of(vstart);
      np = range_of(vstart + n - 1) - rf + 1;
    }
    pv.vreq_rfirst[v] = rf;
    pv.vreq_pieces[v] = np;
    pv.vreq_arrived[v] = 0;
  }
}

// kv_indices[kv_indptr[b] + j] = req_to_token[req_pool_indices[b], j]
__global__ void __launch_bounds__(256) kv_indices_kernel(const int32_t* __restrict__ r2t,
                                                        const void* plan_base, int batch,
                                                        const int32_t* __restrict__ kv_indptr,
                                                        int32_t* __restrict__ out) {
  PlanView pv = plan_view(const_cast<void*>(plan_base), batch);
  for (int b = blockIdx.x; b < batch; b += gridDim.x) {
    const int s = pv.seq_len[b];
    const int32_t* row = r2t + pv.row_off[b];
    int32_t* dst = out + kv_indptr[b];
    for (int j = threadIdx.x; j < s; j += blockDim.x) dst[j] = row[j];
  }
}

}  // namespace b200attn

using namespace b200attn;

// Tuning aids (read once): B200ATTN_DYN_PCT (1..90; default 0 = equal static ranges only: measured on
// B200 the per-SM speed spread is smaller than one dynamic range, so the dynamic tail only adds
// pieces -- profiles/r01_dynamic_ranges.txt), B200ATTN_DYN_BYTES.
static int plan_default_dyn_pct() {
  static const int v = [] { const char* e = getenv("B200ATTN_DYN_PCT"); return e ? atoi(e) : 0; }();
  return v;
}
static int plan_default_dyn_bytes() {
  static const int v = [] { const char* e = getenv("B200ATTN_DYN_BYTES"); return e ? atoi(e) : 512 * 1024; }();
  return v;
}

extern "C" __attribute__((visibility("default"))) size_t b200attn_plan_bytes(int32_t batch) {
  if (batch < 0) return 0;
  return plan_bytes_for(batch);
}

extern "C" __attribute__((visibility("default"))) int b200attn_plan(const b200attn_plan_args* a, b200attn_stream_t stream_) {
  if (!a || !a->plan || a->batch < 0) return B200ATTN_E_INVALID;
  if (a->batch == 0) return B200ATTN_OK;
  const bool csr = a->csr_indptr != nullptr;
  if (!csr && (!a->req_to_token || !a->req_pool_indices || !a->seq_lens)) return B200ATTN_E_INVALID;
  if (a->index_dtype != B200ATTN_I32 && a->index_dtype != B200ATTN_I64) return B200ATTN_E_INVALID;
  if (a->kv_indices && (csr || !a->kv_indptr)) return B200ATTN_E_INVALID;
  if (a->sliding_window < 0 || (csr && a->sliding_window > 0)) return B200ATTN_E_INVALID;
  if (reinterpret_cast<uintptr_t>(a->plan) % 16 != 0) return B200ATTN_E_ALIGN;
  if (a->plan_bytes < plan_bytes_for(a->batch)) return B200ATTN_E_WORKSPACE;
  Geometry g;
  if (!make_geometry(a->num_q_heads, a->num_kv_heads, a->head_dim, a->kv_dtype, &g, a->batch))
    return B200ATTN_E_UNSUPPORTED;

  PlanParams p;
  p.req_to_token = a->req_to_token;
  p.r2t_stride = a->req_to_token_stride;
  p.req_pool_indices = a->req_pool_indices;
  p.seq_lens = a->seq_lens;
  p.csr_indptr = a->csr_indptr;
  p.kv_indptr = a->kv_indptr;
  p.plan = a->plan;
  p.idx64 = a->index_dtype == B200ATTN_I64;
  p.batch = a->batch;
  p.ng = g.ng;
  p.chunk_tokens = g.chunk_tokens;
  int n = sm_count();
  if (a->max_ctas > 0 && a->max_ctas < n) n = a->max_ctas;
  if (n > kMaxRanges / 2) n = kMaxRanges / 2;
  p.max_ctas = n;
  p.window = a->sliding_window;
  p.balance = no_balance() ? 0 : 1;
  // Default floor: >= 64 tokens per piece so the partial write-out stays small
  // next to the KV bytes a piece streams.
  p.min_chunks = a->min_chunks_per_cta > 0 ? a->min_chunks_per_cta : (64 + g.chunk_tokens - 1) / g.chunk_tokens;
  // Optional dynamic tail (off by default): a share of the chunks in ranges of ~512 KB of KV.
  p.dyn_pct = a->dynamic_percent < 0 ? 0 : (a->dynamic_percent == 0 ? plan_default_dyn_pct() : a->dynamic_percent);
  if (p.dyn_pct > 90) p.dyn_pct = 90;
  const int chunk_bytes = 2 * g.chunk_tokens * g.row_bytes;
  p.dyn_chunks = a->dynamic_range_chunks > 0 ? a->dynamic_range_chunks : (plan_default_dyn_bytes() + chunk_bytes - 1) / chunk_bytes;

  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  plan_kernel<<<1, kPlanThreads, 0, stream>>>(p);
  count_launch();
  if (cudaGetLastError() != cudaSuccess) return B200ATTN_E_CUDA;
  if (a->kv_indices) {
    int blocks = a->batch < sm_count() * 8 ? a->batch : sm_count() * 8;
    kv_indices_kernel<<<blocks, 256, 0, stream>>>(a->req_to_token, a->plan, a->batch, a->kv_indptr,
                                                  a->kv_indices);
    count_launch();
    if (cudaGetLastError() != cudaSuccess) return B200ATTN_E_CUDA;
  }
  return B200ATTN_OK;
}

// ---------------------------------------------------------------------------
// Multi-step draft decode (speculative decoding): kv_indptr / kv_indices of all
// `num_steps` draft steps in one launch.  Replaces generate_draft_decode_kv_indices
// (python/sglang/srt/speculative/eagle_utils.py:844-920) as TritonMultiStepDraftBackend
// launches it (layers/attention/triton_backend.py:779-809), same outputs bit for bit:
//   step it (iters = it + 1), sequence s, branch t:
//     kv_indices[it][cum(s) * topk + s * iters * topk + t * (S_s + iters) + j]
//         = req_to_token[req_s, j]                            j < S_s
//         = req_to_token[req_s, ext(s, t) + (j - S_s)]        S_s <= j < S_s + iters
//     kv_indptr[it][z] = sum(positions[0 .. z)) + z * iters    z = s * topk + t, z = 0 writes entry
//                                                              num_seqs * topk instead (the total)
//   ext = S_s + t * num_steps (page_size 1 or topk 1), else the branch's own pages:
//   S_s / P * P + t * ceil((S_s % P + num_steps) / P) * P + S_s % P.
// ---------------------------------------------------------------------------
namespace b200attn {

struct DraftIdxParams {
  const void* req_pool_indices;
  const int32_t* req_to_token;
  int64_t r2t_stride;
  const void* seq_lens;
  const void* positions;
  int32_t* kv_indices;
  int64_t kv_indices_stride;
  int32_t* kv_indptr;
  int64_t kv_indptr_stride;
  int32_t idx64, pos64;
  int32_t num_steps, num_seqs, topk, page_size;
};

__global__ void __launch_bounds__(128) draft_kv_indices_kernel(const DraftIdxParams p) {
  __shared__ long long red[4];
  const int it = blockIdx.x, s = blockIdx.y, t = blockIdx.z;
  const int iters = it + 1;
  const int tid = threadIdx.x;
  auto block_sum = [&](long long v) -> long long {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    return red[0] + red[1] + red[2] + red[3];
  };
  long long part = 0;
  for (int i = tid; i < s; i += 128) part += load_idx(p.seq_lens, i, p.idx64);
  const long long cum = block_sum(part);
  const long long S = load_idx(p.seq_lens, s, p.idx64);
  const int32_t* row = p.req_to_token + load_idx(p.req_pool_indices, s, p.idx64) * p.r2t_stride;
  int32_t* dst = p.kv_indices + p.kv_indices_stride * it + cum * p.topk +
                 static_cast<long long>(s) * iters * p.topk + static_cast<long long>(t) * (S + iters);
  for (long long j = tid; j < S; j += 128) dst[j] = row[j];
  long long ext;
  if (p.page_size == 1 || p.topk == 1) {
    ext = S + static_cast<long long>(t) * p.num_steps;
  } else {
    const long long last = S % p.page_size;
    const long long pages = (last + p.num_steps + p.page_size - 1) / p.page_size;
    ext = S / p.page_size * p.page_size + t * pages * p.page_size + last;
  }
  for (int e = tid; e < iters; e += 128) dst[S + e] = row[ext + e];

  int z = s * p.topk + t;
  if (z == 0) z = p.num_seqs * p.topk;
  part = 0;
  for (int i = tid; i < z; i += 128) part += load_idx(p.positions, i, p.pos64);
  const long long base = block_sum(part);
  if (tid == 0) p.kv_indptr[p.kv_indptr_stride * it + z] = static_cast<int32_t>(base + static_cast<long long>(z) * iters);
}

}  // namespace b200attn

extern "C" __attribute__((visibility("default"))) int b200attn_draft_kv_indices(const b200attn_draft_kv_indices_args* a,
                                                                                b200attn_stream_t stream_) {
  if (!a || a->num_steps < 0 || a->num_seqs < 0 || a->topk < 0) return B200ATTN_E_INVALID;
  if (a->num_steps == 0 || a->num_seqs == 0 || a->topk == 0) return B200ATTN_OK;
  if (!a->req_pool_indices || !a->req_to_token || !a->seq_lens || !a->positions || !a->kv_indices || !a->kv_indptr)
    return B200ATTN_E_INVALID;
  if ((a->index_dtype != B200ATTN_I32 && a->index_dtype != B200ATTN_I64) ||
      (a->positions_dtype != B200ATTN_I32 && a->positions_dtype != B200ATTN_I64) || a->page_size < 1)
    return B200ATTN_E_INVALID;
  if (a->num_seqs > 65535 || a->topk > 65535) return B200ATTN_E_UNSUPPORTED;
  DraftIdxParams p;
  p.req_pool_indices = a->req_pool_indices;
  p.req_to_token = a->req_to_token;
  p.r2t_stride = a->req_to_token_stride;
  p.seq_lens = a->seq_lens;
  p.positions = a->positions;
  p.kv_indices = a->kv_indices;
  p.kv_indices_stride = a->kv_indices_stride;
  p.kv_indptr = a->kv_indptr;
  p.kv_indptr_stride = a->kv_indptr_stride;
  p.idx64 = a->index_dtype == B200ATTN_I64;
  p.pos64 = a->positions_dtype == B200ATTN_I64;
  p.num_steps = a->num_steps;
  p.num_seqs = a->num_seqs;
  p.topk = a->topk;
  p.page_size = a->page_size;
  draft_kv_indices_kernel<<<dim3(a->num_steps, a->num_seqs, a->topk), 128, 0, static_cast<cudaStream_t>(stream_)>>>(p);
  count_launch();
  return cudaGetLastError() == cudaSuccess ? B200ATTN_OK : B200ATTN_E_CUDA;
}
