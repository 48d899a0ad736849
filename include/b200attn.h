/*
 * b200attn.h — C-ABI of the B200-native paged decode-attention path.
 *
 * This is the drop-in boundary: a shared library with plain
 * pointers and sizes, no torch types.  Every entry point enqueues work on the
 * given CUDA stream and returns immediately: no host synchronisation, no
 * allocation, CUDA-graph capture safe.  All memory is owned by the caller.
 *
 * Reference interfaces each entry replaces (paths relative to the reference
 * tree, microsoft/ltp-sglang):
 *
 *   b200attn_kv_store   python/sglang/srt/mem_cache/memory_pool.py:369-407
 *                       (MHATokenToKVPool.set_kv_buffer: optional div_ by
 *                       k/v scale, cast to the pool dtype, index_put by loc)
 *   b200attn_plan       python/sglang/srt/layers/attention/triton_backend.py:160-221
 *                       (cumsum -> kv_indptr, create_flashinfer_kv_indices_triton
 *                       layers/attention/utils.py:10-46, get_num_kv_splits :124-158)
 *   b200attn_decode     python/sglang/srt/layers/attention/triton_ops/decode_attention.py:677-728
 *                       (decode_attention_fwd: split-KV stage 1 :240-488 + stage 2
 *                       LSE merge :491-596), and the math of
 *                       torch_native_backend.py:112-180
 *   b200attn_rope_kv_store  python/sglang/srt/layers/rotary_embedding.py:218-236 +
 *                       sgl-kernel/csrc/elementwise/rope.cu:22-89 (RoPE in place) fused with
 *                       set_kv_buffer (memory_pool.py:369-407); call site models/llama.py:186-189
 *   b200attn_merge_states  sgl-kernel/csrc/attention/merge_attn_states.cu:31-204
 *                       (merge_state_v2: 2-way (o, lse) merge)
 *   b200attn_extend     python/sglang/srt/layers/attention/triton_ops/extend_attention.py:306-438
 *                       (extend_attention_fwd: prefill / extend over cached prefix + new tokens,
 *                       causal, custom-mask (speculative verify), sliding-window and logit-cap
 *                       variants), and the math of torch_native_backend.py:27-110
 *
 * Error convention: 0 on success, a negative B200ATTN_E_* code otherwise.
 * Nothing is launched when an error is returned.  Launch failures reported by
 * cudaGetLastError() right after a launch map to B200ATTN_E_CUDA.
 */
#ifndef B200ATTN_H_
#define B200ATTN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200ATTN_ABI_VERSION 6

/* cudaStream_t without pulling in the CUDA headers. */
typedef void* b200attn_stream_t;

enum {
  B200ATTN_OK = 0,
  B200ATTN_E_INVALID = -1,     /* null pointer, negative size, bad enum */
  B200ATTN_E_UNSUPPORTED = -2, /* shape / dtype combination has no kernel */
  B200ATTN_E_ALIGN = -3,       /* pointer or stride breaks an alignment rule */
  B200ATTN_E_CUDA = -4,        /* the CUDA runtime reported a launch error */
  B200ATTN_E_WORKSPACE = -5,   /* plan / workspace buffer too small */
  B200ATTN_E_STATIC_LIMIT = -6 /* b200attn_decode_static_ws: the call needs more than B200ATTN_STATIC_WS_BYTES */
};

/* Element types. */
enum {
  B200ATTN_BF16 = 0,
  B200ATTN_F16 = 1,
  B200ATTN_FP8_E4M3 = 2, /* torch.float8_e4m3fn, stored as raw bytes (uint8 pool) */
  B200ATTN_F32 = 3,      /* b200attn_merge_states only */
  B200ATTN_FP8_E5M2 = 4  /* ABI 6: torch.float8_e5m2 (--kv-cache-dtype fp8_e5m2, server_args.py:828-833), raw bytes */
};

/* Width of the seq_lens / req_pool_indices arrays (int64 eager, int32 under
 * CUDA graphs: cuda_graph_runner.py:302-305). */
enum { B200ATTN_I32 = 0, B200ATTN_I64 = 1 };

/* How kv_store divides by the per-tensor scale before the fp8 cast. */
enum {
  B200ATTN_SCALE_NONE = 0,      /* set_kv_buffer called without k_scale/v_scale */
  B200ATTN_SCALE_TRUE_DIV = 1,  /* x / s in fp32, rounded to the source dtype (torch CPU div_;
                                   CUDA div_ by a 0-dim CUDA tensor, where torch first casts
                                   the divisor to the source dtype: a *_scale_ptr scale is
                                   therefore rounded to src_dtype before use) */
  B200ATTN_SCALE_RECIP_MUL = 2  /* x * (1.0f / s): what torch's CUDA div_ computes
                                   when the divisor is a Python float */
};

/* ------------------------------------------------------------------ */
/* K1: KV store                                                        */
/* ------------------------------------------------------------------ */
typedef struct {
  /* new K / V of this step: [n_tokens, num_kv_heads, head_dim], last dim
   * contiguous, strides in ELEMENTS (views into the fused qkv buffer). */
  void* k;
  void* v;
  int64_t k_stride_token, k_stride_head;
  int64_t v_stride_token, v_stride_head;
  const int64_t* loc;          /* [n_tokens] pool slot per token (out_cache_loc) */
  /* pool of one layer: [slots, num_kv_heads, head_dim], contiguous rows. */
  void* k_pool;
  void* v_pool;
  int64_t k_pool_stride_slot;  /* elements (of the pool dtype) between slots */
  int64_t v_pool_stride_slot;
  int64_t pool_slots;          /* rows in the pool; a loc outside [0, slots) is skipped */
  int32_t n_tokens;
  int32_t num_kv_heads;
  int32_t head_dim;            /* K head dim */
  int32_t v_head_dim;          /* V head dim */
  int32_t src_dtype;           /* B200ATTN_BF16 | B200ATTN_F16 */
  int32_t pool_dtype;          /* same as src, or B200ATTN_FP8_E4M3 / B200ATTN_FP8_E5M2 */
  int32_t scale_mode;          /* B200ATTN_SCALE_*; ignored unless pool is fp8 */
  int32_t mutate_src;          /* 1: write x/scale back into k/v (div_ is in place) */
  float k_scale, v_scale;      /* host scales (used when *_scale_ptr is NULL) */
  const float* k_scale_ptr;    /* optional device 0-dim fp32 scale (layer.k_scale) */
  const float* v_scale_ptr;
} b200attn_kv_store_args;

int b200attn_kv_store(const b200attn_kv_store_args* args, b200attn_stream_t stream);

/* ------------------------------------------------------------------ */
/* K4: fused RoPE + KV store (producer side of K1)                     */
/* ------------------------------------------------------------------ */
/* Replaces, in one launch, the two steps LlamaAttention.forward runs between the qkv
 * projection and attention (python/sglang/srt/models/llama.py:186-189):
 *   q, k = rotary_emb(positions, q, k)      layers/rotary_embedding.py:218-236 (forward_cuda)
 *                                           -> sgl-kernel/csrc/elementwise/rope.cu:22-89
 *                                           (apply_rope_pos_ids_cos_sin_cache, in place)
 *   set_kv_buffer(layer, loc, k, v, ...)    mem_cache/memory_pool.py:369-407
 * q and k are rotated IN PLACE (fp32 math on the fp32 cos|sin cache, one rounding to the
 * source dtype: the arithmetic of the reference's CUDA path); the rotated k and v are then
 * stored exactly as b200attn_kv_store would store them. */
typedef struct {
  void* q;                      /* [n_tokens, num_q_heads, head_dim], strides in elements */
  int64_t q_stride_token, q_stride_head;
  int32_t num_q_heads;
  int32_t rotary_dim;           /* leading dims of each head that rotate; even, <= head_dim,
                                   multiple of 16 for neox / of 8 for interleaved layout */
  int32_t is_neox;              /* 1: pairs (i, i + rotary_dim/2); 0: pairs (2i, 2i + 1) */
  int32_t pos_dtype;            /* B200ATTN_I32 | B200ATTN_I64 */
  const void* positions;        /* [n_tokens] */
  const float* cos_sin_cache;   /* [max_position, rotary_dim] fp32: cos | sin halves */
  int64_t cache_stride;         /* elements between cache rows */
  int64_t max_position;         /* rows in the cache; positions are clamped into range */
  int32_t save_kv;              /* 1: scatter rotated k and v into the pool; 0: RoPE only */
  int32_t pad_;
  b200attn_kv_store_args kv;    /* k (rotated in place), v, loc, pools, scales: as for
                                   b200attn_kv_store; head_dim must equal v_head_dim here */
} b200attn_rope_kv_store_args;

int b200attn_rope_kv_store(const b200attn_rope_kv_store_args* args, b200attn_stream_t stream);

/* ------------------------------------------------------------------ */
/* K0: plan (once per forward; shared by all layers)                   */
/* ------------------------------------------------------------------ */
typedef struct {
  /* request -> slot map: int32 [max_reqs, stride] (ReqToTokenPool.req_to_token) */
  const int32_t* req_to_token;
  int64_t req_to_token_stride; /* elements between rows */
  const void* req_pool_indices; /* [batch] int32|int64 */
  const void* seq_lens;         /* [batch] int32|int64, includes the token being decoded */
  int32_t index_dtype;          /* B200ATTN_I32 | B200ATTN_I64 (both arrays) */
  int32_t batch;
  /* shape key of the decode calls that will use this plan */
  int32_t num_q_heads;
  int32_t num_kv_heads;
  int32_t head_dim;
  int32_t kv_dtype;             /* pool dtype */
  /* CSR input mode (speculative decode hands kv_indptr/kv_indices in):
   * when csr_indptr != NULL, req_to_token/req_pool_indices/seq_lens are ignored,
   * request b reads csr_indices[csr_indptr[b] .. csr_indptr[b+1]). */
  const int32_t* csr_indptr;
  /* outputs */
  int32_t* kv_indptr;           /* optional [batch+1] int32 (bit-exact with torch.cumsum) */
  int32_t* kv_indices;          /* optional [sum(seq_lens)] int32 (bit-exact with
                                   create_flashinfer_kv_indices_triton) */
  void* plan;                   /* opaque, b200attn_plan_bytes(batch) bytes, 16B aligned */
  size_t plan_bytes;
  int32_t max_ctas;             /* 0 = one per SM; tests use small values to force splits */
  int32_t min_chunks_per_cta;   /* 0 = library default */
  int32_t dynamic_percent;      /* share of the KV chunks CTAs claim dynamically (load balance):
                                   0 = library default (none: equal static ranges), -1 = none */
  int32_t dynamic_range_chunks; /* chunks per dynamic range; 0 = library default (~512 KB of KV) */
  int32_t sliding_window;       /* > 0: request b attends to its last min(S_b, W + 1) tokens only, i.e.
                                   req_to_token[req, S_b - len : S_b] (update_sliding_window_buffer,
                                   triton_backend.py:927-956); kv_indptr / kv_indices then are the
                                   reference's window_kv_indptr / window_kv_indices.  0 = full context.
                                   Not combinable with the CSR input mode. */
  int32_t pad_;
} b200attn_plan_args;

size_t b200attn_plan_bytes(int32_t batch);
int b200attn_plan(const b200attn_plan_args* args, b200attn_stream_t stream);

/* ------------------------------------------------------------------ */
/* K2 + K3: split-KV decode and LSE merge (once per layer)             */
/* ------------------------------------------------------------------ */
typedef struct {
  const void* q;               /* [batch, num_q_heads, head_dim], strides in elements */
  int64_t q_stride_batch, q_stride_head;
  const void* k_pool;          /* [slots, num_kv_heads, head_dim] */
  const void* v_pool;
  int64_t k_pool_stride_slot;  /* elements */
  int64_t v_pool_stride_slot;
  void* o;                     /* [batch, num_q_heads, v_head_dim], q dtype */
  int64_t o_stride_batch, o_stride_head;
  float* lse;                  /* optional [batch, num_q_heads]: ln sum exp(scaled logits) */
  const void* plan;            /* written by b200attn_plan on the same stream */
  const int32_t* indices;      /* req_to_token base, or csr kv_indices in CSR mode */
  float* ws_o;                 /* workspace, b200attn_workspace_bytes() */
  size_t ws_bytes;
  int32_t batch;
  int32_t num_q_heads;
  int32_t num_kv_heads;
  int32_t head_dim;
  int32_t v_head_dim;
  int32_t q_dtype;             /* B200ATTN_BF16 | B200ATTN_F16 */
  int32_t kv_dtype;            /* q_dtype, B200ATTN_FP8_E4M3 or B200ATTN_FP8_E5M2 */
  float sm_scale;              /* layer.scaling */
  float logit_cap;             /* 0 disables; cap * tanh(x / cap) otherwise */
  float k_scale;               /* fp8 dequant: K = fp8 * k_scale (1.0 for 16-bit pools) */
  float v_scale;
  int32_t phases;              /* 0 = one launch: the split-KV kernel also merges the partials (the CTA
                                  that lands a request's last partial merges it); 1 = split-KV kernel
                                  without the merge (K2 alone); 2 = stand-alone merge kernel (K3).
                                  1 followed by 2 equals 0; they exist so each stage can be timed and
                                  tested alone. */
  const int64_t* slot_map;     /* ABI 6, optional (NULL = none): every slot id read from `indices` goes through
                                  slot_map[id] before it addresses the pool.  For the hybrid SWAKVPool
                                  (mem_cache/memory_pool.py:420-553): sliding-window layers live in a smaller pool
                                  whose slots are full_to_swa_index_mapping[full slot]
                                  (translate_loc_from_full_to_swa, :515-517; flashinfer_backend.py:754, 1001
                                  translate a materialised kv_indices list -- here the list is never built). */
  int64_t pool_slots;          /* ABI 6, optional (0 = unknown): rows of k_pool / v_pool.  Only read by the debug build
                                  (make -C csrc debug, -DB200ATTN_DEBUG), whose producers trap on a slot id outside
                                  [0, pool_slots) instead of gathering out of bounds (compute-sanitizer is not
                                  available on every pool). */
} b200attn_decode_args;

/* Bytes of split-KV workspace (partial o + lse) for up to `batch` requests. */
size_t b200attn_workspace_bytes(int32_t batch, int32_t num_q_heads, int32_t num_kv_heads,
                                int32_t head_dim, int32_t v_head_dim, int32_t kv_dtype);
int b200attn_decode(const b200attn_decode_args* args, b200attn_stream_t stream);

/* b200attn_decode for callers that may not allocate device memory (code under CUDA-graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, results and error codes as b200attn_decode, except that the split-KV workspace is a
 * device array inside the library: args->ws_o must be NULL and args->ws_bytes 0 (B200ATTN_E_INVALID otherwise).
 *   - Limit: b200attn_workspace_bytes(batch, heads..., kv_dtype) <= B200ATTN_STATIC_WS_BYTES (64 MiB; e.g. 64
 *     requests x 32 q heads x head_dim 128 need 18.1 MiB, 128 x 64 x 128 need 36.3 MiB); a larger call returns B200ATTN_E_STATIC_LIMIT and
 *     launches nothing.  Every GPU that loads the library holds the 64 MiB until the process exits.
 *   - The call and its set-up (one cudaGetSymbolAddress per device, on first use, with that device current) call
 *     no cudaMalloc / cudaFree / allocating routine.  The plan, q, pools and outputs stay the caller's.
 *   - All calls on a device share the one array: calls that overlap on the GPU (two streams) race.  Nothing in it
 *     must start at zero (every partial row is written before it is read; the counters that order the reads are
 *     in the plan buffer), so it is not cleared between calls and keeps the last call's partials. */
#define B200ATTN_STATIC_WS_BYTES (64u << 20)
int b200attn_decode_static_ws(const b200attn_decode_args* args, b200attn_stream_t stream);

/* ------------------------------------------------------------------ */
/* K5: extend (prefill with cached prefix) attention                   */
/* ------------------------------------------------------------------ */
/* Replaces extend_attention_fwd (triton_ops/extend_attention.py:306-438; kernel :41-303) as
 * TritonAttnBackend.forward_extend calls it (triton_backend.py:632-685) for EXTEND / MIXED /
 * TARGET_VERIFY / DRAFT_EXTEND batches.  Request b owns the query tokens
 * [qo_indptr[b], qo_indptr[b+1]) ("extend" tokens) and attends to
 *   - its cached prefix: Lp = kv_indptr[b+1] - kv_indptr[b] pool slots, listed either in
 *     kv_indices[kv_indptr[b] ...] (CSR, what the reference materialises) or read in place from
 *     req_to_token[req_pool_indices[b], prefix_lens[b] - Lp ...] (no kv_indices needed; with a
 *     sliding window Lp < prefix_lens[b] and the list is the prefix's tail, the reference's
 *     window_kv_indices, triton_backend.py:927-956);
 *   - the extend tokens themselves, read back from the pool at extend_loc[token] (out_cache_loc:
 *     set_kv_buffer / b200attn_kv_store has stored them before this call, :647-650).  The
 *     reference reads them from the contiguous k_extend / v_extend instead; for a 16-bit pool the
 *     bytes are identical.
 * Masks, exactly as the reference kernel applies them (q index i and kv index j are local to the
 * request's extend tokens / to its prefix list):
 *   prefix part:  [custom_mask[i, j] unless skip_prefix_custom_mask]  and
 *                 [i <= j + sliding_window when sliding_window > 0]          (:166-184)
 *   extend part:  custom_mask[i, Lp + j] when a custom mask is given, else j <= i when causal,
 *                 else everything                                            (:245-264)
 * custom_mask is the flattened per-request [Le, Lp + Le] uint8/bool array starting at
 * mask_indptr[b] (:93-95, :169-177).  A row with no visible key yields o = 0, lse = -inf. */
typedef struct {
  const void* q;               /* [tokens, num_q_heads, head_dim], strides in elements */
  int64_t q_stride_token, q_stride_head;
  void* o;                     /* [tokens, num_q_heads, v_head_dim], q dtype */
  int64_t o_stride_token, o_stride_head;
  float* lse;                  /* optional [tokens, num_q_heads]: ln sum exp(scaled logits) */
  const void* k_pool;          /* [slots, num_kv_heads, head_dim] */
  const void* v_pool;
  int64_t k_pool_stride_slot;  /* elements */
  int64_t v_pool_stride_slot;
  const int32_t* qo_indptr;    /* [batch + 1] */
  const int32_t* kv_indptr;    /* [batch + 1] prefix list lengths (CSR) */
  const int32_t* kv_indices;   /* prefix slots (CSR mode), or NULL for the req_to_token mode */
  const int32_t* req_to_token; /* req_to_token mode: [max_reqs, stride] */
  int64_t req_to_token_stride;
  const void* req_pool_indices; /* [batch] int32|int64 */
  const void* prefix_lens;      /* [batch] int32|int64: cached tokens of the request (extend_prefix_lens) */
  int32_t index_dtype;          /* B200ATTN_I32 | B200ATTN_I64 (req_pool_indices, prefix_lens) */
  int32_t mask_indptr_dtype;    /* B200ATTN_I32 | B200ATTN_I64 (the reference keeps it in int64) */
  const int64_t* extend_loc;    /* [tokens] pool slot of every extend token (out_cache_loc) */
  const uint8_t* custom_mask;   /* optional */
  const void* mask_indptr;      /* [batch + 1], required with custom_mask */
  int32_t batch;
  int32_t total_tokens;         /* rows of q */
  int32_t max_extend_len;       /* >= max_b (qo_indptr[b+1] - qo_indptr[b]); sizes the grid */
  int32_t num_q_heads;
  int32_t num_kv_heads;
  int32_t head_dim;
  int32_t v_head_dim;
  int32_t q_dtype;              /* B200ATTN_BF16 | B200ATTN_F16 */
  int32_t kv_dtype;             /* == q_dtype, or B200ATTN_FP8_E4M3 / _E5M2 (raw bytes; strides then count bytes) */
  int32_t causal;               /* 1: extend token i sees extend tokens j <= i */
  int32_t sliding_window;       /* <= 0: off */
  int32_t skip_prefix_custom_mask; /* 1 (the reference's default): custom_mask covers the extend part only */
  float sm_scale;
  float logit_cap;              /* 0 disables */
  float k_scale, v_scale;       /* fp8 pools: K = k_scale * stored, V = v_scale * stored; ignored otherwise */
  int32_t avg_kv_len_hint;      /* ABI 4.  Mean prefix + extend length per request if the host knows it (sglang:
                                   seq_lens_sum / batch_size, forward_batch_info.py:163-311), 0 = unknown.  Only
                                   picks between the two small-q kernels (same results); never read for sizes */
} b200attn_extend_args;

int b200attn_extend(const b200attn_extend_args* args, b200attn_stream_t stream);

/* Index arrays of one extend forward, in one launch (the reference runs 2-3 torch.cumsum launches
 * plus slicing, triton_backend.py:227-318):
 *   qo_indptr[b+1]   = sum_{i<=b} Le_i      Le_i = extend_seq_lens[i], or uniform_extend_len when that is
 *                                           > 0 (TARGET_VERIFY: num_draft_tokens per request, :230-236)
 *   kv_indptr[b+1]   = sum_{i<=b} Lp_i      Lp_i = prefix_lens[i], or min(prefix_lens[i], W + 1) with a
 *                                           sliding window (window_kv_indptr, :927-946)
 *   mask_indptr[b+1] = sum_{i<=b} Le_i * (prefix_lens[i] + Le_i)   (optional, int64; :252-259)
 * No host synchronisation; CUDA-graph capture safe. */
typedef struct {
  const void* extend_seq_lens;  /* [batch] int32|int64, or NULL with uniform_extend_len > 0 */
  const void* prefix_lens;      /* [batch] int32|int64 */
  int32_t index_dtype;          /* B200ATTN_I32 | B200ATTN_I64 */
  int32_t batch;
  int32_t uniform_extend_len;
  int32_t sliding_window;       /* <= 0: off */
  int32_t* qo_indptr;           /* out [batch + 1] */
  int32_t* kv_indptr;           /* out [batch + 1] */
  int64_t* mask_indptr;         /* optional out [batch + 1] */
} b200attn_extend_plan_args;

int b200attn_extend_plan(const b200attn_extend_plan_args* args, b200attn_stream_t stream);

/* ------------------------------------------------------------------ */
/* 2-way state merge (merge_state_v2 semantics)                        */
/* ------------------------------------------------------------------ */
typedef struct {
  const void* o_a;             /* [tokens, heads, head_dim] contiguous */
  const float* lse_a;          /* [tokens, heads] */
  const void* o_b;
  const float* lse_b;
  void* o_out;                 /* [tokens, heads, head_dim] */
  float* lse_out;              /* optional */
  int32_t tokens, heads, head_dim;
  int32_t dtype;               /* B200ATTN_BF16 | B200ATTN_F16 | B200ATTN_F32 (the reference dispatches all
                                  three, merge_attn_states.cu:111-122); head_dim % (16 / sizeof) == 0 */
} b200attn_merge_states_args;

int b200attn_merge_states(const b200attn_merge_states_args* args, b200attn_stream_t stream);

/* ------------------------------------------------------------------ */
/* Multi-step draft decode: kv_indptr / kv_indices of every draft step  */
/* ------------------------------------------------------------------ */
/* Replaces the Triton kernel generate_draft_decode_kv_indices
 * (python/sglang/srt/speculative/eagle_utils.py:844-920) that TritonMultiStepDraftBackend launches once per
 * draft forward with grid (num_steps, num_seqs, topk) (layers/attention/triton_backend.py:779-809).  Step `it`
 * (iters = it + 1), sequence s, branch t own the CSR row z = s * topk + t of step it:
 *   kv_indices[it][sum_{i<s} S_i * topk + s * iters * topk + t * (S_s + iters) + j]
 *       = req_to_token[req_s, j]                       for j < S_s            (the committed prefix)
 *       = req_to_token[req_s, ext + j - S_s]           for S_s <= j < S_s + iters (the branch's draft tokens)
 *     ext = S_s + t * num_steps when page_size == 1 or topk == 1, else
 *           S_s / P * P + t * ceil((S_s % P + num_steps) / P) * P + S_s % P    (each branch owns whole pages)
 *   kv_indptr[it][z] = sum(positions[0 .. z)) + z * iters; row 0 writes entry num_seqs * topk (the total), so
 *                      entry 0 is never written: the caller zero-initialises kv_indptr once.
 * positions[z] is the position of row z's newest token (= S_s for every branch of sequence s). */
typedef struct {
  const void* req_pool_indices;  /* [num_seqs] int32|int64 */
  const int32_t* req_to_token;   /* [max_reqs, stride] */
  int64_t req_to_token_stride;
  const void* seq_lens;          /* [num_seqs] int32|int64 (paged_kernel_lens) */
  int32_t index_dtype;           /* B200ATTN_I32 | B200ATTN_I64 (req_pool_indices, seq_lens) */
  int32_t positions_dtype;       /* B200ATTN_I32 | B200ATTN_I64 */
  const void* positions;         /* [num_seqs * topk] */
  int32_t* kv_indices;           /* out [num_steps, kv_indices_stride] */
  int64_t kv_indices_stride;     /* elements between steps */
  int32_t* kv_indptr;            /* out [num_steps, kv_indptr_stride] */
  int64_t kv_indptr_stride;
  int32_t num_steps, num_seqs, topk, page_size;
} b200attn_draft_kv_indices_args;

int b200attn_draft_kv_indices(const b200attn_draft_kv_indices_args* args, b200attn_stream_t stream);

/* ------------------------------------------------------------------ */
/* misc                                                                */
/* ------------------------------------------------------------------ */
int b200attn_abi_version(void);
const char* b200attn_strerror(int code);
/* Kernels launched by this library in this process (all devices); bench.py
 * reports the delta over the timed region as "gpu_launches". */
uint64_t b200attn_launch_count(void);
/* Decode geometry the library picks for a shape (for tests / DESIGN.md):
 * out[0]=heads per CTA, out[1]=token slices, out[2]=tokens per chunk,
 * out[3]=pipeline stages, out[4]=dynamic smem bytes, out[5]=head groups. */
int b200attn_decode_geometry(int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                             int32_t kv_dtype, int32_t out[6]);
/* ABI 6.  The geometry b200attn_plan / b200attn_decode use for a decode batch of `batch` requests: small
 * batches (fewer than 64 request x head-group pairs) take fewer kv heads per CTA, down to 256-byte row slices,
 * so that a request is cut across 2-3 CTA ranges instead of ~10 and its partials merge in registers.  Both
 * calls derive it from (shape, batch): a plan is only valid for the batch size it was built for (it always
 * was: the plan buffer's layout depends on `batch`).  batch = 0 gives b200attn_decode_geometry's answer. */
int b200attn_decode_geometry_batch(int32_t batch, int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                                   int32_t kv_dtype, int32_t out[6]);

#ifdef __cplusplus
}
#endif
#endif /* B200ATTN_H_ */
