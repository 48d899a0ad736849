// K4: fused rotary embedding + KV store.  One launch for what LlamaAttention.forward does
// between the qkv projection and attention (reference: python/sglang/srt/models/llama.py:186-189):
//   rotary_emb(positions, q, k)   layers/rotary_embedding.py:218-236 -> sgl-kernel/csrc/elementwise/rope.cu:22-89
//   set_kv_buffer(loc, k, v)      mem_cache/memory_pool.py:369-407
// RoPE arithmetic: o1 = x1 cos - x2 sin, o2 = x2 cos + x1 sin in fp32 on the fp32 cos|sin cache
// row of the token's position, one rounding to the source dtype (the reference's CUDA path keeps the
// cache in fp32 for that reason, rotary_embedding.py:100-103).  q and k are updated in place; the
// rotated k and v then go through exactly the store arithmetic of kv_store.cu.
//
// HBM-bound byte work: one thread owns one 16-byte vector (interleaved layout, pass-through dims,
// v) or one pair of vectors (neox layout: columns c and c + rotary_dim / 2).
#include "store_common.cuh"

namespace b200attn {

struct RopeParams {
  void* q;
  int64_t q_st, q_sh;
  void* kv[2];           // k, v
  int64_t kv_st[2], kv_sh[2];
  void* pool[2];
  int64_t pool_stride[2];
  const int64_t* loc;
  int64_t pool_slots;
  const void* positions;
  const float* cache;
  int64_t cache_stride;
  int64_t max_position;
  int32_t n_tokens, hq, hkv, dim, rot;
  int32_t pos64, neox, save_kv;
  int32_t scale_mode, mutate;
  float scale[2];
  const float* scale_ptr[2];
  int32_t e5m2;          // fp8 pools: 1 = float8_e5m2
};

template <typename T>
__device__ __forceinline__ void load8(T (&x)[8], const T* p) {
  *reinterpret_cast<uint4*>(x) = *reinterpret_cast<const uint4*>(p);
}
template <typename T>
__device__ __forceinline__ void store8(T* p, const T (&x)[8]) {
  *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(x);
}

// kv_store.cu's per-vector store: optional scale division (written back when `mutate`), then the
// pool write (16-bit copy or non-saturating e4m3fn cast).
template <typename T, bool FP8>
__device__ __forceinline__ void pool_store8(const RopeParams& p, int which, T (&x)[8], T* src, bool src_dirty,
                                            int64_t slot, int head, int col) {
  if constexpr (FP8) {
    if (p.scale_mode != B200ATTN_SCALE_NONE) {
      const float s = p.scale_ptr[which] ? to_f32<T>(from_f32<T>(*p.scale_ptr[which])) : p.scale[which];
      const float inv = __fdiv_rn(1.0f, s);
      const bool recip = p.scale_mode == B200ATTN_SCALE_RECIP_MUL;
#pragma unroll
      for (int e = 0; e < 8; ++e) x[e] = div_scale<T>(x[e], s, inv, recip);
      src_dirty = src_dirty || p.mutate;
    }
  }
  if (src_dirty) store8(src, x);
  if (!p.save_kv || slot < 0 || slot >= p.pool_slots) return;
  if constexpr (FP8) {
    uint8_t* dst = reinterpret_cast<uint8_t*>(p.pool[which]) + slot * p.pool_stride[which] + head * p.dim + col;
    uint8_t y[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) y[e] = f32_to_fp8(to_f32<T>(x[e]), p.e5m2 != 0);
    *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(y);
  } else {
    T* dst = reinterpret_cast<T*>(p.pool[which]) + slot * p.pool_stride[which] + head * p.dim + col;
    store8(dst, x);
  }
}

template <typename T, bool FP8>
__global__ void __launch_bounds__(256) rope_kv_store_kernel(const RopeParams p) {
  const int ru = p.neox ? p.rot / 16 : p.rot / 8;        // rotate units per head
  const int pu = (p.dim - p.rot) / 8;                    // pass-through units per head
  const int nq = p.hq * ru;
  const int nk = p.hkv * ru;
  const int nkp = p.save_kv ? p.hkv * pu : 0;
  const int nv = p.save_kv ? p.hkv * (p.dim / 8) : 0;
  const int per_token = nq + nk + nkp + nv;
  const int64_t total = static_cast<int64_t>(p.n_tokens) * per_token;
  const int half = p.rot / 2;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int token = static_cast<int>(i / per_token);
    int r = static_cast<int>(i - static_cast<int64_t>(token) * per_token);
    const int64_t slot = p.save_kv ? p.loc[token] : -1;
    if (r < nq + nk) {
      // ---- rotate (q or k) ----
      const bool is_k = r >= nq;
      if (is_k) r -= nq;
      const int head = r / ru;
      const int unit = r - head * ru;
      T* base = is_k ? reinterpret_cast<T*>(p.kv[0]) + token * p.kv_st[0] + head * p.kv_sh[0]
                     : reinterpret_cast<T*>(p.q) + token * p.q_st + head * p.q_sh;
      int64_t pos = p.pos64 ? reinterpret_cast<const int64_t*>(p.positions)[token]
                            : static_cast<int64_t>(reinterpret_cast<const int32_t*>(p.positions)[token]);
      pos = pos < 0 ? 0 : (pos >= p.max_position ? p.max_position - 1 : pos);
      const float* cs = p.cache + pos * p.cache_stride;
      if (p.neox) {
        const int c = unit * 8;
        T x1[8], x2[8];
        load8(x1, base + c);
        load8(x2, base + c + half);
        float co[8], si[8];
        *reinterpret_cast<float4*>(co) = *reinterpret_cast<const float4*>(cs + c);
        *reinterpret_cast<float4*>(co + 4) = *reinterpret_cast<const float4*>(cs + c + 4);
        *reinterpret_cast<float4*>(si) = *reinterpret_cast<const float4*>(cs + half + c);
        *reinterpret_cast<float4*>(si + 4) = *reinterpret_cast<const float4*>(cs + half + c + 4);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float a = to_f32<T>(x1[e]), b = to_f32<T>(x2[e]);
          x1[e] = from_f32<T>(__fsub_rn(__fmul_rn(a, co[e]), __fmul_rn(b, si[e])));
          x2[e] = from_f32<T>(__fadd_rn(__fmul_rn(b, co[e]), __fmul_rn(a, si[e])));
        }
        if (is_k) {
          pool_store8<T, FP8>(p, 0, x1, base + c, true, slot, head, c);
          pool_store8<T, FP8>(p, 0, x2, base + c + half, true, slot, head, c + half);
        } else {
          store8(base + c, x1);
          store8(base + c + half, x2);
        }
      } else {
        const int c = unit * 8;  // pairs (c + 2i, c + 2i + 1) use cos/sin[c / 2 + i]
        T x[8];
        load8(x, base + c);
        float co[4], si[4];
        *reinterpret_cast<float4*>(co) = *reinterpret_cast<const float4*>(cs + c / 2);
        *reinterpret_cast<float4*>(si) = *reinterpret_cast<const float4*>(cs + half + c / 2);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float a = to_f32<T>(x[2 * e]), b = to_f32<T>(x[2 * e + 1]);
          x[2 * e] = from_f32<T>(__fsub_rn(__fmul_rn(a, co[e]), __fmul_rn(b, si[e])));
          x[2 * e + 1] = from_f32<T>(__fadd_rn(__fmul_rn(b, co[e]), __fmul_rn(a, si[e])));
        }
        if (is_k) pool_store8<T, FP8>(p, 0, x, base + c, true, slot, head, c);
        else store8(base + c, x);
      }
      continue;
    }
    r -= nq + nk;
    // ---- pass-through dims of k, and v: plain store ----
    int which, head, col;
    if (r < nkp) {
      which = 0;
      head = r / pu;
      col = p.rot + (r - head * pu) * 8;
    } else {
      r -= nkp;
      which = 1;
      const int vu = p.dim / 8;
      head = r / vu;
      col = (r - head * vu) * 8;
    }
    T* src = reinterpret_cast<T*>(p.kv[which]) + token * p.kv_st[which] + head * p.kv_sh[which] + col;
    T x[8];
    load8(x, src);
    pool_store8<T, FP8>(p, which, x, src, false, slot, head, col);
  }
}

static bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace b200attn

using namespace b200attn;

extern "C" __attribute__((visibility("default"))) int b200attn_rope_kv_store(const b200attn_rope_kv_store_args* a,
                                                                            b200attn_stream_t stream_) {
  if (!a) return B200ATTN_E_INVALID;
  const b200attn_kv_store_args& s = a->kv;
  if (s.n_tokens < 0 || s.num_kv_heads <= 0 || a->num_q_heads <= 0 || s.head_dim <= 0) return B200ATTN_E_INVALID;
  if (s.n_tokens == 0) return B200ATTN_OK;
  if (!a->q || !s.k || !a->positions || !a->cos_sin_cache || a->max_position <= 0) return B200ATTN_E_INVALID;
  if (a->pos_dtype != B200ATTN_I32 && a->pos_dtype != B200ATTN_I64) return B200ATTN_E_INVALID;
  if (s.src_dtype != B200ATTN_BF16 && s.src_dtype != B200ATTN_F16) return B200ATTN_E_UNSUPPORTED;
  const bool save = a->save_kv != 0;
  const bool fp8 = save && (s.pool_dtype == B200ATTN_FP8_E4M3 || s.pool_dtype == B200ATTN_FP8_E5M2);
  if (save) {
    if (!s.v || !s.loc || !s.k_pool || !s.v_pool) return B200ATTN_E_INVALID;
    if (!fp8 && s.pool_dtype != s.src_dtype) return B200ATTN_E_UNSUPPORTED;
    if (s.v_head_dim != s.head_dim) return B200ATTN_E_UNSUPPORTED;
    if (fp8 && (s.scale_mode < B200ATTN_SCALE_NONE || s.scale_mode > B200ATTN_SCALE_RECIP_MUL))
      return B200ATTN_E_INVALID;
  }
  const int d = s.head_dim, rot = a->rotary_dim;
  if (d % 8 != 0 || rot <= 0 || rot > d || rot % (a->is_neox ? 16 : 8) != 0) return B200ATTN_E_UNSUPPORTED;
  // 16-byte vectors everywhere
  if (!al16(a->q) || !al16(s.k) || a->q_stride_token % 8 || a->q_stride_head % 8 || s.k_stride_token % 8 ||
      s.k_stride_head % 8 || !al16(a->cos_sin_cache) || a->cache_stride % 4)
    return B200ATTN_E_ALIGN;
  if (save) {
    const int pa = fp8 ? 8 : 16;
    if (!al16(s.v) || s.v_stride_token % 8 || s.v_stride_head % 8 ||
        reinterpret_cast<uintptr_t>(s.k_pool) % pa || reinterpret_cast<uintptr_t>(s.v_pool) % pa ||
        s.k_pool_stride_slot % 8 || s.v_pool_stride_slot % 8)
      return B200ATTN_E_ALIGN;
  }

  RopeParams p;
  p.q = a->q;
  p.q_st = a->q_stride_token;
  p.q_sh = a->q_stride_head;
  p.kv[0] = s.k;
  p.kv[1] = s.v;
  p.kv_st[0] = s.k_stride_token;
  p.kv_st[1] = s.v_stride_token;
  p.kv_sh[0] = s.k_stride_head;
  p.kv_sh[1] = s.v_stride_head;
  p.pool[0] = s.k_pool;
  p.pool[1] = s.v_pool;
  p.pool_stride[0] = s.k_pool_stride_slot;
  p.pool_stride[1] = s.v_pool_stride_slot;
  p.loc = s.loc;
  p.pool_slots = s.pool_slots;
  p.positions = a->positions;
  p.cache = a->cos_sin_cache;
  p.cache_stride = a->cache_stride;
  p.max_position = a->max_position;
  p.n_tokens = s.n_tokens;
  p.hq = a->num_q_heads;
  p.hkv = s.num_kv_heads;
  p.dim = d;
  p.rot = rot;
  p.pos64 = a->pos_dtype == B200ATTN_I64;
  p.neox = a->is_neox != 0;
  p.save_kv = save;
  p.scale_mode = fp8 ? s.scale_mode : B200ATTN_SCALE_NONE;
  p.e5m2 = save && s.pool_dtype == B200ATTN_FP8_E5M2;
  p.mutate = s.mutate_src;
  p.scale[0] = s.k_scale;
  p.scale[1] = s.v_scale;
  p.scale_ptr[0] = s.k_scale_ptr;
  p.scale_ptr[1] = s.v_scale_ptr;

  const int ru = p.neox ? rot / 16 : rot / 8;
  const int64_t per_token = static_cast<int64_t>(p.hq + p.hkv) * ru +
                            (save ? static_cast<int64_t>(p.hkv) * ((d - rot) / 8 + d / 8) : 0);
  const int64_t total = per_token * p.n_tokens;
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 16;
  if (blocks > cap) blocks = cap;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const dim3 grid(static_cast<unsigned>(blocks)), block(256);
  if (s.src_dtype == B200ATTN_BF16) {
    if (fp8) rope_kv_store_kernel<__nv_bfloat16, true><<<grid, block, 0, stream>>>(p);
    else rope_kv_store_kernel<__nv_bfloat16, false><<<grid, block, 0, stream>>>(p);
  } else {
    if (fp8) rope_kv_store_kernel<__half, true><<<grid, block, 0, stream>>>(p);
    else rope_kv_store_kernel<__half, false><<<grid, block, 0, stream>>>(p);
  }
  count_launch();
  return cudaGetLastError() == cudaSuccess ? B200ATTN_OK : B200ATTN_E_CUDA;
}
