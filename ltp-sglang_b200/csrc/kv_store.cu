// K1: KV store.  Scatters this step's K/V rows into the paged pool through
// out_cache_loc, optionally dividing by the per-tensor scale and casting to
// float8_e4m3fn, bit-exactly with MHATokenToKVPool.set_kv_buffer
// (reference: python/sglang/srt/mem_cache/memory_pool.py:369-407).
//
// HBM-bound byte work: one thread moves one 16-byte vector (8 source
// elements); K and V go in the same launch.
#include "store_common.cuh"

namespace b200attn {

struct StoreParams {
  void* src[2];          // k, v
  int64_t stride_t[2];
  int64_t stride_h[2];
  void* pool[2];
  int64_t pool_stride[2];
  int32_t dim[2];        // head_dim, v_head_dim
  const int64_t* loc;
  int64_t pool_slots;
  int32_t n_tokens;
  int32_t heads;
  int32_t scale_mode;
  int32_t mutate;
  float scale[2];
  const float* scale_ptr[2];
  int32_t e5m2;          // fp8 pools: 1 = float8_e5m2, 0 = float8_e4m3fn
};

template <typename T, bool FP8, int VEC>
__global__ void __launch_bounds__(256) kv_store_kernel(StoreParams p) {
  pdl_launch_dependents();  // decode_kernel's CTAs may come up and initialise; they wait before reading the pool
  const int vk = p.heads * (p.dim[0] / VEC);
  const int vv = p.heads * (p.dim[1] / VEC);
  const int per_token = vk + vv;
  const int64_t total = static_cast<int64_t>(p.n_tokens) * per_token;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int token = static_cast<int>(i / per_token);
    int r = static_cast<int>(i - static_cast<int64_t>(token) * per_token);
    const int which = r >= vk;
    if (which) r -= vk;
    const int vec_per_head = p.dim[which] / VEC;
    const int head = r / vec_per_head;
    const int col = (r - head * vec_per_head) * VEC;
    const int64_t slot = p.loc[token];

    T* src = reinterpret_cast<T*>(p.src[which]) + token * p.stride_t[which] +
             head * p.stride_h[which] + col;
    T x[VEC];
    if constexpr (VEC == 8) {
      *reinterpret_cast<uint4*>(x) = *reinterpret_cast<const uint4*>(src);
    } else {
      x[0] = *src;
    }

    if constexpr (FP8) {
      if (p.scale_mode != B200ATTN_SCALE_NONE) {
        // A device scale is a 0-dim CUDA tensor: torch casts it to the common dtype (the
        // source dtype) before dividing, so the divisor is rounded first.
        const float s = p.scale_ptr[which] ? to_f32<T>(from_f32<T>(*p.scale_ptr[which])) : p.scale[which];
        if (p.scale_mode == B200ATTN_SCALE_RECIP_MUL) {
          const float inv = __fdiv_rn(1.0f, s);
#pragma unroll
          for (int e = 0; e < VEC; ++e) x[e] = from_f32<T>(__fmul_rn(to_f32<T>(x[e]), inv));
        } else {
#pragma unroll
          for (int e = 0; e < VEC; ++e) x[e] = from_f32<T>(__fdiv_rn(to_f32<T>(x[e]), s));
        }
        if (p.mutate) {
          if constexpr (VEC == 8) {
            *reinterpret_cast<uint4*>(src) = *reinterpret_cast<const uint4*>(x);
          } else {
            *src = x[0];
          }
        }
      }
      if (slot < 0 || slot >= p.pool_slots) continue;
      uint8_t* dst = reinterpret_cast<uint8_t*>(p.pool[which]) + slot * p.pool_stride[which] +
                     head * p.dim[which] + col;
      uint8_t y[VEC];
#pragma unroll
      for (int e = 0; e < VEC; ++e) y[e] = f32_to_fp8(to_f32<T>(x[e]), p.e5m2 != 0);
      if constexpr (VEC == 8) {
        *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(y);
      } else {
        dst[0] = y[0];
      }
    } else {
      if (slot < 0 || slot >= p.pool_slots) continue;
      T* dst = reinterpret_cast<T*>(p.pool[which]) + slot * p.pool_stride[which] +
               head * p.dim[which] + col;
      if constexpr (VEC == 8) {
        *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(x);
      } else {
        dst[0] = x[0];
      }
    }
  }
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace b200attn

using namespace b200attn;

extern "C" __attribute__((visibility("default"))) int b200attn_kv_store(const b200attn_kv_store_args* a, b200attn_stream_t stream_) {
  if (!a) return B200ATTN_E_INVALID;
  if (a->n_tokens < 0 || a->num_kv_heads <= 0 || a->head_dim <= 0 || a->v_head_dim <= 0)
    return B200ATTN_E_INVALID;
  if (a->n_tokens == 0) return B200ATTN_OK;
  if (!a->k || !a->v || !a->loc || !a->k_pool || !a->v_pool) return B200ATTN_E_INVALID;
  if (a->src_dtype != B200ATTN_BF16 && a->src_dtype != B200ATTN_F16) return B200ATTN_E_UNSUPPORTED;
  const bool fp8 = a->pool_dtype == B200ATTN_FP8_E4M3 || a->pool_dtype == B200ATTN_FP8_E5M2;
  if (!fp8 && a->pool_dtype != a->src_dtype) return B200ATTN_E_UNSUPPORTED;
  if (fp8 && (a->scale_mode < B200ATTN_SCALE_NONE || a->scale_mode > B200ATTN_SCALE_RECIP_MUL))
    return B200ATTN_E_INVALID;

  StoreParams p;
  p.src[0] = a->k;
  p.src[1] = a->v;
  p.stride_t[0] = a->k_stride_token;
  p.stride_t[1] = a->v_stride_token;
  p.stride_h[0] = a->k_stride_head;
  p.stride_h[1] = a->v_stride_head;
  p.pool[0] = a->k_pool;
  p.pool[1] = a->v_pool;
  p.pool_stride[0] = a->k_pool_stride_slot;
  p.pool_stride[1] = a->v_pool_stride_slot;
  p.dim[0] = a->head_dim;
  p.dim[1] = a->v_head_dim;
  p.loc = a->loc;
  p.pool_slots = a->pool_slots;
  p.n_tokens = a->n_tokens;
  p.heads = a->num_kv_heads;
  p.scale_mode = fp8 ? a->scale_mode : B200ATTN_SCALE_NONE;
  p.mutate = a->mutate_src;
  p.scale[0] = a->k_scale;
  p.scale[1] = a->v_scale;
  p.scale_ptr[0] = a->k_scale_ptr;
  p.scale_ptr[1] = a->v_scale_ptr;
  p.e5m2 = a->pool_dtype == B200ATTN_FP8_E5M2;

  // 16-byte vectors need 8-element granularity everywhere; otherwise go scalar.
  bool vec = a->head_dim % 8 == 0 && a->v_head_dim % 8 == 0 && aligned16(a->k) &&
             aligned16(a->v) && a->k_stride_token % 8 == 0 && a->k_stride_head % 8 == 0 &&
             a->v_stride_token % 8 == 0 && a->v_stride_head % 8 == 0;
  const int pool_align = fp8 ? 8 : 16;
  vec = vec && (reinterpret_cast<uintptr_t>(a->k_pool) % pool_align == 0) &&
        (reinterpret_cast<uintptr_t>(a->v_pool) % pool_align == 0) &&
        a->k_pool_stride_slot % 8 == 0 && a->v_pool_stride_slot % 8 == 0;

  const int v = vec ? 8 : 1;
  const int64_t total =
      static_cast<int64_t>(a->n_tokens) * a->num_kv_heads * ((a->head_dim + a->v_head_dim) / v);
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 16;
  if (blocks > cap) blocks = cap;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const dim3 grid(static_cast<unsigned>(blocks)), block(256);

#define LAUNCH(T, F8, V) kv_store_kernel<T, F8, V><<<grid, block, 0, stream>>>(p)
  if (a->src_dtype == B200ATTN_BF16) {
    if (fp8) { if (vec) LAUNCH(__nv_bfloat16, true, 8); else LAUNCH(__nv_bfloat16, true, 1); }
    else     { if (vec) LAUNCH(__nv_bfloat16, false, 8); else LAUNCH(__nv_bfloat16, false, 1); }
  } else {
    if (fp8) { if (vec) LAUNCH(__half, true, 8); else LAUNCH(__half, true, 1); }
    else     { if (vec) LAUNCH(__half, false, 8); else LAUNCH(__half, false, 1); }
  }
#undef LAUNCH
  count_launch();
  return cudaGetLastError() == cudaSuccess ? B200ATTN_OK : B200ATTN_E_CUDA;
}
