// C-ABI entry points that are not tied to one kernel file: the decode call
// (K2 + K3), geometry / workspace queries, error strings, launch counter.
#include <atomic>
#include <cstdlib>

#include "decode_kernel.cuh"
#include "merge.cuh"

namespace b200attn {

static std::atomic<uint64_t> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

int sm_count() {
  // Per-device cache; no process-global state keyed without the device id.
  static std::atomic<int> cache[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    (void)cudaGetLastError();
    return 148;
  }
  int n = cache[dev].load(std::memory_order_relaxed);
  if (n > 0) return n;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) {
    (void)cudaGetLastError();
    return 148;
  }
  cache[dev].store(n, std::memory_order_relaxed);
  return n;
}

}  // namespace b200attn

using namespace b200attn;

#ifdef B200ATTN_DEC_TRACE
// Trace build only (make trace): where decode_kernel drops its per-CTA clock stamps; [148][16] uint64.
static unsigned long long* g_decode_trace = nullptr;
extern "C" __attribute__((visibility("default"))) void b200attn_debug_set_decode_trace(void* buf) {
  g_decode_trace = static_cast<unsigned long long*>(buf);
}
#endif

extern "C" __attribute__((visibility("default"))) int b200attn_abi_version(void) { return B200ATTN_ABI_VERSION; }

extern "C" __attribute__((visibility("default"))) uint64_t b200attn_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" __attribute__((visibility("default"))) const char* b200attn_strerror(int code) {
  switch (code) {
    case B200ATTN_OK: return "ok";
    case B200ATTN_E_INVALID: return "invalid argument (null pointer, negative size or bad enum)";
    case B200ATTN_E_UNSUPPORTED: return "unsupported shape or dtype combination";
    case B200ATTN_E_ALIGN: return "pointer or stride violates an alignment requirement";
    case B200ATTN_E_CUDA: return "CUDA runtime reported a launch error";
    case B200ATTN_E_WORKSPACE: return "plan or workspace buffer too small";
    case B200ATTN_E_STATIC_LIMIT: return "call needs more scratch than the library's static workspace (B200ATTN_STATIC_WS_BYTES, 64 MiB): use b200attn_decode with a caller-owned workspace";
    default: return "unknown b200attn error code";
  }
}

extern "C" __attribute__((visibility("default"))) int b200attn_decode_geometry(int32_t hq, int32_t hkv, int32_t d, int32_t kv_dtype,
                                        int32_t out[6]) {
  if (!out) return B200ATTN_E_INVALID;
  Geometry g;
  if (!make_geometry(hq, hkv, d, kv_dtype, &g)) return B200ATTN_E_UNSUPPORTED;
  out[0] = g.hc;
  out[1] = g.ts;
  out[2] = g.chunk_tokens;
  out[3] = g.stages;
  out[4] = g.smem_bytes;
  out[5] = g.ng;
  return B200ATTN_OK;
}

extern "C" __attribute__((visibility("default"))) int b200attn_decode_geometry_batch(int32_t batch, int32_t hq, int32_t hkv, int32_t d,
                                        int32_t kv_dtype, int32_t out[6]) {
  if (!out || batch < 0) return B200ATTN_E_INVALID;
  Geometry g;
  if (!make_geometry(hq, hkv, d, kv_dtype, &g, batch)) return B200ATTN_E_UNSUPPORTED;
  out[0] = g.hc;
  out[1] = g.ts;
  out[2] = g.chunk_tokens;
  out[3] = g.stages;
  out[4] = g.smem_bytes;
  out[5] = g.ng;
  return B200ATTN_OK;
}

extern "C" __attribute__((visibility("default"))) size_t b200attn_workspace_bytes(int32_t batch, int32_t hq, int32_t hkv, int32_t d,
                                           int32_t dv, int32_t kv_dtype) {
  Geometry g;
  if (batch < 0 || dv <= 0 || !make_geometry(hq, hkv, d, kv_dtype, &g)) return 0;
  return workspace_floats(batch, hq, dv, g, nullptr) * sizeof(float);
}

extern "C" __attribute__((visibility("default"))) int b200attn_decode(const b200attn_decode_args* a, b200attn_stream_t stream_) {
  if (!a) return B200ATTN_E_INVALID;
  if (a->batch < 0) return B200ATTN_E_INVALID;
  if (a->batch == 0) return B200ATTN_OK;
  if (a->phases < 0 || a->phases > 2) return B200ATTN_E_INVALID;
  if (!a->q || !a->k_pool || !a->v_pool || !a->o || !a->plan || !a->indices || !a->ws_o)
    return B200ATTN_E_INVALID;
  if (a->q_dtype != B200ATTN_BF16 && a->q_dtype != B200ATTN_F16) return B200ATTN_E_UNSUPPORTED;
  const bool fp8 = is_fp8(a->kv_dtype);
  if (!fp8 && a->kv_dtype != a->q_dtype) return B200ATTN_E_UNSUPPORTED;
  if (a->v_head_dim != a->head_dim) return B200ATTN_E_UNSUPPORTED;
  if (!(a->sm_scale > 0.f) || !(a->k_scale > 0.f) || a->logit_cap < 0.f) return B200ATTN_E_INVALID;
  Geometry g;
  if (!make_geometry(a->num_q_heads, a->num_kv_heads, a->head_dim, a->kv_dtype, &g, a->batch))
    return B200ATTN_E_UNSUPPORTED;
  const int e = elem_bytes(a->kv_dtype);
  // bulk copies move 16-byte units; q fragments are read as 32-bit pairs
  if ((reinterpret_cast<uintptr_t>(a->k_pool) | reinterpret_cast<uintptr_t>(a->v_pool)) & 15)
    return B200ATTN_E_ALIGN;
  if ((a->k_pool_stride_slot * e) % 16 != 0 || (a->v_pool_stride_slot * e) % 16 != 0)
    return B200ATTN_E_ALIGN;
  // the producers form a row address as base + slot (u32) x slot bytes (u32)
  if (a->k_pool_stride_slot <= 0 || a->v_pool_stride_slot <= 0 || a->k_pool_stride_slot * e > 0xffffffffll ||
      a->v_pool_stride_slot * e > 0xffffffffll)
    return B200ATTN_E_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(a->q) & 3) || a->q_stride_batch % 2 != 0 || a->q_stride_head % 2 != 0)
    return B200ATTN_E_ALIGN;
  if ((reinterpret_cast<uintptr_t>(a->plan) & 15) || (reinterpret_cast<uintptr_t>(a->ws_o) & 15))
    return B200ATTN_E_ALIGN;

  int n_sm = sm_count();
  if (n_sm > kMaxRanges / 2) n_sm = kMaxRanges / 2;
  size_t lse_off = 0;
  const size_t need = workspace_floats(a->batch, a->num_q_heads, a->v_head_dim, g, &lse_off);
  if (a->ws_bytes < need * sizeof(float)) return B200ATTN_E_WORKSPACE;

  DecodeParams p;
  p.q = a->q;
  p.q_sb = a->q_stride_batch;
  p.q_sh = a->q_stride_head;
  p.k_pool = static_cast<const uint8_t*>(a->k_pool);
  p.v_pool = static_cast<const uint8_t*>(a->v_pool);
  p.k_slot_bytes = a->k_pool_stride_slot * e;
  p.v_slot_bytes = a->v_pool_stride_slot * e;
  p.o = a->o;
  p.o_sb = a->o_stride_batch;
  p.o_sh = a->o_stride_head;
  p.lse = a->lse;
  p.plan = a->plan;
  p.indices = a->indices;
  p.slot_map = a->slot_map;
  p.pool_slots = a->pool_slots;
  p.ws_o = a->ws_o;
  p.ws_lse = a->ws_o + lse_off;
  p.batch = a->batch;
  p.hq = a->num_q_heads;
  p.hc = g.hc;
  p.ts = g.ts;
  p.ng = g.ng;
  p.group = g.group;
  p.row_bytes = g.row_bytes;
  p.row_stride = g.row_stride;
  p.stage_bytes = g.stage_bytes;
  p.stages = g.stages;
  p.chunk_tokens = g.chunk_tokens;
  p.fuse_merge = a->phases == 0;
  p.cp_async = g.cp_async;
  p.swizzle = g.swizzle;
  p.kv_split = g.kv_split;
  p.subs = g.subs;
  p.e5m2 = a->kv_dtype == B200ATTN_FP8_E5M2;
  p.xch_off = a->phases == 0 ? g.xch_off : 0;  // K2 + K3 as separate launches keep every slice in the workspace
  p.unit_shift = 0;
  while ((16 << p.unit_shift) < g.row_bytes) ++p.unit_shift;
  p.sm_scale_k = a->sm_scale * a->k_scale;
  p.scale_log2 = p.sm_scale_k * kLog2e;
  p.logit_cap = a->logit_cap;
  p.v_scale = a->v_scale;
#ifdef B200ATTN_DEC_TRACE
  p.trace = g_decode_trace;
#endif

  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const bool rows16 = g.group > 8;
  const bool half = a->q_dtype == B200ATTN_F16;
  cudaError_t err = cudaSuccess;
  if (a->phases == 2) {
    // merge only
  } else if (half) {
    err = fp8 ? launch_decode<__half, true>(p, a->head_dim, rows16, n_sm, g.smem_bytes, stream)
              : launch_decode<__half, false>(p, a->head_dim, rows16, n_sm, g.smem_bytes, stream);
  } else {
    err = fp8 ? launch_decode<__nv_bfloat16, true>(p, a->head_dim, rows16, n_sm, g.smem_bytes, stream)
              : launch_decode<__nv_bfloat16, false>(p, a->head_dim, rows16, n_sm, g.smem_bytes, stream);
  }
  if (err != cudaSuccess) return B200ATTN_E_CUDA;
  if (a->phases != 2) count_launch();
  if (a->phases != 2) return B200ATTN_OK;  // phases 0: decode_kernel merged the partials itself

  MergeParams m;
  m.plan = a->plan;
  m.ws_o = p.ws_o;
  m.ws_lse = p.ws_lse;
  m.o = a->o;
  m.o_sb = a->o_stride_batch;
  m.o_sh = a->o_stride_head;
  m.lse = a->lse;
  m.batch = a->batch;
  m.hq = a->num_q_heads;
  m.ng = g.ng;
  m.ts = g.ts;
  m.dv = a->v_head_dim;
  if (launch_merge(m, half, stream) != cudaSuccess) return B200ATTN_E_CUDA;
  count_launch();
  return B200ATTN_OK;
}
