// b200attn_decode_static_ws: the decode call (K2 + K3, api.cu) for callers that may not allocate device memory
// -- code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc.  The split-KV
// workspace (partial o + lse, workspace_floats() in common.cuh) lives in a file-scope device array of
// B200ATTN_STATIC_WS_BYTES that every GPU loading the library holds until the process exits.  No counterpart in
// the reference.
//
// * One array per device, shared by every call: two calls that overlap on the GPU (different streams, or one
//   stream per thread) race on it.  Serialise them, or use b200attn_decode with a workspace per stream.
// * Nothing in the workspace has to start at zero: decode_kernel writes every partial row before any CTA reads
//   it, and the arrival counters that order those reads live in the plan buffer, which b200attn_plan resets.
//   The array therefore keeps the previous call's rows and is not cleared between calls.
// * This file calls no cudaMalloc / cudaFree / allocating library routine; the only set-up is one
//   cudaGetSymbolAddress per device (cached), made with that device current.
#include <atomic>

#include "common.cuh"

namespace {

// Host code sees a stand-in for this symbol, not a device address: only cudaGetSymbolAddress's answer may
// be handed to a kernel.
static __device__ __align__(256) unsigned char g_static_ws[B200ATTN_STATIC_WS_BYTES];

constexpr int kMaxDevices = 64;
std::atomic<void*> g_ws_addr[kMaxDevices];

// Device address of g_static_ws on the current device; nullptr when the runtime refuses.
void* static_ws_for_current_device() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    (void)cudaGetLastError();
    return nullptr;
  }
  void* ptr = g_ws_addr[dev].load(std::memory_order_acquire);
  if (ptr) return ptr;
  if (cudaGetSymbolAddress(&ptr, g_static_ws) != cudaSuccess || !ptr) {
    (void)cudaGetLastError();
    return nullptr;
  }
  g_ws_addr[dev].store(ptr, std::memory_order_release);  // same value from every thread that races here
  return ptr;
}

}  // namespace

extern "C" __attribute__((visibility("default"))) int b200attn_decode_static_ws(const b200attn_decode_args* a,
                                                                                b200attn_stream_t stream) {
  if (!a) return B200ATTN_E_INVALID;
  if (a->ws_o != nullptr || a->ws_bytes != 0) return B200ATTN_E_INVALID;  // the library owns the workspace here
  if (a->batch < 0) return B200ATTN_E_INVALID;
  if (a->batch == 0) return B200ATTN_OK;
  const size_t need = b200attn_workspace_bytes(a->batch, a->num_q_heads, a->num_kv_heads, a->head_dim,
                                               a->v_head_dim, a->kv_dtype);
  if (need == 0) return B200ATTN_E_UNSUPPORTED;
  if (need > static_cast<size_t>(B200ATTN_STATIC_WS_BYTES)) return B200ATTN_E_STATIC_LIMIT;
  void* ws = static_ws_for_current_device();
  if (!ws) return B200ATTN_E_CUDA;
  b200attn_decode_args b = *a;
  b.ws_o = static_cast<float*>(ws);
  b.ws_bytes = B200ATTN_STATIC_WS_BYTES;
  return b200attn_decode(&b, stream);
}
