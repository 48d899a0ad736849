// Registration shim for the REFERENCE's CPU decode kernel (test / baseline infrastructure).
//
// oracle/ref_build/Makefile compiles the reference's own translation unit
//   /root/reference/sgl-kernel/csrc/cpu/decode.cpp        (never copied into this repo)
// where it lies and links it with this file into oracle/_ref/.  The only thing added here is
// the operator registration the reference does in torch_extension_cpu.cpp:263-268, under a
// private namespace so it cannot collide with an installed sgl_kernel.
#include <ATen/ATen.h>
#include <torch/library.h>

// defined in the reference: sgl-kernel/csrc/cpu/decode.cpp:1375
void decode_attention_cpu(at::Tensor& query, at::Tensor& k_buffer, at::Tensor& v_buffer, at::Tensor& output,
                          at::Tensor& key, at::Tensor& value, at::Tensor& loc, at::Tensor& attn_logits,
                          at::Tensor& req_to_token, at::Tensor& req_pool_indices, at::Tensor& seq_lens,
                          double sm_scale, double logit_cap);

static void decode_attention_cpu_boxed(at::Tensor query, at::Tensor k_buffer, at::Tensor v_buffer,
                                       at::Tensor output, at::Tensor key, at::Tensor value, at::Tensor loc,
                                       at::Tensor attn_logits, at::Tensor req_to_token,
                                       at::Tensor req_pool_indices, at::Tensor seq_lens, double sm_scale,
                                       double logit_cap) {
  decode_attention_cpu(query, k_buffer, v_buffer, output, key, value, loc, attn_logits, req_to_token,
                       req_pool_indices, seq_lens, sm_scale, logit_cap);
}

TORCH_LIBRARY_FRAGMENT(b200attn_ref, m) {
  m.def(
      "decode_attention_cpu(Tensor query, Tensor(a!) k_cache, Tensor(b!) v_cache, Tensor(c!) output, Tensor key, "
      "Tensor value, Tensor loc, Tensor(d!) attn_logits, Tensor req_to_token, Tensor req_pool_indices, "
      "Tensor seq_lens, float sm_scale, float logit_cap) -> ()");
  m.impl("decode_attention_cpu", c10::kCPU, &decode_attention_cpu_boxed);
}
