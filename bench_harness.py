"""Mock ModelRunner / layer / ForwardBatch around a synthetic workload (``synth.py``), shared by bench.py,
the tools and the tests -- the way the reference's LTP numerical tests fabricate one
(python/sglang/test/numerical_tests/modules/sglang/test_attention.py:164-326).  Not product code, not test
code: the objects here stand in for what sglang's ModelRunner hands an attention backend
(triton_backend.py:62-122)."""
from __future__ import annotations

import types

import torch

import synth

_TORCH_KV = {"bf16": torch.bfloat16, "f16": torch.float16, "fp8_e4m3": torch.float8_e4m3fn,
             "fp8_e5m2": torch.float8_e5m2}


def make_runner(w: synth.Workload, device, tp_size: int = 1):
    """Mock model_runner exposing what attention backends read (triton_backend.py:62-122)."""
    import b200attn

    c = w.cfg
    r2t = b200attn.ReqToTokenPool(w.req_to_token.shape[0], w.req_to_token.shape[1], device)
    r2t.req_to_token.copy_(w.req_to_token)
    pool = b200attn.MHATokenToKVPool(w.pool_size, c.page_size, _TORCH_KV[c.kv_dtype], c.num_kv_heads,
                                     c.head_dim, layer_num=1, device=device)
    assert pool.k_buffer[0].shape == w.k_pool.shape, (pool.k_buffer[0].shape, w.k_pool.shape)
    # adopt the workload's pool storage (no copy: full-size pools are GiBs)
    pool.k_buffer[0] = w.k_pool
    pool.v_buffer[0] = w.v_pool
    mc = types.SimpleNamespace(num_attention_heads=c.num_q_heads * tp_size,
                               get_num_kv_heads=lambda tp: c.num_kv_heads,
                               context_len=w.req_to_token.shape[1] - 4, is_encoder_decoder=False)
    return types.SimpleNamespace(device=str(device), req_to_token_pool=r2t, token_to_kv_pool=pool,
                                 model_config=mc, sliding_window_size=None, attention_tp_size=tp_size)


def make_layer(w: synth.Workload, logit_cap: float = 0.0, with_scales: bool = True):
    import b200attn

    c = w.cfg
    layer = b200attn.AttnLayer(c.num_q_heads, c.head_dim, w.sm_scale, c.num_kv_heads, layer_id=0,
                               logit_cap=logit_cap)
    if c.kv_dtype.startswith("fp8") and with_scales:
        dev = w.k_pool.device
        layer.k_scale = torch.tensor(c.k_scale, dtype=torch.float32, device=dev)
        layer.v_scale = torch.tensor(c.v_scale, dtype=torch.float32, device=dev)
        layer.k_scale_float, layer.v_scale_float = c.k_scale, c.v_scale
    return layer


def make_batch(w: synth.Workload, runner, backend, index_dtype=torch.int64):
    import b200attn

    return b200attn.ForwardBatch(
        forward_mode=b200attn.ForwardMode.DECODE, batch_size=w.cfg.batch,
        req_pool_indices=w.req_pool_indices.to(index_dtype), seq_lens=w.seq_lens.to(index_dtype),
        out_cache_loc=w.out_cache_loc, seq_lens_sum=w.seq_lens_sum,
        req_to_token_pool=runner.req_to_token_pool, token_to_kv_pool=runner.token_to_kv_pool,
        attn_backend=backend)



def make_extend_runner(w, device, sliding_window_size=None, num_draft_tokens=None):
    """Mock model_runner around a ``synth.ExtendWorkload`` (see ``make_runner``)."""
    import b200attn

    c = w.cfg
    r2t = b200attn.ReqToTokenPool(w.req_to_token.shape[0], w.req_to_token.shape[1], device)
    r2t.req_to_token.copy_(w.req_to_token)
    pool = b200attn.MHATokenToKVPool(w.pool_size, c.page_size, _TORCH_KV[c.kv_dtype], c.num_kv_heads,
                                     c.head_dim, layer_num=1, device=device)
    assert pool.k_buffer[0].shape == w.k_pool.shape, (pool.k_buffer[0].shape, w.k_pool.shape)
    pool.k_buffer[0] = w.k_pool
    pool.v_buffer[0] = w.v_pool
    mc = types.SimpleNamespace(num_attention_heads=c.num_q_heads, get_num_kv_heads=lambda tp: c.num_kv_heads,
                               context_len=w.req_to_token.shape[1] - 4, is_encoder_decoder=False)
    sa = types.SimpleNamespace(speculative_num_draft_tokens=num_draft_tokens)
    return types.SimpleNamespace(device=str(device), req_to_token_pool=r2t, token_to_kv_pool=pool,
                                 model_config=mc, sliding_window_size=sliding_window_size,
                                 attention_tp_size=1, server_args=sa)
