"""Shared test/bench harness: a mock ModelRunner around a synthetic workload, the way the
reference's LTP numerical tests fabricate one
(python/sglang/test/numerical_tests/modules/sglang/test_attention.py:164-326)."""
from __future__ import annotations

import os
import types

import numpy as np
import torch

import synth
from bench_harness import make_batch, make_extend_runner, make_layer, make_runner  # noqa: F401  (re-exported)

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

def run_decode(backend, runner, w: synth.Workload, layer=None, save_kv_cache=True,
               index_dtype=torch.int64):
    """init_forward_metadata + RadixAttention-style layer call; returns [B, Hq * D]."""
    layer = layer or make_layer(w)
    fb = make_batch(w, runner, backend, index_dtype)
    backend.init_forward_metadata(fb)
    return layer(w.q, w.k, w.v, fb, save_kv_cache=save_kv_cache)


def oracle_step(w_cpu: synth.Workload, logit_cap: float = 0.0, save_kv: bool = True,
                scale_mode=None, scale_is_cuda_tensor: bool = True):
    """kv_store + decode_attention on the CPU oracle; returns (out, lse, k_pool, v_pool, k, v)."""
    from oracle import decode_oracle as oracle

    c = w_cpu.cfg
    kp, vp = w_cpu.k_pool.clone(), w_cpu.v_pool.clone()
    if c.kv_dtype == "fp8_e5m2":   # fp8 pools are byte tensors; the oracle tells the kind by the view (uint8 = e4m3)
        kp, vp = kp.view(torch.float8_e5m2), vp.view(torch.float8_e5m2)
    k = w_cpu.k.reshape(c.batch, c.num_kv_heads, c.head_dim).clone()
    v = w_cpu.v.reshape(c.batch, c.num_kv_heads, c.head_dim).clone()
    fp8 = c.kv_dtype.startswith("fp8")
    if save_kv:
        oracle.kv_store(kp, vp, w_cpu.out_cache_loc, k, v, c.k_scale if fp8 else None,
                        c.v_scale if fp8 else None,
                        oracle.SCALE_TRUE_DIV if scale_mode is None else scale_mode,
                        scale_is_cuda_tensor)
    out, lse = oracle.decode_attention(
        w_cpu.q.reshape(c.batch, c.num_q_heads, c.head_dim), kp, vp, w_cpu.req_to_token,
        w_cpu.req_pool_indices, w_cpu.seq_lens, w_cpu.sm_scale, logit_cap,
        c.k_scale if fp8 else 1.0, c.v_scale if fp8 else 1.0)
    return out, lse, kp, vp, k, v


def assert_close_out(got: np.ndarray, ref: np.ndarray, atol=2e-2, rtol=1e-2, what="attention output", rel_l2=2e-2):
    """BASELINE.json tolerance: max-abs 2e-2 / rel 1e-2 against the fp32+ oracle -- and, because |o| is only
    0.01-0.1 at contexts of a few hundred tokens and more (so that bar alone would pass a wrong V row or a
    swapped head-dim column), a relative L2 bound per output row (last axis = head_dim):
    ||got - ref|| <= rel_l2 * ||ref||.  A correct bf16 kernel sits at 2e-3 ... 5e-3."""
    g = got.astype(np.float64)
    r = ref.astype(np.float64)
    err = np.abs(g - r)
    bound = atol + rtol * np.abs(r)
    worst = float((err - bound).max()) if err.size else 0.0
    assert worst <= 0, f"{what}: max abs err {err.max():.4e} exceeds atol {atol} + rtol {rtol}"
    if rel_l2 is not None and err.size:
        num = np.sqrt((err ** 2).sum(axis=-1))
        den = np.sqrt((r ** 2).sum(axis=-1))
        ok = num <= rel_l2 * den + 1e-6
        assert ok.all(), (f"{what}: relative L2 error {float((num / np.maximum(den, 1e-30))[~ok].max()):.3e} of a row "
                          f"exceeds {rel_l2} ({int((~ok).sum())} rows)")
    return float(err.max()) if err.size else 0.0


def canon_e4m3_nan(b: np.ndarray) -> np.ndarray:
    """float8_e4m3fn has two NaN encodings (0x7F, 0xFF).  Which one a NaN *input* lands on after
    ``div_`` depends on the platform's bf16 rounding of NaN (torch's AVX512 path writes 0xFFFF,
    its scalar path 0x7FC0, CUDA 0x7FFF), so NaN bytes are compared up to sign."""
    b = np.asarray(b, dtype=np.uint8)
    return np.where((b & 0x7F) == 0x7F, np.uint8(0x7F), b)


# --------------------------------------------------------------------------- #
# extend (prefill with cached prefix) helpers
# --------------------------------------------------------------------------- #
def run_extend(backend, runner, w, layer, mode=None, spec_info=None, index_dtype=torch.int64,
               save_kv_cache=True):
    """init_forward_metadata + layer call for an extend-type batch; returns [tokens, Hq * D]."""
    import b200attn

    mode = mode or b200attn.ForwardMode.EXTEND
    seq_lens = w.seq_lens if not mode.is_target_verify() else w.extend_prefix_lens
    fb = b200attn.ForwardBatch(
        forward_mode=mode, batch_size=len(w.prefix_lens_np),
        req_pool_indices=w.req_pool_indices.to(index_dtype), seq_lens=seq_lens.to(index_dtype),
        out_cache_loc=w.out_cache_loc, seq_lens_sum=int(seq_lens.sum()),
        req_to_token_pool=runner.req_to_token_pool, token_to_kv_pool=runner.token_to_kv_pool,
        attn_backend=backend, spec_info=spec_info,
        extend_prefix_lens=w.extend_prefix_lens.to(index_dtype), extend_seq_lens=w.extend_seq_lens.to(index_dtype),
        extend_seq_lens_cpu=[int(x) for x in w.extend_lens_np])
    backend.init_forward_metadata(fb)
    return layer(w.q, w.k, w.v, fb, save_kv_cache=save_kv_cache)


def oracle_extend(w_cpu, causal=True, logit_cap=0.0, custom_mask=None, sliding_window=-1):
    """kv_store + extend_attention on the CPU oracle; returns (out, lse, k_pool, v_pool)."""
    from oracle import decode_oracle, extend_oracle

    c = w_cpu.cfg
    kp, vp = w_cpu.k_pool.clone(), w_cpu.v_pool.clone()
    if c.kv_dtype == "fp8_e5m2":
        kp, vp = kp.view(torch.float8_e5m2), vp.view(torch.float8_e5m2)
    T = w_cpu.tokens
    fp8 = c.kv_dtype.startswith("fp8")
    decode_oracle.kv_store(kp, vp, w_cpu.out_cache_loc, w_cpu.k.reshape(T, c.num_kv_heads, c.head_dim).clone(),
                           w_cpu.v.reshape(T, c.num_kv_heads, c.head_dim).clone(),
                           c.k_scale if fp8 else None, c.v_scale if fp8 else None)
    ps, es = extend_oracle.slots_from_req_to_token(w_cpu.req_to_token, w_cpu.req_pool_indices,
                                                   w_cpu.prefix_lens_np, w_cpu.extend_lens_np, sliding_window)
    mask_indptr = None
    if custom_mask is not None:
        _, _, mask_indptr = extend_oracle.extend_indptrs(w_cpu.prefix_lens_np, w_cpu.extend_lens_np)
    out, lse = extend_oracle.extend_attention(
        w_cpu.q.reshape(T, c.num_q_heads, c.head_dim), kp, vp, ps, es, w_cpu.sm_scale, causal=causal,
        logit_cap=logit_cap, custom_mask=custom_mask, mask_indptr=mask_indptr, sliding_window=sliding_window,
        k_scale=c.k_scale if fp8 else 1.0, v_scale=c.v_scale if fp8 else 1.0)
    return out, lse, kp, vp
