"""The timed CPU baseline must itself be right: ``oracle/_ref`` (the reference's own
``decode_attention_cpu``, compiled from sgl-kernel/csrc/cpu/decode.cpp where it lies) against the float64
oracle, on the reference's own shape list, seed and tolerance (test/srt/cpu/test_decode.py:62-166:
seq_len 1024, atol 3e-2, cosine > 0.99), with the pools compared after the fused store -- a mis-ordered
argument in ``oracle/ref_cpu_kernel.py`` would otherwise time garbage unnoticed, and bench.py's
``--impl reference`` arm hangs on it.  Also on a bench-shaped paged workload (permuted slots, GQA 4)."""
from __future__ import annotations

import numpy as np
import pytest
import torch

import synth
from oracle import decode_oracle, ref_cpu_kernel


def _load():
    try:
        return ref_cpu_kernel.load()
    except Exception as e:   # not built (no reference tree at build time) or a CPU that cannot run it
        pytest.skip(f"oracle/_ref unavailable: {e}")


# (B, H_Q, H_KV, D, D_V): test/srt/cpu/test_decode.py:145-157
REF_SHAPES = [(2, 16, 16, 64, 64), (2, 16, 1, 16, 16), (2, 32, 8, 33, 55), (2, 16, 1, 64, 64), (2, 64, 1, 13, 13),
              (2, 128, 1, 80, 80), (2, 128, 2, 512, 512), (1, 16, 1, 576, 512), (1, 16, 16, 576, 512),
              (1, 22, 1, 576, 512), (1, 40, 8, 128, 128)]


@pytest.mark.parametrize("shape", REF_SHAPES, ids=["x".join(map(str, s)) for s in REF_SHAPES])
def test_ref_cpu_kernel_matches_oracle_on_reference_shapes(shape):
    _load()
    B, Hq, Hkv, D, Dv = shape
    torch.manual_seed(1234)
    S = 1024
    q = torch.randn(B, Hq, D).to(torch.bfloat16)
    k_pool = torch.randn(B * S, Hkv, D).to(torch.bfloat16)
    v_pool = torch.randn(B * S, Hkv, Dv).to(torch.bfloat16)
    key = torch.randn(B, Hkv, D).to(torch.bfloat16)
    value = torch.randn(B, Hkv, Dv).to(torch.bfloat16)
    loc = torch.randperm(10)[:B].to(torch.int64)           # distinct slots (the reference draws randint(0, 10))
    r2t = torch.arange(B * S, dtype=torch.int32).reshape(B, S)
    rpi = torch.arange(B, dtype=torch.int64)
    sl = torch.full((B,), S, dtype=torch.int64)
    kp_ref, vp_ref = k_pool.clone(), v_pool.clone()
    decode_oracle.kv_store(kp_ref, vp_ref, loc, key.clone(), value.clone())
    ref, _ = decode_oracle.decode_attention(q, kp_ref, vp_ref, r2t, rpi, sl, D ** -0.5)
    kp, vp = k_pool.clone(), v_pool.clone()
    o = ref_cpu_kernel.decode_attention_cpu(q, kp, vp, key, value, loc, r2t, rpi, sl, D ** -0.5)
    assert torch.equal(kp, kp_ref) and torch.equal(vp, vp_ref), "the reference kernel's fused store differs"
    got = o.float().numpy()
    assert np.abs(got - ref).max() <= 3e-2
    cos = float((got * ref).sum() / (np.linalg.norm(got) * np.linalg.norm(ref)))
    assert cos > 0.99, cos


def test_ref_cpu_kernel_on_bench_shaped_workload():
    """The call bench.py makes (cpu_reference_run): paged pools, permuted slots, shared prefixes, int32
    req_to_token, out_cache_loc = each request's last slot."""
    _load()
    cfg = synth.Config("cpu_ref", 6, 0, 32, 8, 128, page_size=16, ragged=(40, 1500), share_groups=2, seed=77)
    w = synth.make_workload(cfg, "cpu")
    c = cfg
    q = w.q.reshape(c.batch, c.num_q_heads, c.head_dim)
    k = w.k.reshape(c.batch, c.num_kv_heads, c.head_dim)
    v = w.v.reshape(c.batch, c.num_kv_heads, c.head_dim)
    kp_ref, vp_ref = w.k_pool.clone(), w.v_pool.clone()
    decode_oracle.kv_store(kp_ref, vp_ref, w.out_cache_loc, k.clone(), v.clone())
    ref, _ = decode_oracle.decode_attention(q, kp_ref, vp_ref, w.req_to_token, w.req_pool_indices, w.seq_lens, w.sm_scale)
    o = ref_cpu_kernel.decode_attention_cpu(q, w.k_pool, w.v_pool, k, v, w.out_cache_loc, w.req_to_token,
                                            w.req_pool_indices, w.seq_lens, w.sm_scale)
    assert torch.equal(w.k_pool, kp_ref) and torch.equal(w.v_pool, vp_ref)
    got = o.float().numpy()
    assert np.abs(got - ref).max() <= 3e-2
    num = np.sqrt(((got - ref) ** 2).sum(-1))
    den = np.sqrt((ref ** 2).sum(-1))
    assert (num <= 3e-2 * den + 1e-6).all(), float((num / den).max())
