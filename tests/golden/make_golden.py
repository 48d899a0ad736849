"""Generates tests/golden/*.npz by running the REFERENCE's own code on CPU.

Run in the build container only (needs the read-only reference tree):

    python tests/golden/make_golden.py

For every case the inputs are rebuilt deterministically from ``synth.make_workload`` (a
sha256 of the input bytes is stored so a drifted RNG is detected, not silently accepted) and
the outputs come from the reference:

* pool rows after ``MHATokenToKVPool.set_kv_buffer``   (mem_cache/memory_pool.py:369-407)
* attention output of ``RadixAttention`` -> ``TorchNativeAttnBackend.forward_decode``
  (layers/radix_attention.py:84-110, layers/attention/torch_native_backend.py:226-267)
* kv_indptr / kv_indices by the Python gather the reference's own test treats as ground
  truth (test/srt/test_create_kvindices.py:43-49)
* extend attention of ``RadixAttention`` -> ``TorchNativeAttnBackend.forward_extend``
  (layers/attention/torch_native_backend.py:27-110, 182-224) on EXTEND batches with cached prefixes
* 2-way state merge by ``merge_state_torch`` (sgl-kernel/tests/test_merge_state_v2.py:101-141),
  extracted from the test file with ``ast`` because the file imports the CUDA-only sgl_kernel.
"""
from __future__ import annotations

import ast
import hashlib
import os
import sys
import types
from typing import Optional

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import synth  # noqa: E402
from oracle import ref_loader  # noqa: E402

CASES = {
    "g1_gqa4_d64_p1": synth.Config("g1", 5, 0, 8, 2, 64, ragged=(3, 200), seed=11),
    "g2_gqa4_d128_p16_shared": synth.Config("g2", 4, 0, 32, 8, 128, page_size=16, ragged=(17, 300),
                                            share_groups=2, seed=12),
    "g3_mha_d128": synth.Config("g3", 3, 0, 4, 4, 128, ragged=(1, 130), seed=13),
    "g4_gqa8_d128_p4": synth.Config("g4", 3, 0, 16, 2, 128, page_size=4, ragged=(30, 260), seed=14),
    "g5_f16_gqa4_d64": synth.Config("g5", 4, 0, 8, 2, 64, kv_dtype="f16", q_dtype="f16",
                                    ragged=(5, 150), seed=15),
    "g7_fp8_gqa4_d128": synth.Config("g7", 4, 0, 32, 8, 128, kv_dtype="fp8_e4m3", ragged=(9, 220),
                                     seed=17),
}

# |x / scale| edge values for the fp8 cast
KAT = [447.0, 448.0, 456.0, 464.0, 465.0, 472.0, 480.0, 500.0, 1e4, float("inf"), float("nan"),
       2.0 ** -9, 2.0 ** -10, 0.51 * 2.0 ** -10, 0.0, -0.0, 1.0, -1.0, 0.3, -0.07, 3.1415, 17.0,
       2.0 ** -6, 2.0 ** -7, 1.5 * 2.0 ** -9, 2.5 * 2.0 ** -9, 240.0, 416.0, 432.0]


def bits(t: torch.Tensor) -> np.ndarray:
    if t.dtype in (torch.bfloat16, torch.float16):
        return t.contiguous().view(torch.int16).numpy().view(np.uint16).copy()
    if t.dtype in (torch.uint8, torch.float8_e4m3fn):
        return t.contiguous().view(torch.uint8).numpy().copy()
    return t.contiguous().numpy().copy()


def digest(w: synth.Workload) -> str:
    h = hashlib.sha256()
    for t in (w.qkv, w.k_pool, w.v_pool, w.req_to_token, w.req_pool_indices, w.seq_lens, w.out_cache_loc):
        h.update(bits(t).tobytes())
    return h.hexdigest()


def run_reference_decode(ref, w: synth.Workload, k_pool: torch.Tensor, v_pool: torch.Tensor,
                         pool_dtype: torch.dtype, save_kv_cache: bool = True):
    c = w.cfg
    pool = ref.MHATokenToKVPool(size=w.pool_size, page_size=c.page_size, dtype=pool_dtype,
                                head_num=c.num_kv_heads, head_dim=c.head_dim, layer_num=1,
                                device="cpu", enable_memory_saver=False)
    assert pool.k_buffer[0].shape == k_pool.shape, (pool.k_buffer[0].shape, k_pool.shape)
    pool.k_buffer[0].copy_(k_pool)
    pool.v_buffer[0].copy_(v_pool)
    r2t = ref.ReqToTokenPool(size=w.req_to_token.shape[0], max_context_len=w.req_to_token.shape[1],
                             device="cpu", enable_memory_saver=False)
    r2t.req_to_token.copy_(w.req_to_token)
    backend = ref.TorchNativeAttnBackend(types.SimpleNamespace(device="cpu"))
    fb = ref.ForwardBatch(forward_mode=ref.ForwardMode.DECODE, batch_size=c.batch,
                          input_ids=torch.zeros(c.batch, dtype=torch.int64),
                          req_pool_indices=w.req_pool_indices, seq_lens=w.seq_lens,
                          out_cache_loc=w.out_cache_loc, seq_lens_sum=w.seq_lens_sum,
                          req_to_token_pool=r2t, token_to_kv_pool=pool, attn_backend=backend)
    layer = ref.RadixAttention(c.num_q_heads, c.head_dim, w.sm_scale, c.num_kv_heads, layer_id=0)
    qkv = w.qkv.clone()
    q, k, v = qkv.split([c.num_q_heads * c.head_dim, c.num_kv_heads * c.head_dim,
                         c.num_kv_heads * c.head_dim], dim=-1)
    out = layer(q, k, v, fb, save_kv_cache=save_kv_cache)
    return out, pool


def gen_decode_cases(ref):
    for name, cfg in CASES.items():
        w = synth.make_workload(cfg, "cpu")
        c = w.cfg
        if c.kv_dtype == "fp8_e4m3":
            # The reference's torch_native backend cannot run SDPA on fp8 tensors and ignores
            # k/v scales (torch_native_backend.py:244-265): pin the fp8 DECODE math on an exactly
            # dequantised bf16 copy of the pool (e4m3 is a subset of bf16), scales = 1, without
            # storing the new token (its fp8 store is pinned separately by the g6 cases).
            kd = w.k_pool.view(torch.float8_e4m3fn).to(torch.bfloat16)
            vd = w.v_pool.view(torch.float8_e4m3fn).to(torch.bfloat16)
            out, pool = run_reference_decode(ref, w, kd, vd, torch.bfloat16, save_kv_cache=False)
        else:
            dt = w.k_pool.dtype
            out, pool = run_reference_decode(ref, w, w.k_pool, w.v_pool, dt)
        loc = w.out_cache_loc
        indptr = np.zeros(c.batch + 1, dtype=np.int32)
        indptr[1:] = torch.cumsum(w.seq_lens, dim=0).numpy().astype(np.int32)
        rpi = w.req_pool_indices.numpy()
        sl = w.seq_lens.numpy()
        indices = torch.cat([w.req_to_token[rpi[i], : sl[i]] for i in range(c.batch)], dim=0).contiguous()
        np.savez(os.path.join(HERE, name + ".npz"),
                 digest=np.frombuffer(digest(w).encode(), dtype=np.uint8),
                 out_bits=bits(out.view(c.batch, c.num_q_heads, c.head_dim)),
                 k_rows_after=bits(pool.k_buffer[0][loc]),
                 v_rows_after=bits(pool.v_buffer[0][loc]),
                 kv_indptr=indptr, kv_indices=indices.numpy().astype(np.int32))
        print(name, "ok", out.shape)


def gen_fp8_store(ref):
    """set_kv_buffer into an fp8-e4m3 pool: KAT edge values and random data, with no scale,
    scale 1.0 and a non-trivial scale pair.  CPU reference => true division semantics."""
    Hkv, D = 2, 64
    for tag, (ks, vs) in {"none": (None, None), "one": (1.0, 1.0), "half_two": (0.5, 2.0),
                          "odd": (0.3, 1.7)}.items():
        for src_dtype, dname in ((torch.bfloat16, "bf16"), (torch.float16, "f16")):
            g = torch.Generator().manual_seed(99)
            n = 24
            k = (torch.randn((n, Hkv, D), generator=g) * 3).to(src_dtype)
            v = (torch.randn((n, Hkv, D), generator=g) * 40).to(src_dtype)
            kat = torch.tensor(KAT + [-x for x in KAT], dtype=torch.float32)
            # place edge values so that x / scale hits them (up to source-dtype rounding)
            k.view(-1)[: kat.numel()] = (kat * (ks or 1.0)).to(src_dtype)
            v.view(-1)[: kat.numel()] = (kat * (vs or 1.0)).to(src_dtype)
            loc = torch.randperm(40, generator=g)[:n].to(torch.int64) + 1
            pool = ref.MHATokenToKVPool(size=48, page_size=1, dtype=torch.float8_e4m3fn, head_num=Hkv,
                                        head_dim=D, layer_num=1, device="cpu", enable_memory_saver=False)
            k_in, v_in = k.clone(), v.clone()
            layer = types.SimpleNamespace(layer_id=0)
            pool.set_kv_buffer(layer, loc, k, v, ks, vs)
            np.savez(os.path.join(HERE, f"g6_fp8_store_{dname}_{tag}.npz"),
                     k_in=bits(k_in), v_in=bits(v_in), loc=loc.numpy(),
                     k_scale=np.float64(-1.0 if ks is None else ks),
                     v_scale=np.float64(-1.0 if vs is None else vs),
                     k_after=bits(k), v_after=bits(v),  # div_ mutates the inputs in place
                     k_pool=bits(pool.k_buffer[0]), v_pool=bits(pool.v_buffer[0]))
    # the whole bf16 -> e4m3fn cast table, from torch itself
    allbf = torch.arange(65536, dtype=torch.int32).to(torch.int16).view(torch.bfloat16)
    np.savez(os.path.join(HERE, "g6_cast_table_bf16_e4m3fn.npz"),
             table=allbf.to(torch.float8_e4m3fn).view(torch.uint8).numpy())
    print("fp8 store ok")


def load_merge_state_torch():
    path = os.path.join(ref_loader.REFERENCE_ROOT, "sgl-kernel", "tests", "test_merge_state_v2.py")
    tree = ast.parse(open(path).read())
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "merge_state_torch")
    ns = {"torch": torch, "Optional": Optional}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), path, "exec"), ns)
    return ns["merge_state_torch"]


def gen_merge():
    f = load_merge_state_torch()
    g = torch.Generator().manual_seed(7)
    T, H, D = 37, 8, 128
    for dt, dname in ((torch.bfloat16, "bf16"), (torch.float16, "f16")):
        la = torch.randn((T, H), generator=g)
        lb = torch.randn((T, H), generator=g)
        ma = torch.rand((T, H), generator=g) < 0.1
        mb = (torch.rand((T, H), generator=g) < 0.1) & ~ma
        la[ma] = float("inf")
        lb[mb] = float("inf")
        oa = torch.randn((T, H, D), generator=g).to(dt)
        ob = torch.randn((T, H, D), generator=g).to(dt)
        out, lse = f(oa.clone(), la.clone(), ob.clone(), lb.clone())
        np.savez(os.path.join(HERE, f"g8_merge_{dname}.npz"), o_a=bits(oa), o_b=bits(ob),
                 lse_a=la.numpy(), lse_b=lb.numpy(), out=out.to(torch.float32).numpy(),
                 lse=lse.numpy())
    print("merge ok")


# (head_size, rotary_dim, max_position, base, is_neox, tokens, q heads, kv heads): the shape list of
# the reference's own RoPE test (sgl-kernel/tests/test_rotary_embedding.py:146-151), fewer tokens.
ROPE_CASES = {
    "g9_rope_neox_d64": (64, 64, 32, 8000, True, 24, 1, 1),
    "g9_rope_neox_d256_r128": (256, 128, 4096, 10000, True, 12, 4, 2),
    "g9_rope_gptj_d128": (128, 128, 2048, 10000, False, 10, 16, 4),
    "g9_rope_gptj_d512_r128": (512, 128, 311, 10000, False, 9, 4, 2),
    "g9_rope_neox_d128_llama": (128, 128, 8192, 500000, True, 11, 32, 8),
}


def gen_rope():
    """``RotaryEmbedding.forward_native`` (layers/rotary_embedding.py:138-166) of the reference, on
    fp32 inputs (exact fp32 arithmetic on the fp32 cache: pins the oracle) and on bf16 inputs (the
    reference's bf16 arithmetic: the 1e-2 comparison its own test makes)."""
    ref_loader.load()
    sys.path.insert(0, os.path.join(ref_loader.REFERENCE_ROOT, "python"))
    from sglang.srt.layers.rotary_embedding import RotaryEmbedding
    sys.path.pop(0)
    for name, (hs, rot, maxpos, base, neox, n, hq, hkv) in ROPE_CASES.items():
        r = RotaryEmbedding.__new__(RotaryEmbedding)   # __init__ probes platform-specific kernels
        torch.nn.Module.__init__(r)
        r.head_size, r.rotary_dim, r.max_position_embeddings, r.base = hs, rot, maxpos, base
        r.is_neox_style, r.dtype = neox, torch.float32
        r.register_buffer("cos_sin_cache", r._compute_cos_sin_cache(), persistent=False)
        g = torch.Generator().manual_seed(hs + rot + n)
        pos = torch.randint(0, maxpos, (n,), generator=g, dtype=torch.int64)
        q = torch.randn((n, hq * hs), generator=g).to(torch.bfloat16)
        k = torch.randn((n, hkv * hs), generator=g).to(torch.bfloat16)
        q32, k32 = r.forward_native(pos, q.float(), k.float())
        q16, k16 = r.forward_native(pos, q.clone(), k.clone())
        np.savez_compressed(os.path.join(HERE, name + ".npz"), positions=pos.numpy(), q_in=bits(q), k_in=bits(k),
                 cache_rows=r.cos_sin_cache[pos].numpy(), q_out_f32=q32.numpy(), k_out_f32=k32.numpy(),
                 q_out_bf16=bits(q16), k_out_bf16=bits(k16),
                 meta=np.array([hs, rot, maxpos, base, int(neox), n, hq, hkv], dtype=np.int64))
    print("rope ok")


# name -> (Config, prefix_lens, extend_lens, causal)
EXTEND_CASES = {
    "g10_extend_gqa4_d128": (synth.Config("g10a", 3, 0, 8, 2, 128, seed=21), (0, 37, 130), (70, 33, 1), True),
    "g10_extend_gqa8_d64_p16": (synth.Config("g10b", 2, 0, 16, 2, 64, page_size=16, seed=22), (16, 100), (50, 129), True),
    "g10_extend_mha_d128_f16": (synth.Config("g10c", 2, 0, 4, 4, 128, kv_dtype="f16", q_dtype="f16", seed=23),
                                (5, 0), (40, 150), True),
    "g10_extend_gqa4_d128_noncausal": (synth.Config("g10d", 2, 0, 8, 2, 128, seed=24), (20, 0), (45, 64), False),
}


def digest_extend(w) -> str:
    h = hashlib.sha256()
    for t in (w.qkv, w.k_pool, w.v_pool, w.req_to_token, w.req_pool_indices, w.seq_lens, w.out_cache_loc):
        h.update(bits(t).tobytes())
    return h.hexdigest()


def gen_extend_cases(ref):
    sys.path.insert(0, os.path.join(ref_loader.REFERENCE_ROOT, "python"))
    from sglang.srt.layers.radix_attention import AttentionType
    sys.path.pop(0)
    for name, (cfg, pl, el, causal) in EXTEND_CASES.items():
        w = synth.make_extend_workload(cfg, pl, el, "cpu")
        c = w.cfg
        pool = ref.MHATokenToKVPool(size=w.pool_size, page_size=c.page_size, dtype=w.k_pool.dtype,
                                    head_num=c.num_kv_heads, head_dim=c.head_dim, layer_num=1,
                                    device="cpu", enable_memory_saver=False)
        pool.k_buffer[0].copy_(w.k_pool)
        pool.v_buffer[0].copy_(w.v_pool)
        r2t = ref.ReqToTokenPool(size=w.req_to_token.shape[0], max_context_len=w.req_to_token.shape[1],
                                 device="cpu", enable_memory_saver=False)
        r2t.req_to_token.copy_(w.req_to_token)
        backend = ref.TorchNativeAttnBackend(types.SimpleNamespace(device="cpu"))
        B = len(pl)
        fb = ref.ForwardBatch(forward_mode=ref.ForwardMode.EXTEND, batch_size=B,
                              input_ids=torch.zeros(w.tokens, dtype=torch.int64),
                              req_pool_indices=w.req_pool_indices, seq_lens=w.seq_lens,
                              out_cache_loc=w.out_cache_loc, seq_lens_sum=int(w.seq_lens.sum()),
                              req_to_token_pool=r2t, token_to_kv_pool=pool, attn_backend=backend,
                              extend_prefix_lens=w.extend_prefix_lens, extend_seq_lens=w.extend_seq_lens)
        layer = ref.RadixAttention(c.num_q_heads, c.head_dim, w.sm_scale, c.num_kv_heads, layer_id=0,
                                   attn_type=AttentionType.DECODER if causal else AttentionType.ENCODER_ONLY)
        qkv = w.qkv.clone()
        q, k, v = qkv.split([c.num_q_heads * c.head_dim, c.num_kv_heads * c.head_dim,
                             c.num_kv_heads * c.head_dim], dim=-1)
        out = layer(q, k, v, fb, save_kv_cache=True)
        loc = w.out_cache_loc
        np.savez_compressed(os.path.join(HERE, name + ".npz"),
                            digest=np.frombuffer(digest_extend(w).encode(), dtype=np.uint8),
                            prefix_lens=np.asarray(pl, dtype=np.int64), extend_lens=np.asarray(el, dtype=np.int64),
                            causal=np.int64(causal),
                            out_bits=bits(out.view(w.tokens, c.num_q_heads, c.head_dim)),
                            k_rows_after=bits(pool.k_buffer[0][loc]), v_rows_after=bits(pool.v_buffer[0][loc]))
        print(name, "ok", out.shape)


if __name__ == "__main__":
    torch.manual_seed(0)
    only = sys.argv[1] if len(sys.argv) > 1 else "all"
    ref = ref_loader.load()
    if only in ("all", "decode"):
        gen_decode_cases(ref)
    if only in ("all", "fp8"):
        gen_fp8_store(ref)
    if only in ("all", "merge"):
        gen_merge()
    if only in ("all", "rope"):
        gen_rope()
    if only in ("all", "extend"):
        gen_extend_cases(ref)
