"""b200attn_decode_static_ws (include/b200attn.h): the decode call whose split-KV workspace is a device array
inside the library, for callers that may not allocate.  It has no counterpart in the reference, so the check is
against the existing entry point, b200attn_decode, on the same plan and inputs -- through the C-ABI, on two calls
in a row (the array keeps the first call's partials), with the tolerance of the parity tests and, because both run
the same kernel over the same ranges, bit for bit.

Collected last on purpose; passes alone in a fresh process and after the others.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest
import torch

import synth
from tests.harness import assert_close_out

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _plan(capi, w, cfg, plan_buf, kv_code):
    a = capi.PlanArgs()
    a.req_to_token = w.req_to_token.data_ptr()
    a.req_to_token_stride = w.req_to_token.stride(0)
    a.req_pool_indices = w.req_pool_indices.data_ptr()
    a.seq_lens = w.seq_lens.data_ptr()
    a.index_dtype = capi.I64
    a.batch = cfg.batch
    a.num_q_heads, a.num_kv_heads, a.head_dim, a.kv_dtype = cfg.num_q_heads, cfg.num_kv_heads, cfg.head_dim, kv_code
    a.plan = plan_buf.data_ptr()
    a.plan_bytes = plan_buf.numel()
    capi.check("b200attn_plan", capi.lib().b200attn_plan(C.byref(a), _stream()))


def _decode_args(capi, w, cfg, q3, o, lse, plan_buf, kv_code):
    a = capi.DecodeArgs()
    a.q = q3.data_ptr()
    a.q_stride_batch, a.q_stride_head = q3.stride(0), q3.stride(1)
    a.k_pool, a.v_pool = w.k_pool.data_ptr(), w.v_pool.data_ptr()
    a.k_pool_stride_slot, a.v_pool_stride_slot = w.k_pool.stride(0), w.v_pool.stride(0)
    a.o = o.data_ptr()
    a.o_stride_batch, a.o_stride_head = o.stride(0), o.stride(1)
    a.lse = lse.data_ptr()
    a.plan = plan_buf.data_ptr()
    a.indices = w.req_to_token.data_ptr()
    a.ws_o, a.ws_bytes = None, 0
    a.batch, a.num_q_heads, a.num_kv_heads = cfg.batch, cfg.num_q_heads, cfg.num_kv_heads
    a.head_dim = a.v_head_dim = cfg.head_dim
    a.q_dtype = capi.BF16
    a.kv_dtype = kv_code
    a.sm_scale, a.logit_cap = w.sm_scale, 0.0
    a.k_scale = a.v_scale = 1.0
    a.phases = 0
    a.slot_map = None
    a.pool_slots = w.k_pool.shape[0]
    return a


# (kv dtype, batch, q heads, kv heads, TP degree): one slice per CTA, token slices of a small batch, an fp8 pool,
# and a one-head-per-CTA shard whose requests are cut across many ranges (the workspace-heavy case)
CASES = [("bf16", 70, 32, 8, 1), ("bf16", 37, 32, 8, 1), ("fp8_e4m3", 24, 32, 8, 1), ("bf16", 9, 32, 8, 8)]


@pytest.mark.parametrize("kv,batch,hq,hkv,tp", CASES)
def test_decode_static_ws_matches_decode(kv, batch, hq, hkv, tp):
    from b200attn import capi

    lib = capi.lib()
    cfg = synth.shard(synth.Config("noalloc", batch, 0, hq, hkv, 128, page_size=1, ragged=(1, 5000), seed=77,
                                   kv_dtype=kv), tp)
    w = synth.make_workload(cfg, DEV)
    B, Hq, D = cfg.batch, cfg.num_q_heads, cfg.head_dim
    kv_code = capi.FP8_E4M3 if kv.startswith("fp8") else capi.BF16
    need = lib.b200attn_workspace_bytes(B, Hq, cfg.num_kv_heads, D, D, kv_code)
    assert 0 < need <= capi.STATIC_WS_BYTES

    # everything the two entry points need is allocated up front, by the caller
    plan_buf = torch.zeros(lib.b200attn_plan_bytes(B), dtype=torch.uint8, device=DEV)
    workspace = torch.empty(need // 4, dtype=torch.float32, device=DEV)
    qs = [w.q.view(B, Hq, D).clone(), (1.7 * torch.randn_like(w.q)).view(B, Hq, D)]
    outs = [[torch.full((B, Hq, D), float("nan"), dtype=w.q.dtype, device=DEV) for _ in range(2)] for _ in range(2)]
    lses = [[torch.full((B * Hq,), float("nan"), dtype=torch.float32, device=DEV) for _ in range(2)] for _ in range(2)]
    _plan(capi, w, cfg, plan_buf, kv_code)
    torch.cuda.synchronize()

    for call in range(2):   # two calls in a row: the second finds the first one's partials in the array
        q3 = qs[call]
        a_ref = _decode_args(capi, w, cfg, q3, outs[call][0], lses[call][0], plan_buf, kv_code)
        a_ref.ws_o, a_ref.ws_bytes = workspace.data_ptr(), workspace.numel() * 4
        a_new = _decode_args(capi, w, cfg, q3, outs[call][1], lses[call][1], plan_buf, kv_code)
        with torch.cuda.device(DEV):
            capi.check("b200attn_decode", lib.b200attn_decode(C.byref(a_ref), _stream()))
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
            launches = capi.launch_count()
            capi.check("b200attn_decode_static_ws", lib.b200attn_decode_static_ws(C.byref(a_new), _stream()))
            torch.cuda.synchronize()
            if call == 1:   # (the first call may load the module that holds the array)
                assert torch.cuda.mem_get_info()[0] == free_before, "the second call changed the free device memory"
        assert capi.launch_count() == launches + 1
        o_ref, o_new = outs[call][0].float().cpu().numpy(), outs[call][1].float().cpu().numpy()
        assert np.isfinite(o_new).all()
        assert_close_out(o_new, o_ref, what=f"call {call}: decode_static_ws vs decode")
        assert np.abs(lses[call][1].cpu().numpy() - lses[call][0].cpu().numpy()).max() <= 1e-3
        # same kernel, same plan, same ranges: the two agree bit for bit
        assert torch.equal(outs[call][1], outs[call][0]), f"call {call}: outputs differ"
        assert torch.equal(lses[call][1], lses[call][0]), f"call {call}: lse differs"
    assert not torch.equal(outs[0][1], outs[1][1])   # the two calls really had different inputs


def test_decode_static_ws_rejects():
    """Sizes beyond the static array fail with their own code before anything is launched; a caller workspace is
    refused (the library owns it in this entry point)."""
    from b200attn import capi

    lib = capi.lib()
    a = capi.DecodeArgs()
    a.batch, a.num_q_heads, a.num_kv_heads, a.head_dim, a.v_head_dim = 4096, 64, 8, 128, 128
    a.q_dtype = a.kv_dtype = capi.BF16
    assert lib.b200attn_workspace_bytes(4096, 64, 8, 128, 128, capi.BF16) > capi.STATIC_WS_BYTES
    launches = capi.launch_count()
    assert lib.b200attn_decode_static_ws(C.byref(a), None) == capi.E_STATIC_LIMIT
    assert "static workspace" in capi.strerror(capi.E_STATIC_LIMIT)
    a.batch = 8
    buf = torch.empty(16, dtype=torch.float32, device=DEV)
    a.ws_o, a.ws_bytes = buf.data_ptr(), 64
    assert lib.b200attn_decode_static_ws(C.byref(a), None) == capi.E_INVALID
    a.ws_o, a.ws_bytes = None, 0
    a.head_dim = a.v_head_dim = 72   # no kernel for this head dim
    assert lib.b200attn_decode_static_ws(C.byref(a), None) == capi.E_UNSUPPORTED
    assert capi.launch_count() == launches
