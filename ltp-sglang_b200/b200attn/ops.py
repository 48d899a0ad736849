"""torch custom ops ``torch.ops.b200attn.*`` over the C-ABI (``capi``).

The schemas follow the sgl-kernel convention of explicit out-parameters marked
mutable (reference: sgl-kernel/csrc/common_extension.cc:36-39, 56-63).  The ops
are registered for CUDA only; calling them on CPU tensors raises -- there is no
fallback path.  Every op enqueues on ``torch.cuda.current_stream()`` and is
CUDA-graph capture safe (no sync, no allocation).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import capi

_DT = {
    torch.bfloat16: capi.BF16,
    torch.float16: capi.F16,
    torch.float8_e4m3fn: capi.FP8_E4M3,
    torch.uint8: capi.FP8_E4M3,  # fp8 pools are stored as uint8 (memory_pool.py:114-118); an e5m2 pool must be
    #                              passed viewed as torch.float8_e5m2 (the backend does)
    torch.float8_e5m2: capi.FP8_E5M2,
}
_IDX = {torch.int32: capi.I32, torch.int64: capi.I64}


def dtype_code(dt: torch.dtype) -> int:
    try:
        return _DT[dt]
    except KeyError:
        raise TypeError(f"b200attn: unsupported dtype {dt}") from None


def _cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "b200attn ops run on CUDA tensors only (sm_100a kernels; there is no CPU path)"
            )


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _stream(t: torch.Tensor):
    return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


# --------------------------------------------------------------------------- #
# implementations (ctypes -> C-ABI)
# --------------------------------------------------------------------------- #
def _kv_store(k, v, loc, k_pool, v_pool, k_scale, v_scale, k_scale_t, v_scale_t, scale_mode):
    _cuda(k, v, loc, k_pool, v_pool, k_scale_t, v_scale_t)
    if k.dim() != 3 or v.dim() != 3 or k_pool.dim() != 3 or v_pool.dim() != 3:
        raise ValueError("kv_store: k, v, k_pool, v_pool must be [tokens|slots, heads, dim]")
    if k.stride(2) != 1 or v.stride(2) != 1 or not k_pool.is_contiguous() or not v_pool.is_contiguous():
        raise ValueError("kv_store: last dim must be contiguous; pools must be contiguous")
    if loc.dtype != torch.int64 or not loc.is_contiguous():
        raise TypeError("kv_store: loc must be a contiguous int64 tensor (out_cache_loc)")
    if loc.numel() != k.shape[0] or v.shape[0] != k.shape[0]:
        raise ValueError("kv_store: loc, k, v disagree on the number of tokens")
    if k.shape[1] != k_pool.shape[1] or k.shape[2] != k_pool.shape[2] or v.shape[2] != v_pool.shape[2]:
        raise ValueError("kv_store: k/v shape does not match the pool")
    a = capi.KvStoreArgs()
    _fill_store_args(a, k, v, loc, k_pool, v_pool, k_scale, v_scale, k_scale_t, v_scale_t, scale_mode)
    with torch.cuda.device(k.device):
        capi.check("b200attn_kv_store", capi.lib().b200attn_kv_store(C.byref(a), _stream(k)))


def _fill_store_args(a, k, v, loc, k_pool, v_pool, k_scale, v_scale, k_scale_t, v_scale_t, scale_mode):
    a.k, a.v = k.data_ptr(), v.data_ptr()
    a.k_stride_token, a.k_stride_head = k.stride(0), k.stride(1)
    a.v_stride_token, a.v_stride_head = v.stride(0), v.stride(1)
    a.loc = loc.data_ptr()
    a.k_pool, a.v_pool = k_pool.data_ptr(), v_pool.data_ptr()
    a.k_pool_stride_slot, a.v_pool_stride_slot = k_pool.stride(0), v_pool.stride(0)
    a.pool_slots = k_pool.shape[0]
    a.n_tokens, a.num_kv_heads = k.shape[0], k.shape[1]
    a.head_dim, a.v_head_dim = k.shape[2], v.shape[2]
    a.src_dtype = dtype_code(k.dtype)
    a.pool_dtype = dtype_code(k_pool.dtype)
    a.scale_mode = scale_mode
    a.mutate_src = 1
    a.k_scale = 1.0 if k_scale is None else k_scale
    a.v_scale = 1.0 if v_scale is None else v_scale
    a.k_scale_ptr, a.v_scale_ptr = _ptr(k_scale_t), _ptr(v_scale_t)


def _rope_kv_store(q, k, v, positions, cos_sin_cache, is_neox, loc, k_pool, v_pool, k_scale, v_scale,
                   k_scale_t, v_scale_t, scale_mode):
    """RoPE on q and k in place (+ the KV store when ``loc`` / pools are given)."""
    save = loc is not None
    _cuda(q, k, v, positions, cos_sin_cache, loc, k_pool, v_pool, k_scale_t, v_scale_t)
    if q.dim() != 3 or k.dim() != 3 or q.stride(2) != 1 or k.stride(2) != 1:
        raise ValueError("rope_kv_store: q / k must be [tokens, heads, dim] with a contiguous last dim")
    if q.shape[0] != k.shape[0] or q.shape[2] != k.shape[2] or q.dtype != k.dtype:
        raise ValueError("rope_kv_store: q and k disagree on tokens, head_dim or dtype")
    if cos_sin_cache.dtype != torch.float32 or cos_sin_cache.dim() != 2 or cos_sin_cache.stride(1) != 1:
        raise ValueError("rope_kv_store: cos_sin_cache should be float32 [max_position, rotary_dim]")
    if positions.dtype not in _IDX or not positions.is_contiguous() or positions.numel() != q.shape[0]:
        raise TypeError("rope_kv_store: positions must be contiguous int32/int64 [tokens]")
    a = capi.RopeKvStoreArgs()
    a.q = q.data_ptr()
    a.q_stride_token, a.q_stride_head = q.stride(0), q.stride(1)
    a.num_q_heads = q.shape[1]
    a.rotary_dim = cos_sin_cache.shape[1]
    a.is_neox = 1 if is_neox else 0
    a.pos_dtype = _IDX[positions.dtype]
    a.positions = positions.data_ptr()
    a.cos_sin_cache = cos_sin_cache.data_ptr()
    a.cache_stride = cos_sin_cache.stride(0)
    a.max_position = cos_sin_cache.shape[0]
    a.save_kv = 1 if save else 0
    if save:
        if v is None or k_pool is None or v_pool is None:
            raise ValueError("rope_kv_store: v, k_pool and v_pool are needed to store")
        if v.dim() != 3 or v.stride(2) != 1 or v.shape != k.shape:
            raise ValueError("rope_kv_store: v must match k ([tokens, kv_heads, dim])")
        if loc.dtype != torch.int64 or not loc.is_contiguous() or loc.numel() != k.shape[0]:
            raise TypeError("rope_kv_store: loc must be a contiguous int64 [tokens] tensor (out_cache_loc)")
        if not k_pool.is_contiguous() or not v_pool.is_contiguous() or k_pool.dim() != 3:
            raise ValueError("rope_kv_store: pools must be contiguous [slots, kv_heads, dim]")
        if k.shape[1] != k_pool.shape[1] or k.shape[2] != k_pool.shape[2] or v.shape[2] != v_pool.shape[2]:
            raise ValueError("rope_kv_store: k/v shape does not match the pool")
        _fill_store_args(a.kv, k, v, loc, k_pool, v_pool, k_scale, v_scale, k_scale_t, v_scale_t, scale_mode)
    else:
        a.kv.k = k.data_ptr()
        a.kv.k_stride_token, a.kv.k_stride_head = k.stride(0), k.stride(1)
        a.kv.n_tokens, a.kv.num_kv_heads = k.shape[0], k.shape[1]
        a.kv.head_dim = a.kv.v_head_dim = k.shape[2]
        a.kv.src_dtype = a.kv.pool_dtype = dtype_code(k.dtype)
    with torch.cuda.device(q.device):
        capi.check("b200attn_rope_kv_store", capi.lib().b200attn_rope_kv_store(C.byref(a), _stream(q)))


def _plan(req_to_token, req_pool_indices, seq_lens, plan_buf, kv_indptr, kv_indices,
          num_q_heads, num_kv_heads, head_dim, kv_dtype, max_ctas, min_chunks, sliding_window=0):
    _cuda(req_to_token, req_pool_indices, seq_lens, plan_buf, kv_indptr, kv_indices)
    if req_to_token.dtype != torch.int32 or req_to_token.dim() != 2 or req_to_token.stride(1) != 1:
        raise TypeError("plan: req_to_token must be int32 [max_reqs, max_context_len]")
    if req_pool_indices.dtype != seq_lens.dtype or seq_lens.dtype not in _IDX:
        raise TypeError("plan: req_pool_indices and seq_lens must both be int32 or both int64")
    if not (req_pool_indices.is_contiguous() and seq_lens.is_contiguous()):
        raise ValueError("plan: req_pool_indices / seq_lens must be contiguous")
    batch = seq_lens.numel()
    if req_pool_indices.numel() != batch:
        raise ValueError("plan: req_pool_indices and seq_lens differ in length")
    for name, t in (("kv_indptr", kv_indptr), ("kv_indices", kv_indices)):
        if t is not None and (t.dtype != torch.int32 or not t.is_contiguous()):
            raise TypeError(f"plan: {name} must be contiguous int32")
    if kv_indptr is not None and kv_indptr.numel() < batch + 1:
        raise ValueError("plan: kv_indptr needs batch + 1 entries")
    a = capi.PlanArgs()
    a.req_to_token = req_to_token.data_ptr()
    a.req_to_token_stride = req_to_token.stride(0)
    a.req_pool_indices = req_pool_indices.data_ptr()
    a.seq_lens = seq_lens.data_ptr()
    a.index_dtype = _IDX[seq_lens.dtype]
    a.batch = batch
    a.num_q_heads, a.num_kv_heads, a.head_dim, a.kv_dtype = num_q_heads, num_kv_heads, head_dim, kv_dtype
    a.csr_indptr = None
    a.kv_indptr, a.kv_indices = _ptr(kv_indptr), _ptr(kv_indices)
    a.plan = plan_buf.data_ptr()
    a.plan_bytes = plan_buf.numel() * plan_buf.element_size()
    a.max_ctas, a.min_chunks_per_cta = max_ctas, min_chunks
    a.sliding_window = sliding_window
    with torch.cuda.device(plan_buf.device):
        capi.check("b200attn_plan", capi.lib().b200attn_plan(C.byref(a), _stream(plan_buf)))


def _plan_csr(csr_indptr, plan_buf, num_q_heads, num_kv_heads, head_dim, kv_dtype, max_ctas, min_chunks):
    _cuda(csr_indptr, plan_buf)
    if csr_indptr.dtype != torch.int32 or not csr_indptr.is_contiguous() or csr_indptr.numel() < 1:
        raise TypeError("plan_csr: kv_indptr must be contiguous int32 [batch + 1]")
    a = capi.PlanArgs()
    a.index_dtype = capi.I32
    a.batch = csr_indptr.numel() - 1
    a.num_q_heads, a.num_kv_heads, a.head_dim, a.kv_dtype = num_q_heads, num_kv_heads, head_dim, kv_dtype
    a.csr_indptr = csr_indptr.data_ptr()
    a.plan = plan_buf.data_ptr()
    a.plan_bytes = plan_buf.numel() * plan_buf.element_size()
    a.max_ctas, a.min_chunks_per_cta = max_ctas, min_chunks
    with torch.cuda.device(plan_buf.device):
        capi.check("b200attn_plan", capi.lib().b200attn_plan(C.byref(a), _stream(plan_buf)))


def _decode(q, k_pool, v_pool, o, lse, plan_buf, indices, workspace, sm_scale, logit_cap, k_scale, v_scale,
            phases=0, slot_map=None):
    _cuda(q, k_pool, v_pool, o, lse, plan_buf, indices, workspace, slot_map)
    if slot_map is not None and (slot_map.dtype != torch.int64 or not slot_map.is_contiguous()):
        raise TypeError("decode: slot_map (SWAKVPool.full_to_swa_index_mapping) must be a contiguous int64 tensor")
    if q.dim() != 3 or o.dim() != 3 or k_pool.dim() != 3 or v_pool.dim() != 3:
        raise ValueError("decode: q/o must be [batch, heads, dim]; pools [slots, kv_heads, dim]")
    if q.stride(2) != 1 or o.stride(2) != 1 or k_pool.stride(2) != 1 or v_pool.stride(2) != 1:
        raise ValueError("decode: last dim must be contiguous")
    if k_pool.stride(1) != k_pool.shape[2] or v_pool.stride(1) != v_pool.shape[2]:
        raise ValueError("decode: kv heads of a slot must be adjacent ([slots, kv_heads, dim] layout)")
    if o.dtype != q.dtype or o.shape[0] != q.shape[0] or o.shape[1] != q.shape[1]:
        raise ValueError("decode: o must match q in dtype, batch and heads")
    if indices.dtype != torch.int32:
        raise TypeError("decode: indices (req_to_token or kv_indices) must be int32")
    if lse is not None and (lse.dtype != torch.float32 or not lse.is_contiguous()
                            or lse.numel() < q.shape[0] * q.shape[1]):
        raise TypeError("decode: lse must be contiguous float32 [batch, heads]")
    if workspace.dtype != torch.float32:
        raise TypeError("decode: workspace must be float32")
    a = capi.DecodeArgs()
    a.q = q.data_ptr()
    a.q_stride_batch, a.q_stride_head = q.stride(0), q.stride(1)
    a.k_pool, a.v_pool = k_pool.data_ptr(), v_pool.data_ptr()
    a.k_pool_stride_slot, a.v_pool_stride_slot = k_pool.stride(0), v_pool.stride(0)
    a.o = o.data_ptr()
    a.o_stride_batch, a.o_stride_head = o.stride(0), o.stride(1)
    a.lse = _ptr(lse)
    a.plan = plan_buf.data_ptr()
    a.indices = indices.data_ptr()
    a.ws_o = workspace.data_ptr()
    a.ws_bytes = workspace.numel() * 4
    a.batch, a.num_q_heads = q.shape[0], q.shape[1]
    a.num_kv_heads = k_pool.shape[1]
    a.head_dim, a.v_head_dim = q.shape[2], v_pool.shape[2]
    a.q_dtype = dtype_code(q.dtype)
    a.kv_dtype = dtype_code(k_pool.dtype)
    a.sm_scale, a.logit_cap, a.k_scale, a.v_scale = sm_scale, logit_cap, k_scale, v_scale
    a.phases = phases
    a.slot_map = _ptr(slot_map)
    a.pool_slots = min(k_pool.shape[0], v_pool.shape[0])
    with torch.cuda.device(q.device):
        capi.check("b200attn_decode", capi.lib().b200attn_decode(C.byref(a), _stream(q)))


def _extend(q, o, lse, k_pool, v_pool, qo_indptr, kv_indptr, kv_indices, req_to_token, req_pool_indices,
            prefix_lens, extend_loc, custom_mask, mask_indptr, max_extend_len, causal, sliding_window,
            skip_prefix_custom_mask, sm_scale, logit_cap, k_scale=1.0, v_scale=1.0, avg_kv_len=0):
    """Extend / prefill attention over cached prefix + new tokens (``extend_attention_fwd``,
    triton_ops/extend_attention.py:306-438); the new tokens' K/V are read from the pool at
    ``extend_loc`` (``out_cache_loc``), where ``kv_store`` has put them.  An fp8-e4m3 pool is read as
    ``k_scale * K`` / ``v_scale * V`` (the per-tensor scales ``kv_store`` divided by)."""
    _cuda(q, o, lse, k_pool, v_pool, qo_indptr, kv_indptr, kv_indices, req_to_token, req_pool_indices,
          prefix_lens, extend_loc, custom_mask, mask_indptr)
    if q.dim() != 3 or o.dim() != 3 or k_pool.dim() != 3 or v_pool.dim() != 3:
        raise ValueError("extend: q/o must be [tokens, heads, dim]; pools [slots, kv_heads, dim]")
    if q.stride(2) != 1 or o.stride(2) != 1 or k_pool.stride(2) != 1 or v_pool.stride(2) != 1:
        raise ValueError("extend: last dim must be contiguous")
    if k_pool.stride(1) != k_pool.shape[2] or v_pool.stride(1) != v_pool.shape[2]:
        raise ValueError("extend: kv heads of a slot must be adjacent ([slots, kv_heads, dim] layout)")
    if o.dtype != q.dtype or o.shape[0] != q.shape[0] or o.shape[1] != q.shape[1]:
        raise ValueError("extend: o must match q in dtype, tokens and heads")
    for name, t in (("qo_indptr", qo_indptr), ("kv_indptr", kv_indptr), ("kv_indices", kv_indices)):
        if t is not None and (t.dtype != torch.int32 or not t.is_contiguous()):
            raise TypeError(f"extend: {name} must be contiguous int32")
    batch = qo_indptr.numel() - 1
    if kv_indptr.numel() < batch + 1:
        raise ValueError("extend: kv_indptr needs batch + 1 entries")
    if extend_loc.dtype != torch.int64 or not extend_loc.is_contiguous() or extend_loc.numel() < q.shape[0]:
        raise TypeError("extend: extend_loc must be contiguous int64 [tokens] (out_cache_loc)")
    a = capi.ExtendArgs()
    if kv_indices is None:
        if req_to_token is None or req_pool_indices is None or prefix_lens is None:
            raise ValueError("extend: give kv_indices, or req_to_token + req_pool_indices + prefix_lens")
        if req_to_token.dtype != torch.int32 or req_to_token.dim() != 2 or req_to_token.stride(1) != 1:
            raise TypeError("extend: req_to_token must be int32 [max_reqs, max_context_len]")
        if req_pool_indices.dtype != prefix_lens.dtype or prefix_lens.dtype not in _IDX:
            raise TypeError("extend: req_pool_indices and prefix_lens must both be int32 or both int64")
        if not (req_pool_indices.is_contiguous() and prefix_lens.is_contiguous()):
            raise ValueError("extend: req_pool_indices / prefix_lens must be contiguous")
        if req_pool_indices.numel() < batch or prefix_lens.numel() < batch:
            raise ValueError("extend: req_pool_indices / prefix_lens shorter than the batch")
        a.req_to_token = req_to_token.data_ptr()
        a.req_to_token_stride = req_to_token.stride(0)
        a.req_pool_indices, a.prefix_lens = req_pool_indices.data_ptr(), prefix_lens.data_ptr()
        a.index_dtype = _IDX[prefix_lens.dtype]
    if custom_mask is not None:
        if custom_mask.dtype not in (torch.uint8, torch.bool) or not custom_mask.is_contiguous():
            raise TypeError("extend: custom_mask must be a contiguous bool / uint8 tensor")
        if mask_indptr is None or mask_indptr.dtype not in _IDX or not mask_indptr.is_contiguous():
            raise TypeError("extend: mask_indptr must be contiguous int32/int64 [batch + 1]")
        a.custom_mask, a.mask_indptr = custom_mask.data_ptr(), mask_indptr.data_ptr()
        a.mask_indptr_dtype = _IDX[mask_indptr.dtype]
    if lse is not None and (lse.dtype != torch.float32 or not lse.is_contiguous()
                            or lse.numel() < q.shape[0] * q.shape[1]):
        raise TypeError("extend: lse must be contiguous float32 [tokens, heads]")
    a.q, a.o, a.lse = q.data_ptr(), o.data_ptr(), _ptr(lse)
    a.q_stride_token, a.q_stride_head = q.stride(0), q.stride(1)
    a.o_stride_token, a.o_stride_head = o.stride(0), o.stride(1)
    a.k_pool, a.v_pool = k_pool.data_ptr(), v_pool.data_ptr()
    a.k_pool_stride_slot, a.v_pool_stride_slot = k_pool.stride(0), v_pool.stride(0)
    a.qo_indptr, a.kv_indptr, a.kv_indices = qo_indptr.data_ptr(), kv_indptr.data_ptr(), _ptr(kv_indices)
    a.extend_loc = extend_loc.data_ptr()
    a.batch, a.total_tokens, a.max_extend_len = batch, q.shape[0], max_extend_len
    a.num_q_heads, a.num_kv_heads = q.shape[1], k_pool.shape[1]
    a.head_dim, a.v_head_dim = q.shape[2], v_pool.shape[2]
    a.q_dtype, a.kv_dtype = dtype_code(q.dtype), dtype_code(k_pool.dtype)
    a.causal = 1 if causal else 0
    a.sliding_window = sliding_window
    a.skip_prefix_custom_mask = 1 if skip_prefix_custom_mask else 0
    a.sm_scale, a.logit_cap, a.k_scale, a.v_scale = sm_scale, logit_cap, k_scale, v_scale
    a.avg_kv_len_hint = max(0, min(int(avg_kv_len), 2**31 - 1))
    with torch.cuda.device(q.device):
        capi.check("b200attn_extend", capi.lib().b200attn_extend(C.byref(a), _stream(q)))


def _extend_plan(extend_seq_lens, prefix_lens, qo_indptr, kv_indptr, mask_indptr, uniform_extend_len,
                 sliding_window):
    """qo_indptr / kv_indptr / mask_indptr of one extend forward in one launch."""
    _cuda(extend_seq_lens, prefix_lens, qo_indptr, kv_indptr, mask_indptr)
    if prefix_lens.dtype not in _IDX or not prefix_lens.is_contiguous():
        raise TypeError("extend_plan: prefix_lens must be contiguous int32/int64")
    batch = prefix_lens.numel()
    if extend_seq_lens is not None and (extend_seq_lens.dtype != prefix_lens.dtype
                                        or not extend_seq_lens.is_contiguous()
                                        or extend_seq_lens.numel() != batch):
        raise TypeError("extend_plan: extend_seq_lens must match prefix_lens in dtype and length")
    if extend_seq_lens is None and uniform_extend_len <= 0:
        raise ValueError("extend_plan: give extend_seq_lens or uniform_extend_len > 0")
    for name, t in (("qo_indptr", qo_indptr), ("kv_indptr", kv_indptr)):
        if t.dtype != torch.int32 or not t.is_contiguous() or t.numel() < batch + 1:
            raise TypeError(f"extend_plan: {name} must be contiguous int32 [batch + 1]")
    if mask_indptr is not None and (mask_indptr.dtype != torch.int64 or not mask_indptr.is_contiguous()
                                    or mask_indptr.numel() < batch + 1):
        raise TypeError("extend_plan: mask_indptr must be contiguous int64 [batch + 1]")
    a = capi.ExtendPlanArgs()
    a.extend_seq_lens, a.prefix_lens = _ptr(extend_seq_lens), prefix_lens.data_ptr()
    a.index_dtype = _IDX[prefix_lens.dtype]
    a.batch = batch
    a.uniform_extend_len, a.sliding_window = uniform_extend_len, sliding_window
    a.qo_indptr, a.kv_indptr, a.mask_indptr = qo_indptr.data_ptr(), kv_indptr.data_ptr(), _ptr(mask_indptr)
    with torch.cuda.device(prefix_lens.device):
        capi.check("b200attn_extend_plan", capi.lib().b200attn_extend_plan(C.byref(a), _stream(prefix_lens)))


def _draft_kv_indices(req_pool_indices, req_to_token, seq_lens, positions, kv_indices, kv_indptr, topk, page_size):
    """kv_indptr / kv_indices of all draft-decode steps (``generate_draft_decode_kv_indices``,
    speculative/eagle_utils.py:844-920).  ``kv_indices`` ``[steps, n]`` and ``kv_indptr``
    ``[steps, rows + 1]`` int32 are written in place; ``kv_indptr[:, 0]`` is left alone."""
    _cuda(req_pool_indices, req_to_token, seq_lens, positions, kv_indices, kv_indptr)
    if req_to_token.dtype != torch.int32 or req_to_token.dim() != 2 or req_to_token.stride(1) != 1:
        raise TypeError("draft_kv_indices: req_to_token must be int32 [max_reqs, max_context_len]")
    if req_pool_indices.dtype != seq_lens.dtype or seq_lens.dtype not in _IDX:
        raise TypeError("draft_kv_indices: req_pool_indices and seq_lens must both be int32 or both int64")
    if positions.dtype not in _IDX or not positions.is_contiguous():
        raise TypeError("draft_kv_indices: positions must be contiguous int32/int64")
    if not (req_pool_indices.is_contiguous() and seq_lens.is_contiguous()):
        raise ValueError("draft_kv_indices: req_pool_indices / seq_lens must be contiguous")
    for name, t in (("kv_indices", kv_indices), ("kv_indptr", kv_indptr)):
        if t.dtype != torch.int32 or t.dim() != 2 or t.stride(1) != 1:
            raise TypeError(f"draft_kv_indices: {name} must be int32 [steps, n] with a contiguous last dim")
    num_seqs = seq_lens.numel()
    if kv_indices.shape[0] != kv_indptr.shape[0]:
        raise ValueError("draft_kv_indices: kv_indices and kv_indptr disagree on the number of steps")
    if positions.numel() < num_seqs * topk or kv_indptr.shape[1] < num_seqs * topk + 1:
        raise ValueError("draft_kv_indices: positions / kv_indptr shorter than num_seqs * topk")
    a = capi.DraftKvIndicesArgs()
    a.req_pool_indices, a.seq_lens = req_pool_indices.data_ptr(), seq_lens.data_ptr()
    a.req_to_token, a.req_to_token_stride = req_to_token.data_ptr(), req_to_token.stride(0)
    a.index_dtype, a.positions_dtype = _IDX[seq_lens.dtype], _IDX[positions.dtype]
    a.positions = positions.data_ptr()
    a.kv_indices, a.kv_indices_stride = kv_indices.data_ptr(), kv_indices.stride(0)
    a.kv_indptr, a.kv_indptr_stride = kv_indptr.data_ptr(), kv_indptr.stride(0)
    a.num_steps, a.num_seqs, a.topk, a.page_size = kv_indices.shape[0], num_seqs, topk, page_size
    with torch.cuda.device(kv_indices.device):
        capi.check("b200attn_draft_kv_indices", capi.lib().b200attn_draft_kv_indices(C.byref(a), _stream(kv_indices)))


def _merge_states(o_a, lse_a, o_b, lse_b, o_out, lse_out):
    _cuda(o_a, lse_a, o_b, lse_b, o_out, lse_out)
    if o_a.dim() != 3 or o_a.shape != o_b.shape or o_out.shape != o_a.shape:
        raise ValueError("merge_states: o tensors must be [tokens, heads, dim] of equal shape")
    for t in (o_a, o_b, o_out, lse_a, lse_b):
        if not t.is_contiguous():
            raise ValueError("merge_states: tensors must be contiguous")
    if lse_a.dtype != torch.float32 or lse_b.dtype != torch.float32:
        raise TypeError("merge_states: lse must be float32")
    a = capi.MergeStatesArgs()
    a.o_a, a.lse_a, a.o_b, a.lse_b = o_a.data_ptr(), lse_a.data_ptr(), o_b.data_ptr(), lse_b.data_ptr()
    a.o_out, a.lse_out = o_out.data_ptr(), _ptr(lse_out)
    a.tokens, a.heads, a.head_dim = o_a.shape
    a.dtype = capi.F32 if o_a.dtype == torch.float32 else dtype_code(o_a.dtype)
    with torch.cuda.device(o_a.device):
        capi.check("b200attn_merge_states", capi.lib().b200attn_merge_states(C.byref(a), _stream(o_a)))


# --------------------------------------------------------------------------- #
# registration: torch.ops.b200attn.*
# --------------------------------------------------------------------------- #
_LIB = torch.library.Library("b200attn", "DEF")
_LIB.define(
    "kv_store(Tensor(a!) k, Tensor(b!) v, Tensor loc, Tensor(c!) k_pool, Tensor(d!) v_pool, "
    "float? k_scale, float? v_scale, Tensor? k_scale_t, Tensor? v_scale_t, int scale_mode) -> ()"
)
_LIB.define(
    "rope_kv_store(Tensor(a!) q, Tensor(b!) k, Tensor(c!)? v, Tensor positions, Tensor cos_sin_cache, "
    "bool is_neox, Tensor? loc, Tensor(d!)? k_pool, Tensor(e!)? v_pool, float? k_scale, float? v_scale, "
    "Tensor? k_scale_t, Tensor? v_scale_t, int scale_mode) -> ()"
)
_LIB.define(
    "plan(Tensor req_to_token, Tensor req_pool_indices, Tensor seq_lens, Tensor(a!) plan_buf, "
    "Tensor(b!)? kv_indptr, Tensor(c!)? kv_indices, int num_q_heads, int num_kv_heads, int head_dim, "
    "int kv_dtype, int max_ctas, int min_chunks, int sliding_window=0) -> ()"
)
_LIB.define(
    "plan_csr(Tensor kv_indptr, Tensor(a!) plan_buf, int num_q_heads, int num_kv_heads, int head_dim, "
    "int kv_dtype, int max_ctas, int min_chunks) -> ()"
)
_LIB.define(
    "decode(Tensor q, Tensor k_pool, Tensor v_pool, Tensor(a!) o, Tensor(b!)? lse, Tensor plan_buf, "
    "Tensor indices, Tensor(c!) workspace, float sm_scale, float logit_cap, float k_scale, "
    "float v_scale, int phases=0, Tensor? slot_map=None) -> ()"
)
_LIB.define(
    "extend(Tensor q, Tensor(a!) o, Tensor(b!)? lse, Tensor k_pool, Tensor v_pool, Tensor qo_indptr, "
    "Tensor kv_indptr, Tensor? kv_indices, Tensor? req_to_token, Tensor? req_pool_indices, "
    "Tensor? prefix_lens, Tensor extend_loc, Tensor? custom_mask, Tensor? mask_indptr, int max_extend_len, "
    "bool causal, int sliding_window, bool skip_prefix_custom_mask, float sm_scale, float logit_cap, "
    "float k_scale=1.0, float v_scale=1.0, int avg_kv_len=0) -> ()"
)
_LIB.define(
    "extend_plan(Tensor? extend_seq_lens, Tensor prefix_lens, Tensor(a!) qo_indptr, Tensor(b!) kv_indptr, "
    "Tensor(c!)? mask_indptr, int uniform_extend_len, int sliding_window) -> ()"
)
_LIB.define(
    "merge_states(Tensor o_a, Tensor lse_a, Tensor o_b, Tensor lse_b, Tensor(a!) o_out, "
    "Tensor(b!)? lse_out) -> ()"
)
_LIB.define(
    "draft_kv_indices(Tensor req_pool_indices, Tensor req_to_token, Tensor seq_lens, Tensor positions, "
    "Tensor(a!) kv_indices, Tensor(b!) kv_indptr, int topk, int page_size) -> ()"
)
_LIB.impl("draft_kv_indices", _draft_kv_indices, "CUDA")
_LIB.impl("kv_store", _kv_store, "CUDA")
_LIB.impl("rope_kv_store", _rope_kv_store, "CUDA")
_LIB.impl("plan", _plan, "CUDA")
_LIB.impl("plan_csr", _plan_csr, "CUDA")
_LIB.impl("decode", _decode, "CUDA")
_LIB.impl("extend", _extend, "CUDA")
_LIB.impl("extend_plan", _extend_plan, "CUDA")
_LIB.impl("merge_states", _merge_states, "CUDA")


def plan_bytes(batch: int) -> int:
    return int(capi.lib().b200attn_plan_bytes(batch))


def workspace_bytes(batch: int, num_q_heads: int, num_kv_heads: int, head_dim: int,
                    v_head_dim: int, kv_dtype: int) -> int:
    n = int(capi.lib().b200attn_workspace_bytes(batch, num_q_heads, num_kv_heads, head_dim,
                                                v_head_dim, kv_dtype))
    if n == 0:
        raise capi.B200AttnError("b200attn_workspace_bytes", capi.E_UNSUPPORTED,
                                 capi.strerror(capi.E_UNSUPPORTED))
    return n
