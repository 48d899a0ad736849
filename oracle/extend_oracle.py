"""CPU oracle for the extend (prefill with cached prefix) attention path.  TEST INFRASTRUCTURE ONLY.

A numpy float64 restatement of what the reference computes for EXTEND / MIXED / TARGET_VERIFY /
DRAFT_EXTEND batches.  Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of the bench scripts may import this module; the product
(``ltp-sglang_b200/``) never does.

Parity status:
* causal / non-causal extend over a cached prefix: PINNED.  ``tests/golden/g10_extend_*.npz`` were
  produced by the reference's own ``TorchNativeAttnBackend.forward_extend`` (imported from the
  read-only tree by ``tests/golden/make_golden.py``); ``tests/test_oracle.py`` checks
  ``extend_attention`` against them.
* custom mask, sliding window, logit cap: the reference's only implementation is the Triton kernel
  (GPU-only; no stored vectors), so these follow its source line by line and are pinned by the
  properties the reference's own test asserts (test/srt/test_triton_attention_kernels.py:189-224:
  a custom mask equal to the causal mask reproduces the causal result) -- otherwise "parity unpinned".

What each function follows (paths relative to the reference tree):

* ``extend_attention``   python/sglang/srt/layers/attention/triton_ops/extend_attention.py:41-303
                         (prefix stage :131-204 with its window / custom masks, extend stage
                         :206-287 with custom mask | causal | none) and
                         layers/attention/torch_native_backend.py:27-110 (gather through
                         req_to_token, causal SDPA, keep the extend rows)
* ``extend_indptrs``     layers/attention/triton_backend.py:227-318 (qo_indptr, kv_indptr,
                         window_kv_indptr :927-956, mask_indptr :252-259)
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from .decode_oracle import pool_to_f32


def extend_indptrs(prefix_lens: np.ndarray, extend_lens: np.ndarray, sliding_window: int = -1):
    """(qo_indptr int32, kv_indptr int32, mask_indptr int64).  With ``sliding_window > 0`` kv_indptr is
    the reference's window_kv_indptr: cumsum(min(prefix, W + 1))."""
    pl = np.asarray(prefix_lens, dtype=np.int64)
    el = np.asarray(extend_lens, dtype=np.int64)
    lp = np.minimum(pl, sliding_window + 1) if sliding_window > 0 else pl
    qo = np.zeros(len(pl) + 1, dtype=np.int32)
    kv = np.zeros(len(pl) + 1, dtype=np.int32)
    mi = np.zeros(len(pl) + 1, dtype=np.int64)
    qo[1:] = np.cumsum(el)
    kv[1:] = np.cumsum(lp)
    mi[1:] = np.cumsum(el * (pl + el))
    return qo, kv, mi


def extend_attention(q: torch.Tensor, k_pool: torch.Tensor, v_pool: torch.Tensor,
                     prefix_slots: Sequence[np.ndarray], extend_slots: Sequence[np.ndarray],
                     sm_scale: float, causal: bool = True, logit_cap: float = 0.0,
                     custom_mask: Optional[np.ndarray] = None, mask_indptr: Optional[np.ndarray] = None,
                     sliding_window: int = -1, skip_prefix_custom_mask: bool = True,
                     k_scale: float = 1.0, v_scale: float = 1.0) -> Tuple[np.ndarray, np.ndarray]:
    """``q`` [tokens, Hq, D] (requests back to back); request b attends to the pool rows
    ``prefix_slots[b]`` (its cached prefix list, already windowed if a window is used) followed by
    ``extend_slots[b]`` (its own new tokens, one per query row).  Returns (o [tokens, Hq, D] float64,
    lse [tokens, Hq] float64 = ln sum exp(scaled, capped logits); -inf and o = 0 for a row that sees
    no key)."""
    T, Hq, D = q.shape
    Hkv = k_pool.shape[1]
    G = Hq // Hkv
    qf = q.detach().to(torch.float32).cpu().numpy().astype(np.float64)
    kf = pool_to_f32(k_pool)
    vf = pool_to_f32(v_pool)
    out = np.zeros((T, Hq, D), dtype=np.float64)
    lse = np.full((T, Hq), -np.inf, dtype=np.float64)
    t0 = 0
    for b, (ps, es) in enumerate(zip(prefix_slots, extend_slots)):
        ps = np.asarray(ps, dtype=np.int64)
        es = np.asarray(es, dtype=np.int64)
        Lp, Le = len(ps), len(es)
        if Le == 0:
            continue
        slots = np.concatenate([ps, es])
        K = kf[slots].astype(np.float64) * k_scale      # [L, Hkv, D]
        V = vf[slots].astype(np.float64) * v_scale
        i = np.arange(Le)[:, None]
        j = np.arange(Lp + Le)[None, :]
        vis = np.ones((Le, Lp + Le), dtype=bool)
        m = None
        if custom_mask is not None:
            ms = int(mask_indptr[b])
            m = np.asarray(custom_mask[ms: ms + Le * (Lp + Le)]).reshape(Le, Lp + Le) != 0
        # prefix part (extend_attention.py:166-184)
        if sliding_window > 0:
            vis[:, :Lp] &= (i <= j[:, :Lp] + sliding_window)
        if m is not None and not skip_prefix_custom_mask:
            vis[:, :Lp] &= m[:, :Lp]
        # extend part (:245-264)
        if m is not None:
            vis[:, Lp:] &= m[:, Lp:]
        elif causal:
            vis[:, Lp:] &= (j[:, Lp:] - Lp <= i)
        for h in range(Hq):
            s = qf[t0:t0 + Le, h] @ K[:, h // G].T * sm_scale   # [Le, L]
            if logit_cap > 0:
                s = logit_cap * np.tanh(s / logit_cap)
            s = np.where(vis, s, -np.inf)
            mx = s.max(axis=1, keepdims=True)
            dead = ~np.isfinite(mx[:, 0])
            mx = np.where(np.isfinite(mx), mx, 0.0)
            p = np.exp(s - mx)
            den = p.sum(axis=1, keepdims=True)
            with np.errstate(divide="ignore", invalid="ignore"):
                o = np.where(den > 0, (p @ V[:, h // G]) / np.where(den > 0, den, 1.0), 0.0)
                l = np.where(den[:, 0] > 0, mx[:, 0] + np.log(np.where(den[:, 0] > 0, den[:, 0], 1.0)), -np.inf)
            o[dead] = 0.0
            out[t0:t0 + Le, h] = o
            lse[t0:t0 + Le, h] = l
        t0 += Le
    return out, lse


def slots_from_req_to_token(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, prefix_lens,
                            extend_lens, sliding_window: int = -1):
    """Per-request (prefix list, extend list) the way the reference backends derive them:
    prefix = req_to_token[req, start : prefix] with start = prefix - min(prefix, W + 1) under a window
    (triton_backend.py:286-311, 927-956), extend = req_to_token[req, prefix : prefix + extend]
    (== out_cache_loc of those tokens; torch_native_backend.py:83-86 gathers both at once)."""
    r2t = req_to_token.cpu().numpy()
    rpi = req_pool_indices.cpu().numpy()
    ps, es = [], []
    for b in range(len(rpi)):
        pl, el = int(prefix_lens[b]), int(extend_lens[b])
        start = pl - min(pl, sliding_window + 1) if sliding_window > 0 else 0
        ps.append(r2t[rpi[b], start:pl].astype(np.int64))
        es.append(r2t[rpi[b], pl:pl + el].astype(np.int64))
    return ps, es
