"""CPU oracle for the paged decode-attention path.  TEST INFRASTRUCTURE ONLY.

A plain numpy / torch-CPU restatement of what the reference computes on this
path.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` leg may import this module; the product
(``ltp-sglang_b200/``) never does.

Parity status: PINNED.  ``tests/golden/*.npz`` were produced by running the
reference's own Python (``TorchNativeAttnBackend.forward_decode`` +
``MHATokenToKVPool.set_kv_buffer``, imported from the read-only reference tree
by ``tests/golden/make_golden.py``) and ``tests/test_oracle.py`` checks every
function below against them; the fp8 cast is additionally checked against
torch's own ``.to(float8_e4m3fn)`` on all 65536 bf16 bit patterns.

What each function follows (paths relative to the reference tree):

* ``kv_store``            python/sglang/srt/mem_cache/memory_pool.py:369-407
* ``kv_indices``          python/sglang/srt/layers/attention/utils.py:10-46 and
                          test/srt/test_create_kvindices.py:43-49 (the Python gather
                          the reference's own test uses as ground truth)
* ``decode_attention``    python/sglang/srt/layers/attention/torch_native_backend.py:112-180
                          (per-request gather + softmax(scale q.K^T) V), in float64,
                          plus lse = ln sum exp(scale q.K^T) as defined by
                          triton_ops/decode_attention.py:385-401
* ``torch_native_decode`` the same loop through torch SDPA in the model dtype (what the
                          reference backend literally executes; used as the timed CPU port)
* ``merge_states``        sgl-kernel/csrc/attention/merge_attn_states.cu:31-106
* ``cos_sin_cache`` / ``rope``  python/sglang/srt/layers/rotary_embedding.py:113-166
                          (``_compute_cos_sin_cache`` and ``forward_native`` with
                          ``_apply_rotary_emb`` :49-75), in float64 on the fp32 cache: the
                          arithmetic of the reference's CUDA path
                          (sgl-kernel/csrc/elementwise/rope.cu:22-89 keeps the cache in fp32)
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np
import torch

# --------------------------------------------------------------------------- #
# bit-level casts
# --------------------------------------------------------------------------- #


def f32_to_bf16_bits(x: np.ndarray) -> np.ndarray:
    """float32 -> bfloat16 bits, round-to-nearest-even (NaN -> 0x7FC0 like torch CPU)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    b = x.view(np.uint32)
    nan = np.isnan(x)
    rounded = ((b + (np.uint32(0x7FFF) + ((b >> np.uint32(16)) & np.uint32(1)))) >> np.uint32(16)).astype(np.uint16)
    return np.where(nan, np.uint16(0x7FC0), rounded)


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (np.ascontiguousarray(b, dtype=np.uint16).astype(np.uint32) << np.uint32(16)).view(np.float32)


def f32_to_e4m3fn_bits(x: np.ndarray) -> np.ndarray:
    """float32 -> float8_e4m3fn bits the way torch casts: RNE, NOT saturating.

    |x| >= 480 and NaN -> 0x7F | sign; (464, 480) rounds up into 0x7F; [448, 464] -> 0x7E;
    subnormals kept (2^-9 -> 0x01, 2^-10 ties to even -> 0x00).  SURVEY.md Appendix A.7.
    """
    x = np.ascontiguousarray(x, dtype=np.float32)
    bits = x.view(np.uint32).copy()
    sign = bits & np.uint32(0x80000000)
    bits ^= sign
    out = np.zeros(bits.shape, dtype=np.uint8)

    over = bits >= np.uint32(1087 << 20)
    sub = (~over) & (bits < np.uint32(121 << 23))
    norm = ~(over | sub)

    out[over] = 0x7F
    magic = np.uint32(141 << 23)
    if sub.any():
        a = bits[sub].view(np.float32) + np.array([magic], dtype=np.uint32).view(np.float32)[0]
        out[sub] = (a.astype(np.float32).view(np.uint32) - magic).astype(np.uint8)
    if norm.any():
        nb = bits[norm].astype(np.uint64)
        odd = (nb >> 20) & 1
        nb = (nb + np.uint64(((7 - 127) << 23) & 0xFFFFFFFF) + np.uint64(0x7FFFF) + odd) & np.uint64(0xFFFFFFFF)
        out[norm] = (nb >> 20).astype(np.uint8)
    return out | (sign >> np.uint32(24)).astype(np.uint8)


def f32_to_e5m2_bits(x: np.ndarray) -> np.ndarray:
    """float32 -> float8_e5m2 bits the way torch casts (c10/util/Float8_e5m2.h, what
    ``MHATokenToKVPool.set_kv_buffer`` runs for ``--kv-cache-dtype fp8_e5m2``, memory_pool.py:385-391): RNE,
    NOT saturating -- |x| >= 61440 rounds into infinity (0x7C), NaN -> 0x7F, subnormals kept.  Checked against
    torch on all 65 536 bf16 patterns by tests/test_oracle.py."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    bits = x.view(np.uint32).copy()
    sign = bits & np.uint32(0x80000000)
    bits ^= sign
    out = np.zeros(bits.shape, dtype=np.uint8)
    over = bits >= np.uint32(143 << 23)
    sub = (~over) & (bits < np.uint32(113 << 23))
    norm = ~(over | sub)
    out[over] = np.where(bits[over] > np.uint32(255 << 23), 0x7F, 0x7C).astype(np.uint8)
    magic = np.uint32(134 << 23)
    if sub.any():
        a = bits[sub].view(np.float32) + np.array([magic], dtype=np.uint32).view(np.float32)[0]
        out[sub] = (a.astype(np.float32).view(np.uint32) - magic).astype(np.uint8)
    if norm.any():
        nb = bits[norm].astype(np.uint64)
        odd = (nb >> 21) & 1
        nb = (nb + np.uint64(((15 - 127) << 23) & 0xFFFFFFFF) + np.uint64(0xFFFFF) + odd) & np.uint64(0xFFFFFFFF)
        out[norm] = (nb >> 21).astype(np.uint8)
    return out | (sign >> np.uint32(24)).astype(np.uint8)


def e5m2_bits_to_f32(b: np.ndarray) -> np.ndarray:
    """float8_e5m2 bits -> float32: the byte is the upper half of an IEEE half (exact)."""
    h = (np.ascontiguousarray(b, dtype=np.uint8).astype(np.uint16) << np.uint16(8)).view(np.float16)
    return h.astype(np.float32)


_E4M3_TABLE: Optional[np.ndarray] = None


def e4m3fn_bits_to_f32(b: np.ndarray) -> np.ndarray:
    """float8_e4m3fn bits -> float32 (exact; 0x7F / 0xFF are NaN)."""
    global _E4M3_TABLE
    if _E4M3_TABLE is None:
        t = np.zeros(256, dtype=np.float32)
        for i in range(256):
            s = -1.0 if i & 0x80 else 1.0
            e = (i >> 3) & 0xF
            m = i & 7
            if e == 0xF and m == 7:
                v = np.nan
            elif e == 0:
                v = s * (m / 8.0) * 2.0 ** -6
            else:
                v = s * (1.0 + m / 8.0) * 2.0 ** (e - 7)
            t[i] = v
        _E4M3_TABLE = t
    return _E4M3_TABLE[np.ascontiguousarray(b, dtype=np.uint8)]


def _to_f32(t: torch.Tensor) -> np.ndarray:
    return t.detach().to(torch.float32).cpu().numpy()


# --------------------------------------------------------------------------- #
# KV store
# --------------------------------------------------------------------------- #
SCALE_NONE, SCALE_TRUE_DIV, SCALE_RECIP_MUL = 0, 1, 2


def _round_to(x32: np.ndarray, dtype: torch.dtype) -> np.ndarray:
    """Round float32 values to the source dtype and return them as float32."""
    if dtype == torch.bfloat16:
        return bf16_bits_to_f32(f32_to_bf16_bits(x32))
    if dtype == torch.float16:
        return x32.astype(np.float16).astype(np.float32)
    raise TypeError(dtype)


def kv_store(k_pool: torch.Tensor, v_pool: torch.Tensor, loc: torch.Tensor, cache_k: torch.Tensor,
             cache_v: torch.Tensor, k_scale: Optional[float] = None, v_scale: Optional[float] = None,
             scale_mode: int = SCALE_TRUE_DIV, scale_is_cuda_tensor: bool = False) -> None:
    """In-place ``set_kv_buffer`` on CPU tensors.

    ``k_pool``/``v_pool``: one layer's ``[slots, heads, dim]`` pool, either the source dtype or
    ``uint8`` (fp8-e4m3 bytes).  With an fp8 pool and a scale: ``x <- round_src(x / s)`` is
    written back into ``cache_k``/``cache_v`` (``div_`` is in place, memory_pool.py:385-389),
    then cast.  ``scale_mode`` picks true division (torch CPU; CUDA with a tensor divisor) or
    multiplication by the fp32 reciprocal (torch CUDA with a Python-float divisor).
    ``scale_is_cuda_tensor``: the divisor is a 0-dim CUDA tensor, which torch casts to the source
    dtype before dividing.
    """
    loc_np = loc.detach().cpu().numpy().astype(np.int64)
    for pool, cache, scale in ((k_pool, cache_k, k_scale), (v_pool, cache_v, v_scale)):
        if pool.dtype in (torch.uint8, torch.float8_e4m3fn, torch.float8_e5m2):   # (uint8 = e4m3 bytes)
            x = _to_f32(cache)
            if scale is not None and scale_mode != SCALE_NONE:
                s = np.float32(scale)
                if scale_is_cuda_tensor:
                    s = _round_to(np.array([s], dtype=np.float32), cache.dtype)[0]
                if scale_mode == SCALE_RECIP_MUL:
                    x = x * (np.float32(1.0) / s)
                else:
                    x = x / s
                x = _round_to(x.astype(np.float32), cache.dtype)
                cache.copy_(torch.from_numpy(x).to(cache.dtype))
            bits = f32_to_e5m2_bits(x) if pool.dtype == torch.float8_e5m2 else f32_to_e4m3fn_bits(x)
            pool_np = pool.view(torch.uint8).numpy()
            pool_np[loc_np] = bits  # last write wins for duplicate slots, like index_put
        else:
            assert pool.dtype == cache.dtype
            pool[torch.from_numpy(loc_np)] = cache


# --------------------------------------------------------------------------- #
# page indexing
# --------------------------------------------------------------------------- #
def kv_indices(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
               seq_lens: torch.Tensor, sliding_window: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """(kv_indptr[B+1], kv_indices[sum S]) as int32 numpy arrays.  With ``sliding_window`` W > 0 the
    window_kv_indptr / window_kv_indices of update_sliding_window_buffer (triton_backend.py:927-956):
    lens = min(S, W + 1), start = S - lens."""
    r2t = req_to_token.detach().cpu().numpy()
    rpi = req_pool_indices.detach().cpu().numpy().astype(np.int64)
    sl = seq_lens.detach().cpu().numpy().astype(np.int64)
    wl = np.minimum(sl, sliding_window + 1) if sliding_window > 0 else sl
    indptr = np.zeros(len(sl) + 1, dtype=np.int32)
    indptr[1:] = np.cumsum(wl).astype(np.int32)
    parts = [r2t[rpi[i], sl[i] - wl[i]: sl[i]] for i in range(len(sl))]
    flat = np.concatenate(parts).astype(np.int32) if parts else np.zeros(0, dtype=np.int32)
    return indptr, flat


def draft_decode_kv_indices(req_to_token, req_pool_indices, seq_lens, positions, num_steps: int, topk: int,
                            page_size: int, kv_indices_width: int):
    """All draft-decode steps' (kv_indptr [steps, bs*topk + 1], kv_indices [steps, width]) -- a loop-level
    restatement of the Triton kernel ``generate_draft_decode_kv_indices``
    (python/sglang/srt/speculative/eagle_utils.py:844-920; grid (num_steps, num_seqs, topk)).  Parity
    UNPINNED: the reference kernel needs Triton on a GPU and holds no golden vectors; this follows its source
    line by line.  Unwritten entries stay 0 (the caller's zero-initialised buffers)."""
    r2t = req_to_token.detach().cpu().numpy()
    rpi = req_pool_indices.detach().cpu().numpy().astype(np.int64)
    sl = seq_lens.detach().cpu().numpy().astype(np.int64)
    pos = positions.detach().cpu().numpy().astype(np.int64)
    n = len(sl)
    indptr = np.zeros((num_steps, n * topk + 1), dtype=np.int32)
    indices = np.zeros((num_steps, kv_indices_width), dtype=np.int32)
    for it in range(num_steps):
        iters = it + 1
        for b in range(n):
            S = int(sl[b])
            cum = int(sl[:b].sum())
            row = r2t[rpi[b]]
            for t in range(topk):
                off = cum * topk + b * iters * topk + t * (S + iters)
                indices[it, off: off + S] = row[:S]
                if page_size == 1 or topk == 1:
                    ext = S + t * num_steps
                else:
                    last = S % page_size
                    pages = (last + num_steps + page_size - 1) // page_size
                    ext = S // page_size * page_size + t * pages * page_size + last
                indices[it, off + S: off + S + iters] = row[ext: ext + iters]
                z = b * topk + t
                if z == 0:
                    z = n * topk
                indptr[it, z] = int(pos[:z].sum()) + z * iters
    return indptr, indices


# --------------------------------------------------------------------------- #
# decode attention (float64 SDPA + lse)
# --------------------------------------------------------------------------- #
def pool_to_f32(pool: torch.Tensor) -> np.ndarray:
    if pool.dtype == torch.float8_e5m2:
        return e5m2_bits_to_f32(pool.view(torch.uint8).cpu().numpy())
    if pool.dtype in (torch.uint8, torch.float8_e4m3fn):
        return e4m3fn_bits_to_f32(pool.view(torch.uint8).cpu().numpy())
    return _to_f32(pool)


def decode_attention(q: torch.Tensor, k_pool: torch.Tensor, v_pool: torch.Tensor,
                     req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, seq_lens: torch.Tensor,
                     sm_scale: float, logit_cap: float = 0.0, k_scale: float = 1.0,
                     v_scale: float = 1.0, sliding_window: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """Returns (out [B, Hq, Dv] float32, lse [B, Hq] float32), computed in float64.

    Request b attends to slots ``req_to_token[req_pool_indices[b], :seq_lens[b]]``
    (torch_native_backend.py:160-163); q head h uses kv head ``h // (Hq / Hkv)`` (SDPA
    ``enable_gqa``).  fp8 pools are dequantised as ``fp8 * scale``
    (layers/quantization/kv_cache.py:21-23).  ``logit_cap``: ``cap * tanh(x / cap)`` after the
    scale (triton_ops/decode_attention.py:349-352).  ``sliding_window`` W > 0: only the last
    ``min(S, W + 1)`` tokens, ``req_to_token[req, S - len : S]`` (update_sliding_window_buffer,
    layers/attention/triton_backend.py:927-956).
    """
    qf = _to_f32(q).astype(np.float64)
    B, Hq, D = qf.shape
    kf = pool_to_f32(k_pool)
    vf = pool_to_f32(v_pool)
    Hkv, Dv = kf.shape[1], vf.shape[2]
    G = Hq // Hkv
    r2t = req_to_token.detach().cpu().numpy()
    rpi = req_pool_indices.detach().cpu().numpy().astype(np.int64)
    sl = seq_lens.detach().cpu().numpy().astype(np.int64)
    out = np.zeros((B, Hq, Dv), dtype=np.float32)
    lse = np.full((B, Hq), -np.inf, dtype=np.float32)
    for b in range(B):
        S = int(sl[b])
        if S == 0:
            continue
        wl = min(S, sliding_window + 1) if sliding_window > 0 else S
        slots = r2t[rpi[b], S - wl:S].astype(np.int64)
        K = kf[slots].astype(np.float64) * k_scale  # [S, Hkv, D]
        V = vf[slots].astype(np.float64) * v_scale
        qg = qf[b].reshape(Hkv, G, D)
        logits = np.einsum("hgd,shd->hgs", qg, K) * sm_scale
        if logit_cap > 0:
            logits = logit_cap * np.tanh(logits / logit_cap)
        m = logits.max(axis=-1, keepdims=True)
        p = np.exp(logits - m)
        l = p.sum(axis=-1, keepdims=True)
        o = np.einsum("hgs,shd->hgd", p / l, V)
        out[b] = o.reshape(Hq, Dv).astype(np.float32)
        lse[b] = (m + np.log(l)).reshape(Hq).astype(np.float32)
    return out, lse


def torch_native_decode(q: torch.Tensor, k_pool: torch.Tensor, v_pool: torch.Tensor,
                        req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
                        seq_lens: torch.Tensor, sm_scale: float) -> torch.Tensor:
    """What the reference's torch_native backend executes for decode: one SDPA per request on
    gathered K/V, in the model dtype (torch_native_backend.py:147-178).  Returns [B, Hq, Dv]."""
    B, Hq, _ = q.shape
    Hkv = k_pool.shape[1]
    out = torch.empty((B, Hq, v_pool.shape[2]), dtype=q.dtype)
    qh = q.movedim(0, 1)  # [Hq, B, D]
    for b in range(B):
        S = int(seq_lens[b])
        slots = req_to_token[int(req_pool_indices[b]), :S].to(torch.int64)
        K = k_pool[slots].movedim(0, 1)
        V = v_pool[slots].movedim(0, 1)
        o = torch.nn.functional.scaled_dot_product_attention(
            qh[:, b : b + 1, :].unsqueeze(0), K.unsqueeze(0), V.unsqueeze(0),
            enable_gqa=Hq != Hkv, scale=sm_scale, is_causal=False)
        out[b] = o.squeeze(0).movedim(1, 0)[0]
    return out


# --------------------------------------------------------------------------- #
# state merge
# --------------------------------------------------------------------------- #
def merge_states(o_a: np.ndarray, lse_a: np.ndarray, o_b: np.ndarray,
                 lse_b: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """2-way (o, lse) merge; +inf lse is treated as -inf (merge_attn_states.cu:68-69)."""
    a = np.where(np.isinf(lse_a), -np.inf, lse_a).astype(np.float64)
    b = np.where(np.isinf(lse_b), -np.inf, lse_b).astype(np.float64)
    m = np.maximum(a, b)
    with np.errstate(invalid="ignore"):
        ea, eb = np.exp(a - m), np.exp(b - m)
    s = ea + eb
    out = o_a.astype(np.float64) * (ea / s)[..., None] + o_b.astype(np.float64) * (eb / s)[..., None]
    return out.astype(np.float32), (np.log(s) + m).astype(np.float32)


def merge_n(o: np.ndarray, lse: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """N-way merge over axis 0 (stage-2 semantics, triton_ops/decode_attention.py:529-552)."""
    l = lse.astype(np.float64)
    m = l.max(axis=0)
    w = np.exp(l - m)
    s = w.sum(axis=0)
    out = (o.astype(np.float64) * w[..., None]).sum(axis=0) / s[..., None]
    return out.astype(np.float32), (m + np.log(s)).astype(np.float32)


# --------------------------------------------------------------------------- #
# rotary embedding
# --------------------------------------------------------------------------- #
def cos_sin_cache(rotary_dim: int, max_position: int, base: float = 10000.0) -> torch.Tensor:
    """fp32 [max_position, rotary_dim] cache, cos | sin halves (rotary_embedding.py:113-136)."""
    inv_freq = 1.0 / (base ** (torch.arange(0, rotary_dim, 2, dtype=torch.float) / rotary_dim))
    t = torch.arange(max_position, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1)


def rope(positions: torch.Tensor, x: torch.Tensor, cache: torch.Tensor, is_neox: bool = True) -> np.ndarray:
    """Rotates ``x`` [tokens, heads, head_dim] (rotary_embedding.py:138-166, :49-75); returns float64.

    The first ``rotary_dim = cache.shape[1]`` dims of each head rotate, the rest pass through.
    neox: pairs (i, i + rotary_dim/2); otherwise (2i, 2i+1).  The caller rounds to the model dtype.
    """
    xf = _to_f32(x).astype(np.float64)
    rot = cache.shape[1]
    cs = cache.detach().cpu().numpy().astype(np.float64)[positions.detach().cpu().numpy().astype(np.int64)]
    cos, sin = cs[:, None, : rot // 2], cs[:, None, rot // 2:]
    out = xf.copy()
    if is_neox:
        x1, x2 = xf[..., : rot // 2], xf[..., rot // 2: rot]
        out[..., : rot // 2] = x1 * cos - x2 * sin
        out[..., rot // 2: rot] = x2 * cos + x1 * sin
    else:
        x1, x2 = xf[..., 0:rot:2], xf[..., 1:rot:2]
        out[..., 0:rot:2] = x1 * cos - x2 * sin
        out[..., 1:rot:2] = x2 * cos + x1 * sin
    return out


def round_to_dtype(x: np.ndarray, dtype: torch.dtype) -> torch.Tensor:
    """float64/float32 array -> torch tensor of ``dtype`` with one RNE rounding from float32."""
    return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).to(dtype)
