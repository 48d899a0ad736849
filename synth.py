"""Synthetic paged-KV workloads for tests and bench.

The five named configurations of BASELINE.json, plus a free-form builder.  Index
structures (slot permutation, request rows, shared prefixes) are drawn on the
host with numpy so they are identical on every rank and device; pool contents
are drawn on the target device.

Slot layout rules of the reference that the generator honours:
* slot = page * page_size + offset, tokens of a page consecutive
  (python/sglang/srt/mem_cache/allocator.py:444-447);
* page 0 is the dummy page padded tokens write to (allocator.py:124-128), so
  real pages start at 1 and the pool has ``size + page_size`` rows
  (memory_pool.py:222-239);
* ``seq_lens`` already counts the token being decoded and
  ``req_to_token[req, S-1] == out_cache_loc`` (schedule_batch.py:1562-1590).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field, replace
from typing import Dict, Optional

import numpy as np
import torch


@dataclass(frozen=True)
class Config:
    name: str
    batch: int
    ctx: int                      # uniform context length (ignored when ragged)
    num_q_heads: int
    num_kv_heads: int
    head_dim: int
    kv_dtype: str = "bf16"        # "bf16" | "f16" | "fp8_e4m3" | "fp8_e5m2" (fp8 pools are uint8 byte tensors)
    q_dtype: str = "bf16"
    page_size: int = 1
    ragged: Optional[tuple] = None  # (min_len, max_len): log-uniform lengths
    share_groups: int = 0         # requests per prefix-sharing group (0 = none)
    k_scale: float = 1.0
    v_scale: float = 1.0
    seed: int = 1234
    layers_equiv: int = 32        # for the model-equivalent tokens/s figure


CONFIGS: Dict[str, Config] = {
    # configs[0]: Llama-3.2-1B-shaped, the reference's CPU-runnable case
    "C1": Config("C1", 16, 2048, 32, 8, 64, seed=1235, layers_equiv=16),
    # configs[1]: Llama-3.1-8B-shaped, the configuration the metric is quoted on
    "C2": Config("C2", 128, 4096, 32, 8, 128, seed=1236, layers_equiv=32),
    # configs[2]: 8B-shaped with FP8-e4m3 KV cache and per-tensor scales
    "C3": Config("C3", 256, 8192, 32, 8, 128, kv_dtype="fp8_e4m3", k_scale=0.5, v_scale=2.0,
                 seed=1237, layers_equiv=32),
    # configs[3]: Llama-3.1-70B-shaped, KV heads sharded over TP ranks
    "C4": Config("C4", 256, 8192, 64, 8, 128, seed=1238, layers_equiv=80),
    # configs[4]: ragged 512 requests, ctx 128..32k, pages of 16 (64 as a variant), shared prefixes
    "C5": Config("C5", 512, 0, 32, 8, 128, page_size=16, ragged=(128, 32768), share_groups=8,
                 seed=1239, layers_equiv=32),
}

_TORCH_DT = {"bf16": torch.bfloat16, "f16": torch.float16}
_FP8_DT = {"fp8_e4m3": torch.float8_e4m3fn, "fp8_e5m2": torch.float8_e5m2}


def shard(cfg: Config, tp_size: int) -> Config:
    """Per-rank shape under KV-head tensor parallelism (models/llama.py:119-132)."""
    if tp_size == 1:
        return cfg
    if cfg.num_kv_heads % tp_size != 0 or cfg.num_q_heads % tp_size != 0:
        raise ValueError(f"{cfg.name}: {cfg.num_kv_heads} kv heads do not shard over {tp_size} ranks")
    return replace(cfg, num_q_heads=cfg.num_q_heads // tp_size, num_kv_heads=cfg.num_kv_heads // tp_size)


def draw_seq_lens(cfg: Config, rng: np.random.Generator) -> np.ndarray:
    if cfg.ragged is None:
        return np.full(cfg.batch, cfg.ctx, dtype=np.int64)
    lo, hi = cfg.ragged
    u = rng.uniform(math.log(lo), math.log(hi), size=cfg.batch)
    return np.clip(np.rint(np.exp(u)), lo, hi).astype(np.int64)


@dataclass
class Workload:
    cfg: Config
    seq_lens_np: np.ndarray
    req_to_token: torch.Tensor       # int32 [max_reqs, max_ctx]
    req_pool_indices: torch.Tensor   # int64 [B]
    seq_lens: torch.Tensor           # int64 [B]
    out_cache_loc: torch.Tensor      # int64 [B]
    qkv: torch.Tensor                # [B, (Hq + 2 Hkv) * D] fused projection output
    q: torch.Tensor                  # views of qkv
    k: torch.Tensor
    v: torch.Tensor
    k_pool: torch.Tensor             # [slots, Hkv, D] (uint8 for fp8)
    v_pool: torch.Tensor
    pool_size: int                   # slots excluding the dummy page
    sm_scale: float
    unique_tokens: int
    extras: dict = field(default_factory=dict)

    @property
    def seq_lens_sum(self) -> int:
        return int(self.seq_lens_np.sum())

    def algorithmic_bytes(self, save_kv: bool = True) -> int:
        """KV read + q + o + (store: read new k/v, write pool rows) + index words."""
        c = self.cfg
        e_kv = 1 if c.kv_dtype.startswith("fp8") else 2
        total = self.seq_lens_sum * c.num_kv_heads * 2 * c.head_dim * e_kv
        total += 2 * c.batch * c.num_q_heads * c.head_dim * 2
        if save_kv:
            total += c.batch * c.num_kv_heads * 2 * c.head_dim * (2 + e_kv)
        total += 4 * self.seq_lens_sum
        return total


def make_workload(cfg: Config, device="cpu", max_reqs: Optional[int] = None, slack_pages: int = 8,
                  pool_fill: str = "randn") -> Workload:
    """Build pools + request tables for ``cfg`` on ``device``."""
    rng = np.random.default_rng(cfg.seed)
    P = cfg.page_size
    B = cfg.batch
    sl = draw_seq_lens(cfg, rng)
    max_ctx = int(sl.max()) + 4  # req_to_token is [max_reqs, context_len + 4] (model_runner.py:1170)
    max_reqs = max_reqs or max(B + 3, 8)

    pages_per_req = (sl + P - 1) // P
    # shared prefixes: groups of `share_groups` consecutive requests share their first pages
    shared = np.zeros(B, dtype=np.int64)
    if cfg.share_groups > 1:
        for g0 in range(0, B, cfg.share_groups):
            g = slice(g0, min(g0 + cfg.share_groups, B))
            shared[g] = int(0.5 * sl[g].min()) // P
    n_private = int((pages_per_req - shared).sum())
    n_shared = int(shared[:: max(cfg.share_groups, 1)].sum()) if cfg.share_groups > 1 else 0
    n_pages = n_private + n_shared + slack_pages
    perm = rng.permutation(n_pages) + 1  # page 0 is reserved
    cursor = 0

    r2t = np.zeros((max_reqs, max_ctx), dtype=np.int32)
    rpi = rng.permutation(max_reqs)[:B].astype(np.int64)
    offs = np.arange(P, dtype=np.int64)
    group_pages = None
    for b in range(B):
        npg = int(pages_per_req[b])
        pages = np.empty(npg, dtype=np.int64)
        ns = int(shared[b])
        if ns:
            if b % cfg.share_groups == 0:
                group_pages = perm[cursor : cursor + ns]
                cursor += ns
            pages[:ns] = group_pages
        pages[ns:] = perm[cursor : cursor + npg - ns]
        cursor += npg - ns
        slots = (pages[:, None] * P + offs[None, :]).reshape(-1)[: sl[b]]
        r2t[rpi[b], : sl[b]] = slots
    pool_size = n_pages * P
    rows = pool_size + P
    loc = r2t[rpi, sl - 1].astype(np.int64)

    dev = torch.device(device)
    gen = torch.Generator(device=dev)
    gen.manual_seed(cfg.seed)
    qdt = _TORCH_DT[cfg.q_dtype]
    Hq, Hkv, D = cfg.num_q_heads, cfg.num_kv_heads, cfg.head_dim
    qkv = torch.randn((B, (Hq + 2 * Hkv) * D), generator=gen, device=dev, dtype=torch.float32).to(qdt)
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)

    def fill(scale: float) -> torch.Tensor:
        if cfg.kv_dtype.startswith("fp8"):
            out = torch.empty((rows, Hkv, D), dtype=torch.uint8, device=dev)
        else:
            out = torch.empty((rows, Hkv, D), dtype=_TORCH_DT[cfg.kv_dtype], device=dev)
        step = max(1, (1 << 26) // (Hkv * D))  # draw in slabs to bound fp32 temporaries
        for r0 in range(0, rows, step):
            r1 = min(rows, r0 + step)
            if pool_fill == "zeros":
                x = torch.zeros((r1 - r0, Hkv, D), device=dev, dtype=torch.float32)
            else:
                x = torch.randn((r1 - r0, Hkv, D), generator=gen, device=dev, dtype=torch.float32)
            if cfg.kv_dtype.startswith("fp8"):
                out[r0:r1] = (x / scale).to(torch.bfloat16).to(_FP8_DT[cfg.kv_dtype]).view(torch.uint8)
            else:
                out[r0:r1] = x.to(out.dtype)
        return out

    k_pool = fill(cfg.k_scale)
    v_pool = fill(cfg.v_scale)

    unique = int(np.unique(np.concatenate([r2t[rpi[b], : sl[b]] for b in range(B)])).size)
    return Workload(
        cfg=cfg, seq_lens_np=sl,
        req_to_token=torch.from_numpy(r2t).to(dev),
        req_pool_indices=torch.from_numpy(rpi).to(dev),
        seq_lens=torch.from_numpy(sl).to(dev),
        out_cache_loc=torch.from_numpy(loc).to(dev),
        qkv=qkv, q=q, k=k, v=v, k_pool=k_pool, v_pool=v_pool, pool_size=pool_size,
        sm_scale=float(D) ** -0.5, unique_tokens=unique)


def shard_workload(w: Workload, rank: int, tp_size: int) -> Workload:
    """Rank ``rank``'s head slice of a global workload under KV-head tensor parallelism: kv heads
    ``[rank * Hkv / tp, (rank + 1) * Hkv / tp)`` with the ``Hq / Hkv`` q heads attached to each
    (models/llama.py:119-132; pool sized per rank, model_runner.py:1251-1263).  Request tables,
    ``seq_lens`` and ``out_cache_loc`` are shared by all ranks (same scheduler decisions)."""
    c = w.cfg
    rc = shard(c, tp_size)
    Hq, Hkv, D = c.num_q_heads, c.num_kv_heads, c.head_dim
    hq, hkv = rc.num_q_heads, rc.num_kv_heads
    q = w.q.reshape(c.batch, Hq, D)[:, rank * hq:(rank + 1) * hq]
    k = w.k.reshape(c.batch, Hkv, D)[:, rank * hkv:(rank + 1) * hkv]
    v = w.v.reshape(c.batch, Hkv, D)[:, rank * hkv:(rank + 1) * hkv]
    qkv = torch.cat([q.reshape(c.batch, -1), k.reshape(c.batch, -1), v.reshape(c.batch, -1)], dim=-1).contiguous()
    q, k, v = qkv.split([hq * D, hkv * D, hkv * D], dim=-1)
    return replace(w, cfg=rc, qkv=qkv, q=q, k=k, v=v,
                   k_pool=w.k_pool[:, rank * hkv:(rank + 1) * hkv].contiguous(),
                   v_pool=w.v_pool[:, rank * hkv:(rank + 1) * hkv].contiguous())


def small(name: str, **kw) -> Config:
    """A reduced copy of a named config for oracle-speed tests."""
    return replace(CONFIGS[name], **kw)


def cos_sin_cache(rotary_dim: int, max_position: int, base: float = 10000.0) -> torch.Tensor:
    """Synthetic-input helper: the fp32 ``[max_position, rotary_dim]`` cos | sin table a
    ``RotaryEmbedding`` holds (python/sglang/srt/layers/rotary_embedding.py:113-136)."""
    inv_freq = 1.0 / (base ** (torch.arange(0, rotary_dim, 2, dtype=torch.float) / rotary_dim))
    freqs = torch.einsum("i,j -> ij", torch.arange(max_position, dtype=torch.float), inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1)


# --------------------------------------------------------------------------- #
# extend (prefill with cached prefix) workloads
# --------------------------------------------------------------------------- #
@dataclass
class ExtendWorkload:
    """An EXTEND batch: request b has ``prefix_lens[b]`` cached tokens and ``extend_lens[b]`` new
    ones whose q/k/v rows sit back to back in ``qkv`` (forward_batch_info.py:226-236:
    extend_prefix_lens / extend_seq_lens; seq_lens = prefix + extend)."""
    cfg: Config
    prefix_lens_np: np.ndarray
    extend_lens_np: np.ndarray
    req_to_token: torch.Tensor       # int32 [max_reqs, max_ctx]
    req_pool_indices: torch.Tensor   # int64 [B]
    seq_lens: torch.Tensor           # int64 [B] prefix + extend
    extend_prefix_lens: torch.Tensor  # int64 [B]
    extend_seq_lens: torch.Tensor    # int64 [B]
    out_cache_loc: torch.Tensor      # int64 [tokens]: slot of every new token
    qkv: torch.Tensor                # [tokens, (Hq + 2 Hkv) * D]
    q: torch.Tensor
    k: torch.Tensor
    v: torch.Tensor
    k_pool: torch.Tensor
    v_pool: torch.Tensor
    pool_size: int
    sm_scale: float

    @property
    def tokens(self) -> int:
        return int(self.extend_lens_np.sum())

    def flops(self, causal: bool = True) -> int:
        """4 * D * (visible q-k pairs) * Hq: the algorithmic flops of QK^T and PV."""
        c = self.cfg
        pl, el = self.prefix_lens_np.astype(np.int64), self.extend_lens_np.astype(np.int64)
        pairs = pl * el + (el * (el + 1) // 2 if causal else el * el)
        return int(4 * c.head_dim * c.num_q_heads * pairs.sum())


def make_extend_workload(cfg: Config, prefix_lens, extend_lens, device="cpu", max_reqs: Optional[int] = None,
                         slack_pages: int = 4) -> ExtendWorkload:
    rng = np.random.default_rng(cfg.seed)
    P = cfg.page_size
    pl = np.asarray(prefix_lens, dtype=np.int64)
    el = np.asarray(extend_lens, dtype=np.int64)
    B = len(pl)
    sl = pl + el
    max_ctx = int(sl.max()) + 4
    max_reqs = max_reqs or max(B + 3, 8)
    pages_per_req = (sl + P - 1) // P
    n_pages = int(pages_per_req.sum()) + slack_pages
    perm = rng.permutation(n_pages) + 1
    r2t = np.zeros((max_reqs, max_ctx), dtype=np.int32)
    rpi = rng.permutation(max_reqs)[:B].astype(np.int64)
    offs = np.arange(P, dtype=np.int64)
    cursor = 0
    locs = []
    for b in range(B):
        npg = int(pages_per_req[b])
        pages = perm[cursor: cursor + npg]
        cursor += npg
        slots = (pages[:, None] * P + offs[None, :]).reshape(-1)[: sl[b]]
        r2t[rpi[b], : sl[b]] = slots
        locs.append(slots[pl[b]: sl[b]])
    loc = np.concatenate(locs).astype(np.int64) if locs else np.zeros(0, dtype=np.int64)
    pool_size = n_pages * P
    rows = pool_size + P

    dev = torch.device(device)
    gen = torch.Generator(device=dev)
    gen.manual_seed(cfg.seed)
    qdt = _TORCH_DT[cfg.q_dtype]
    Hq, Hkv, D = cfg.num_q_heads, cfg.num_kv_heads, cfg.head_dim
    T = int(el.sum())
    qkv = torch.randn((T, (Hq + 2 * Hkv) * D), generator=gen, device=dev, dtype=torch.float32).to(qdt)
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    fp8 = cfg.kv_dtype.startswith("fp8")
    kvdt = torch.uint8 if fp8 else _TORCH_DT[cfg.kv_dtype]   # fp8 pools are stored as raw bytes
    k_pool = torch.empty((rows, Hkv, D), dtype=kvdt, device=dev)
    v_pool = torch.empty((rows, Hkv, D), dtype=kvdt, device=dev)
    step = max(1, (1 << 26) // (Hkv * D))
    for t, scale in ((k_pool, cfg.k_scale), (v_pool, cfg.v_scale)):
        for r0 in range(0, rows, step):
            r1 = min(rows, r0 + step)
            x = torch.randn((r1 - r0, Hkv, D), generator=gen, device=dev, dtype=torch.float32)
            if fp8:
                t[r0:r1] = (x / scale).to(torch.bfloat16).to(_FP8_DT[cfg.kv_dtype]).view(torch.uint8)
            else:
                t[r0:r1] = x.to(kvdt)
    return ExtendWorkload(
        cfg=cfg, prefix_lens_np=pl, extend_lens_np=el,
        req_to_token=torch.from_numpy(r2t).to(dev), req_pool_indices=torch.from_numpy(rpi).to(dev),
        seq_lens=torch.from_numpy(sl).to(dev), extend_prefix_lens=torch.from_numpy(pl).to(dev),
        extend_seq_lens=torch.from_numpy(el).to(dev), out_cache_loc=torch.from_numpy(loc).to(dev),
        qkv=qkv, q=q, k=k, v=v, k_pool=k_pool, v_pool=v_pool, pool_size=pool_size,
        sm_scale=float(D) ** -0.5)
