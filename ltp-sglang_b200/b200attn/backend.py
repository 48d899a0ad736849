"""``B200AttnBackend``: the sm_100a paged attention backend (decode + extend).

Drop-in for the reference's ``AttentionBackend`` plug-in surface
(python/sglang/srt/layers/attention/base_attn_backend.py:14-117); host logic
mirrors ``TritonAttnBackend`` (layers/attention/triton_backend.py:41-732):

* ``init_forward_metadata``      one ``plan`` launch per forward (kv_indptr scan,
  optional kv_indices, split plan) instead of cumsum + Triton index kernel +
  Triton split heuristic (:160-221);
* ``forward_decode``             ``kv_store`` + ``decode`` (2 + 1 launches)
  instead of 2-6 ATen launches + 2 Triton launches (:687-732);
* CUDA-graph hooks               all buffers are allocated once in ``__init__``,
  so capture / replay only re-run the plan (:338-627);
* ``forward_extend``             ``kv_store`` + ``extend`` (tcgen05 / TMEM kernel for head_dim 128,
  mma.sync kernel for head_dim 64) over cached prefix + new tokens, for EXTEND / MIXED /
  TARGET_VERIFY / DRAFT_EXTEND batches (:227-318, :632-685); with ``prefill_backend=`` it
  delegates instead, the way ``HybridAttnBackend`` does (hybrid_attn_backend.py:88-100).

Registration in sglang is one ``elif`` in
``ModelRunner._get_attention_backend_from_str`` (model_runner.py:1374-1460) and
one choice string (server_args.py:1263-1279); see INTEGRATION.md.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Optional

import torch

from . import capi, ops
from .interfaces import AttentionBackend, ForwardMode


def store_kv(k_pool: torch.Tensor, v_pool: torch.Tensor, loc: torch.Tensor, cache_k: torch.Tensor,
             cache_v: torch.Tensor, k_scale=None, v_scale=None, scale_mode: Optional[int] = None) -> None:
    """``set_kv_buffer`` semantics (memory_pool.py:369-407) on raw layer pools.

    Scales apply only when the pool dtype differs from the source dtype (fp8).
    The arithmetic follows what the reference's ``cache_k.div_(k_scale)`` does on
    CUDA: a Python float multiplies by the fp32 reciprocal (torch's CPU-scalar
    fast path); a 0-dim CUDA tensor (``layer.k_scale``) is first cast to the
    source dtype, then divides in fp32.  ``scale_mode`` overrides the choice
    (``capi.SCALE_TRUE_DIV`` with floats reproduces torch's CPU result).
    ``cache_k`` / ``cache_v`` are updated in place like ``div_`` does.
    """
    if loc.dtype != torch.int64:
        loc = loc.to(torch.int64)
    fp8 = k_pool.dtype in (torch.uint8, torch.float8_e4m3fn, torch.float8_e5m2)
    ks = vs = kt = vt = None
    mode = capi.SCALE_NONE
    if fp8 and (k_scale is not None or v_scale is not None):
        tensors = isinstance(k_scale, torch.Tensor) or isinstance(v_scale, torch.Tensor)
        mode = capi.SCALE_TRUE_DIV if tensors else capi.SCALE_RECIP_MUL
        if scale_mode is not None:
            mode = scale_mode
        if tensors:
            kt = _scale_tensor(k_scale, cache_k.device)
            vt = _scale_tensor(v_scale, cache_v.device)
        else:
            ks = 1.0 if k_scale is None else float(k_scale)
            vs = 1.0 if v_scale is None else float(v_scale)
    torch.ops.b200attn.kv_store(cache_k, cache_v, loc, k_pool, v_pool, ks, vs, kt, vt, mode)


def rope_store_kv(q: torch.Tensor, k: torch.Tensor, v: Optional[torch.Tensor], positions: torch.Tensor,
                  cos_sin_cache: torch.Tensor, is_neox: bool = True, loc: Optional[torch.Tensor] = None,
                  k_pool: Optional[torch.Tensor] = None, v_pool: Optional[torch.Tensor] = None,
                  k_scale=None, v_scale=None, scale_mode: Optional[int] = None) -> None:
    """One launch for ``rotary_emb(positions, q, k)`` + ``set_kv_buffer(layer, loc, k, v, ...)``
    (models/llama.py:186-189; layers/rotary_embedding.py:218-236; memory_pool.py:369-407).

    ``q`` / ``k`` ``[tokens, heads, head_dim]`` are rotated in place (views of the fused qkv buffer are
    fine); with ``loc`` and the layer's pools the rotated k and v are stored with ``store_kv``
    semantics, and attention is then called with ``save_kv_cache=False``.  ``cos_sin_cache`` is the
    fp32 ``RotaryEmbedding.cos_sin_cache``; its width is the rotary dim.
    """
    ks = vs = kt = vt = None
    mode = capi.SCALE_NONE
    if loc is not None:
        if loc.dtype != torch.int64:
            loc = loc.to(torch.int64)
        fp8 = k_pool.dtype in (torch.uint8, torch.float8_e4m3fn, torch.float8_e5m2)
        if fp8 and (k_scale is not None or v_scale is not None):
            tensors = isinstance(k_scale, torch.Tensor) or isinstance(v_scale, torch.Tensor)
            mode = capi.SCALE_TRUE_DIV if tensors else capi.SCALE_RECIP_MUL
            if scale_mode is not None:
                mode = scale_mode
            if tensors:
                kt = _scale_tensor(k_scale, k.device)
                vt = _scale_tensor(v_scale, k.device)
            else:
                ks = 1.0 if k_scale is None else float(k_scale)
                vs = 1.0 if v_scale is None else float(v_scale)
    torch.ops.b200attn.rope_kv_store(q, k, v, positions, cos_sin_cache, bool(is_neox), loc, k_pool, v_pool,
                                     ks, vs, kt, vt, mode)


def _scale_tensor(s, device) -> torch.Tensor:
    if s is None:
        return torch.ones((), dtype=torch.float32, device=device)
    if isinstance(s, torch.Tensor):
        if s.dtype != torch.float32 or s.device != device:
            s = s.to(device=device, dtype=torch.float32)
        return s
    return torch.full((), float(s), dtype=torch.float32, device=device)


@dataclass
class ForwardMetadata:
    batch: int
    indices: torch.Tensor            # req_to_token, or kv_indices in CSR mode
    kv_indptr: Optional[torch.Tensor]
    kv_indices: Optional[torch.Tensor]
    # sliding-window layers (interleaved with full-attention layers, e.g. Gemma-3) use a second plan
    window_kv_indptr: Optional[torch.Tensor] = None
    window_kv_indices: Optional[torch.Tensor] = None


@dataclass
class ExtendMetadata:
    """Index arrays of one EXTEND / MIXED / TARGET_VERIFY / DRAFT_EXTEND forward (the extend half of
    the reference's ``ForwardMetadata``, triton_backend.py:25-38)."""
    batch: int
    qo_indptr: torch.Tensor
    kv_indptr: torch.Tensor
    kv_indices: Optional[torch.Tensor]        # None: prefix slots are read from req_to_token in place
    req_pool_indices: Optional[torch.Tensor]
    prefix_lens: Optional[torch.Tensor]
    custom_mask: Optional[torch.Tensor]
    mask_indptr: Optional[torch.Tensor]
    max_extend_len: Optional[int]             # None: bounded by the token count at forward time
    window_kv_indptr: Optional[torch.Tensor] = None
    avg_kv_len: int = 0                       # seq_lens_sum / batch when the host knows it (eager forwards): kernel choice only


class B200AttnBackend(AttentionBackend):
    def __init__(self, model_runner: Any, skip_prefill: bool = False, kv_indptr_buf: Optional[torch.Tensor] = None,
                 prefill_backend: Optional[AttentionBackend] = None,
                 materialize_kv_indices: bool = False, max_ctas: int = 0, min_chunks: int = 0,
                 return_lse: bool = False):
        """``skip_prefill`` / ``kv_indptr_buf``: the two constructor arguments the multi-step draft wrapper
        passes (triton_backend.py:41-61, 761-769): no extend-side buffers, and a caller-owned ``kv_indptr`` row
        of ``[max_bs * topk + 1]`` entries."""
        super().__init__()
        self.skip_prefill = skip_prefill
        capi.lib()  # fail here, loudly, if the CUDA library has not been built
        self.device = torch.device(model_runner.device)
        if self.device.type != "cuda":
            raise RuntimeError("B200AttnBackend needs a CUDA device (sm_100a); there is no CPU path")
        self.prefill_backend = prefill_backend
        self.req_to_token = model_runner.req_to_token_pool.req_to_token
        self.max_bs = model_runner.req_to_token_pool.size
        self.token_to_kv_pool = model_runner.token_to_kv_pool

        mc = model_runner.model_config
        try:  # inside sglang the attention TP size is process-global state
            from sglang.srt.layers.dp_attention import get_attention_tp_size

            tp = int(get_attention_tp_size())
        except Exception:
            tp = int(getattr(model_runner, "attention_tp_size", 1) or 1)
        self.num_head = mc.num_attention_heads // tp
        self.num_kv_head = mc.get_num_kv_heads(tp)
        k0 = self.token_to_kv_pool.get_key_buffer(0)
        v0 = self.token_to_kv_pool.get_value_buffer(0)
        self.head_dim = k0.shape[-1]
        self.v_head_dim = v0.shape[-1]
        self.kv_dtype = ops.dtype_code(k0.dtype)
        sw = getattr(model_runner, "sliding_window_size", None)
        self.sliding_window_size = int(sw) if sw is not None and sw > 0 else None
        # hybrid pool of sliding-window models (SWAKVPool, memory_pool.py:420-553): window layers read and write
        # the swa pool through full_to_swa_index_mapping (see _layer_pools)
        self.is_swa_pool = hasattr(self.token_to_kv_pool, "full_to_swa_index_mapping")
        if getattr(mc, "is_encoder_decoder", False):
            raise NotImplementedError("B200AttnBackend: cross attention is not supported yet")
        if self.head_dim != self.v_head_dim:
            raise NotImplementedError("B200AttnBackend: qk_head_dim != v_head_dim is not supported yet")
        # raises B200AttnError(E_UNSUPPORTED) for shapes without a kernel
        self.geometry = capi.decode_geometry(self.num_head, self.num_kv_head, self.head_dim, self.kv_dtype)

        self.materialize_kv_indices = materialize_kv_indices
        self.max_ctas = max_ctas
        self.min_chunks = min_chunks
        self.return_lse = return_lse
        self.max_context_len = getattr(mc, "context_len", self.req_to_token.shape[1])

        dev = self.device
        sa = getattr(model_runner, "server_args", None)
        self.num_draft_tokens = getattr(sa, "speculative_num_draft_tokens", None)
        self.speculative_num_steps = getattr(sa, "speculative_num_steps", None)
        self.speculative = getattr(sa, "speculative_algorithm", None) is not None or kv_indptr_buf is not None
        # Rows one decode launch may carry: a speculative draft step runs topk branches per request through the
        # CSR plan (kv_indptr.numel() - 1 = bs * topk rows, triton_backend.py:206-208), so every per-row buffer
        # is sized for that, not for req_to_token_pool.size alone.
        topk = int(getattr(sa, "speculative_eagle_topk", None) or 1)
        self.max_rows = self.max_bs * max(1, topk)
        if kv_indptr_buf is not None:
            self.max_rows = max(self.max_rows, kv_indptr_buf.numel() - 1)
        self.plan_buf = torch.zeros(ops.plan_bytes(self.max_rows), dtype=torch.uint8, device=dev)
        ws = ops.workspace_bytes(self.max_rows, self.num_head, self.num_kv_head, self.head_dim,
                                 self.v_head_dim, self.kv_dtype)
        self.workspace = torch.empty(ws // 4, dtype=torch.float32, device=dev)
        if kv_indptr_buf is None:
            self.kv_indptr = torch.zeros(self.max_rows + 1, dtype=torch.int32, device=dev)
        else:
            if kv_indptr_buf.dtype != torch.int32 or not kv_indptr_buf.is_contiguous():
                raise TypeError("kv_indptr_buf must be a contiguous int32 tensor")
            self.kv_indptr = kv_indptr_buf
        self.window_plan_buf = self.window_kv_indptr = None
        if self.sliding_window_size is not None:   # two sets of buffers, as triton_backend.py:77-90
            self.window_plan_buf = torch.zeros_like(self.plan_buf)
            self.window_kv_indptr = torch.zeros_like(self.kv_indptr)
        self.cuda_graph_window_kv_indices: Optional[torch.Tensor] = None
        self.lse = torch.empty(self.max_rows * self.num_head, dtype=torch.float32, device=dev) if return_lse else None
        self.cuda_graph_kv_indices: Optional[torch.Tensor] = None
        self.forward_metadata: Optional[ForwardMetadata] = None
        # extend (prefill) path: index buffers as triton_backend.py:92-99
        self.qo_indptr = self.extend_kv_indptr = self.extend_window_kv_indptr = self.mask_indptr = None
        if not skip_prefill:
            self.qo_indptr = torch.zeros(self.max_bs + 1, dtype=torch.int32, device=dev)
            self.extend_kv_indptr = torch.zeros(self.max_bs + 1, dtype=torch.int32, device=dev)
            self.extend_window_kv_indptr = torch.zeros_like(self.extend_kv_indptr) \
                if self.sliding_window_size is not None else None
            self.mask_indptr = torch.zeros(self.max_bs + 1, dtype=torch.int64, device=dev)
        self.extend_metadata: Optional[ExtendMetadata] = None
        # CUDA-graph state (init_cuda_graph_state): static copies of everything a captured kernel reads that the
        # runner hands over afresh on every replay (spec_info tensors, triton_backend.py:445-501, 575-619)
        self.cuda_graph_custom_mask: Optional[torch.Tensor] = None
        self.cg_req_pool_indices = self.cg_prefix_lens = self.cg_extend_lens = None
        self.cg_max_bs = self.cg_max_num_tokens = 0

    # ------------------------------------------------------------------ #
    # metadata
    # ------------------------------------------------------------------ #
    def _ensure_graph_kv_indices(self, need: int = 0) -> None:
        """``cuda_graph_kv_indices`` (triton_backend.py:357-364).  Decode reads req_to_token in place, so the
        buffer only exists where something consumes a flat list: materialised indices, or speculative CSR
        input.  Allocated by init_cuda_graph_state when that is known up front, else by the first capture
        hook that needs it (the hook runs before the capture begins, cuda_graph_runner.py:618)."""
        if self.cuda_graph_kv_indices is None:
            n = max(need, max(self.cg_max_num_tokens, 1) * self.max_context_len)
            self.cuda_graph_kv_indices = torch.zeros(n, dtype=torch.int32, device=self.device)

    def _plan(self, bs: int, req_pool_indices: torch.Tensor, seq_lens: torch.Tensor,
              seq_lens_sum: Optional[int], spec_info: Any, graph: bool) -> None:
        if bs > self.max_bs:
            raise ValueError(f"batch {bs} exceeds req_to_token_pool.size {self.max_bs}")
        if spec_info is not None:
            # speculative draft decode: the CSR lists come from spec_info (triton_backend.py:206-208)
            kv_indptr, kv_indices = spec_info.kv_indptr, spec_info.kv_indices
            rows = kv_indptr.numel() - 1
            if rows > self.max_rows:
                raise ValueError(f"speculative decode batch of {rows} rows exceeds the {self.max_rows} rows this backend "
                                 "was sized for (req_to_token_pool.size * speculative_eagle_topk)")
            if graph:
                # A captured decode kernel keeps the addresses it was launched with, and CudaGraphRunner hands the
                # replay hook a fresh spec_info every step (cuda_graph_runner.py:760-768): the lists are copied
                # into the backend's static buffers, as the reference does (triton_backend.py:574-576).
                self._ensure_graph_kv_indices(kv_indices.numel())
                if kv_indices.numel() > self.cuda_graph_kv_indices.numel():
                    raise ValueError("spec_info.kv_indices is longer than cuda_graph_kv_indices")
                if kv_indptr.data_ptr() != self.kv_indptr.data_ptr():
                    self.kv_indptr[: rows + 1].copy_(kv_indptr)
                if kv_indices.data_ptr() != self.cuda_graph_kv_indices.data_ptr():
                    self.cuda_graph_kv_indices[: kv_indices.numel()].copy_(kv_indices)
                kv_indptr, kv_indices = self.kv_indptr[: rows + 1], self.cuda_graph_kv_indices
            torch.ops.b200attn.plan_csr(kv_indptr, self.plan_buf, self.num_head, self.num_kv_head,
                                        self.head_dim, self.kv_dtype, self.max_ctas, self.min_chunks)
            self.forward_metadata = ForwardMetadata(rows, kv_indices, kv_indptr, kv_indices)
            return
        kv_indices = window_kv_indices = None
        if self.materialize_kv_indices:
            if graph:
                kv_indices = self.cuda_graph_kv_indices
                window_kv_indices = self.cuda_graph_window_kv_indices
            else:
                if seq_lens_sum is None:
                    seq_lens_sum = int(seq_lens[:bs].sum().item())
                kv_indices = torch.empty(seq_lens_sum, dtype=torch.int32, device=self.device)
                if self.sliding_window_size is not None:
                    n = min(seq_lens_sum, bs * (self.sliding_window_size + 1))
                    window_kv_indices = torch.empty(n, dtype=torch.int32, device=self.device)
        torch.ops.b200attn.plan(self.req_to_token, req_pool_indices[:bs], seq_lens[:bs], self.plan_buf,
                                self.kv_indptr, kv_indices, self.num_head, self.num_kv_head,
                                self.head_dim, self.kv_dtype, self.max_ctas, self.min_chunks, 0)
        md = ForwardMetadata(bs, self.req_to_token, self.kv_indptr[: bs + 1], kv_indices)
        if self.sliding_window_size is not None:
            # second plan over the last min(S, W + 1) tokens (update_sliding_window_buffer, triton_backend.py:927-956)
            torch.ops.b200attn.plan(self.req_to_token, req_pool_indices[:bs], seq_lens[:bs], self.window_plan_buf,
                                    self.window_kv_indptr, window_kv_indices, self.num_head, self.num_kv_head,
                                    self.head_dim, self.kv_dtype, self.max_ctas, self.min_chunks,
                                    self.sliding_window_size)
            md.window_kv_indptr = self.window_kv_indptr[: bs + 1]
            md.window_kv_indices = window_kv_indices
        self.forward_metadata = md

    def init_forward_metadata(self, forward_batch):
        mode = forward_batch.forward_mode
        if mode.is_decode_or_idle():
            if forward_batch.batch_size == 0:
                self.forward_metadata = ForwardMetadata(0, self.req_to_token, self.kv_indptr[:1], None)
                return
            self._plan(forward_batch.batch_size, forward_batch.req_pool_indices, forward_batch.seq_lens,
                       forward_batch.seq_lens_sum, forward_batch.spec_info, graph=False)
        elif self.prefill_backend is not None:
            self.prefill_backend.init_forward_metadata(forward_batch)
        else:
            self._plan_extend(forward_batch.batch_size, forward_batch.req_pool_indices, forward_batch.seq_lens,
                              mode, forward_batch.spec_info, forward_batch)

    def _plan_extend(self, bs: int, req_pool_indices: torch.Tensor, seq_lens: torch.Tensor, mode,
                     spec_info: Any, forward_batch: Any = None, graph: Optional[str] = None) -> None:
        """Metadata of the native extend path: what ``TritonAttnBackend.init_forward_metadata`` builds
        for its three extend modes (triton_backend.py:227-318), in one launch and without
        materialising ``kv_indices`` for the cached prefix.  ``graph`` = "capture" / "replay": every array
        a captured kernel reads lives in a buffer this backend owns and is refilled here
        (triton_backend.py:445-501, 579-623)."""
        if self.skip_prefill:
            raise RuntimeError("B200AttnBackend was built with skip_prefill=True: no extend path")
        if bs > self.max_bs:
            raise ValueError(f"batch {bs} exceeds req_to_token_pool.size {self.max_bs}")
        if bs == 0:
            self.extend_metadata = None
            return
        qo, kv, win = self.qo_indptr[: bs + 1], self.extend_kv_indptr[: bs + 1], None
        rpi = req_pool_indices[:bs]
        # host-side length hint (forward_batch.seq_lens_sum, forward_batch_info.py:163-311); a captured graph
        # has none -- its kernel choice is fixed at capture
        sls = getattr(forward_batch, "seq_lens_sum", None) if forward_batch is not None else None
        avg = int(sls) // bs if sls else 0
        if graph is not None:
            if self.cg_req_pool_indices is None or bs > self.cg_req_pool_indices.numel():
                raise RuntimeError("init_cuda_graph_state was not run (or max_bs was smaller than this batch)")
            # static int64 copies: copy_ converts the runner's int32 buffers without allocating
            self.cg_req_pool_indices[:bs].copy_(rpi)
            self.cg_prefix_lens[:bs].copy_(seq_lens[:bs])
            rpi, pl_static = self.cg_req_pool_indices[:bs], self.cg_prefix_lens[:bs]
        if mode.is_target_verify():
            # prefix = the whole committed sequence, extend = num_draft_tokens per request, visibility
            # inside the draft tree from spec_info.custom_mask (:227-263)
            n = int(getattr(spec_info, "draft_token_num", None) or self.num_draft_tokens)
            mask = spec_info.custom_mask
            if graph is not None:
                pl = pl_static
                if mask is not None:
                    if mask.numel() > self.cuda_graph_custom_mask.numel():
                        raise ValueError("spec_info.custom_mask is longer than cuda_graph_custom_mask")
                    self.cuda_graph_custom_mask[: mask.numel()].copy_(mask.reshape(-1))   # :467-468, :602-603
                    mask = self.cuda_graph_custom_mask
            else:
                pl = seq_lens[:bs]
                if pl.dtype != rpi.dtype:
                    pl = pl.to(rpi.dtype)
            mi = self.mask_indptr[: bs + 1]
            torch.ops.b200attn.extend_plan(None, pl, qo, kv, mi, n, 0)
            self.extend_metadata = ExtendMetadata(bs, qo, kv, None, rpi, pl, mask, mi, n, None, avg)
        elif mode.is_draft_extend():
            if graph is None:
                kv_indices, kv_indptr, qo_indptr, custom_mask = spec_info.generate_attn_arg_prefill(
                    req_pool_indices, seq_lens, None, self.req_to_token)   # :265-279
                self.extend_metadata = ExtendMetadata(bs, qo_indptr, kv_indptr, kv_indices, None, None,
                                                      custom_mask, None, None, None, avg)
                return
            # Under a graph generate_attn_arg_prefill cannot be used (it allocates): the same lists are built
            # into static buffers.  Prefix list = req_to_token[req, :seq_len] read in place (the reference
            # flattens it into cuda_graph_kv_indices, :486-497, :612-623); extend tokens per request =
            # speculative_num_steps + 1 at capture (:477-485), spec_info.accept_length at replay (:608-611).
            per = int(self.speculative_num_steps or 0) + 1
            if graph == "capture":
                torch.ops.b200attn.extend_plan(None, pl_static, qo, kv, None, per, 0)
            else:
                self.cg_extend_lens[:bs].copy_(spec_info.accept_length[:bs])
                torch.ops.b200attn.extend_plan(self.cg_extend_lens[:bs], pl_static, qo, kv, None, 0, 0)
            self.extend_metadata = ExtendMetadata(bs, qo, kv, None, rpi, pl_static, None, None, per, None, avg)
        else:
            if graph is not None:
                raise ValueError(f"Invalid forward mode for CUDA graph capture / replay: {mode!r}")   # :504-507
            pl, el = forward_batch.extend_prefix_lens[:bs], forward_batch.extend_seq_lens[:bs]
            if pl.dtype != rpi.dtype:
                pl = pl.to(rpi.dtype)
            if el.dtype != pl.dtype:
                el = el.to(pl.dtype)
            torch.ops.b200attn.extend_plan(el, pl, qo, kv, None, 0, 0)
            if self.sliding_window_size is not None:
                win = self.extend_window_kv_indptr[: bs + 1]
                torch.ops.b200attn.extend_plan(el, pl, qo, win, None, 0, self.sliding_window_size)
            cpu = getattr(forward_batch, "extend_seq_lens_cpu", None)
            self.extend_metadata = ExtendMetadata(bs, qo, kv, None, rpi, pl, None, None,
                                                  max(cpu) if cpu else None, win, avg)

    def init_cuda_graph_state(self, max_bs: int, max_num_tokens: int, kv_indices_buf: Optional[torch.Tensor] = None):
        """Static buffers of the CUDA-graph path (triton_backend.py:338-388).  The decode kernel reads
        req_to_token in place and its plan / workspace already have fixed addresses, so only the flat index
        list (speculative CSR input, materialised indices) and the extend-side spec buffers are added."""
        self.cg_max_bs, self.cg_max_num_tokens = max_bs, max_num_tokens
        if kv_indices_buf is not None:
            if kv_indices_buf.dtype != torch.int32 or not kv_indices_buf.is_contiguous():
                raise TypeError("kv_indices_buf must be a contiguous int32 tensor")
            self.cuda_graph_kv_indices = kv_indices_buf                       # :363-364
        elif self.materialize_kv_indices or self.speculative:
            self.cuda_graph_kv_indices = torch.zeros(max_num_tokens * self.max_context_len,
                                                     dtype=torch.int32, device=self.device)
        if self.materialize_kv_indices and self.sliding_window_size is not None:   # :373-380
            self.cuda_graph_window_kv_indices = torch.zeros(
                max_num_tokens * (self.sliding_window_size + 1), dtype=torch.int32, device=self.device)
        if not self.skip_prefill and self.prefill_backend is None:
            if self.speculative or self.num_draft_tokens:
                self.cuda_graph_custom_mask = torch.zeros(max_num_tokens * self.max_context_len,
                                                          dtype=torch.uint8, device=self.device)    # :366-371
            n = max(max_bs, 1)
            self.cg_req_pool_indices = torch.zeros(n, dtype=torch.int64, device=self.device)
            self.cg_prefix_lens = torch.zeros(n, dtype=torch.int64, device=self.device)
            self.cg_extend_lens = torch.zeros(n, dtype=torch.int64, device=self.device)
        if self.prefill_backend is not None and hasattr(self.prefill_backend, "init_cuda_graph_state"):
            try:
                self.prefill_backend.init_cuda_graph_state(max_bs, max_num_tokens)
            except NotImplementedError:
                pass

    def init_forward_metadata_capture_cuda_graph(self, bs, num_tokens, req_pool_indices, seq_lens,
                                                 encoder_lens, forward_mode, spec_info):
        assert encoder_lens is None, "cross attention is not supported"
        if not forward_mode.is_decode_or_idle():
            if self.prefill_backend is not None:
                return self.prefill_backend.init_forward_metadata_capture_cuda_graph(
                    bs, num_tokens, req_pool_indices, seq_lens, encoder_lens, forward_mode, spec_info)
            if not (forward_mode.is_target_verify() or forward_mode.is_draft_extend()):
                raise ValueError(f"Invalid forward mode for CUDA graph capture: {forward_mode!r}")  # triton_backend.py:504-507
            if forward_mode.is_target_verify() and self.cuda_graph_custom_mask is None \
                    and getattr(spec_info, "custom_mask", None) is not None:
                # not announced through server_args: allocate now, before the capture begins
                self.cuda_graph_custom_mask = torch.zeros(max(self.cg_max_num_tokens, num_tokens) * self.max_context_len,
                                                          dtype=torch.uint8, device=self.device)
            return self._plan_extend(bs, req_pool_indices, seq_lens, forward_mode, spec_info, graph="capture")
        self._plan(bs, req_pool_indices, seq_lens, None, spec_info, graph=True)

    def init_forward_metadata_replay_cuda_graph(self, bs, req_pool_indices, seq_lens, seq_lens_sum,
                                                encoder_lens, forward_mode, spec_info, seq_lens_cpu):
        if not forward_mode.is_decode_or_idle():
            if self.prefill_backend is not None:
                return self.prefill_backend.init_forward_metadata_replay_cuda_graph(
                    bs, req_pool_indices, seq_lens, seq_lens_sum, encoder_lens, forward_mode, spec_info,
                    seq_lens_cpu)
            if not (forward_mode.is_target_verify() or forward_mode.is_draft_extend()):
                raise ValueError(f"Invalid forward mode for CUDA graph replay: {forward_mode!r}")   # :624-627
            return self._plan_extend(bs, req_pool_indices, seq_lens, forward_mode, spec_info, graph="replay")
        self._plan(bs, req_pool_indices, seq_lens, seq_lens_sum, spec_info, graph=True)

    def get_cuda_graph_seq_len_fill_value(self):
        return 1

    # ------------------------------------------------------------------ #
    # forward
    # ------------------------------------------------------------------ #
    def forward_decode(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True):
        if getattr(layer, "is_cross_attention", False):
            raise NotImplementedError("B200AttnBackend: cross attention is not supported")
        sw = getattr(layer, "sliding_window_size", -1)
        windowed = sw is not None and sw > -1   # triton_backend.py:711-716
        if windowed and self.window_plan_buf is None:
            raise RuntimeError("B200AttnBackend: layer has a sliding window but model_runner.sliding_window_size is unset")
        if windowed and self.forward_metadata is not None and self.forward_metadata.window_kv_indptr is None \
                and self.forward_metadata.batch > 0:
            raise NotImplementedError("B200AttnBackend: sliding-window layers with speculative (CSR) metadata")
        if layer.qk_head_dim != layer.v_head_dim:
            raise NotImplementedError("B200AttnBackend: qk_head_dim != v_head_dim is not supported")

        q = q.reshape(-1, layer.tp_q_head_num * layer.qk_head_dim)
        o = torch.empty_like(q)
        k_pool, v_pool, slot_map = self._layer_pools(forward_batch.token_to_kv_pool, layer)
        if save_kv_cache and k is not None:
            loc = forward_batch.out_cache_loc
            if slot_map is not None:
                loc = slot_map[loc]                       # translate_loc_from_full_to_swa (memory_pool.py:531-533)
            store_kv(k_pool, v_pool, loc,
                     k.view(-1, layer.tp_k_head_num, layer.qk_head_dim),
                     v.view(-1, layer.tp_v_head_num, layer.v_head_dim),
                     getattr(layer, "k_scale", None), getattr(layer, "v_scale", None))
        md = self.forward_metadata
        bs = q.shape[0]
        if bs == 0:
            return o
        if md is None or md.batch != bs:
            raise RuntimeError("B200AttnBackend.forward_decode: init_forward_metadata was not run for this batch")
        k_scale = getattr(layer, "k_scale_float", None)
        v_scale = getattr(layer, "v_scale_float", None)
        fp8 = self.kv_dtype in (capi.FP8_E4M3, capi.FP8_E5M2)
        torch.ops.b200attn.decode(
            q.view(bs, layer.tp_q_head_num, layer.qk_head_dim), k_pool, v_pool,
            o.view(bs, layer.tp_q_head_num, layer.v_head_dim),
            self.lse[: bs * self.num_head] if self.lse is not None else None,
            self.window_plan_buf if windowed else self.plan_buf,
            md.indices, self.workspace, float(layer.scaling),
            float(getattr(layer, "logit_cap", 0.0) or 0.0),
            float(k_scale) if (fp8 and k_scale is not None) else 1.0,
            float(v_scale) if (fp8 and v_scale is not None) else 1.0, 0, slot_map)
        return o

    def _layer_pools(self, pool, layer):
        """(k_pool, v_pool, slot_map) of a layer.  ``MHATokenToKVPool``: the layer's buffers, no map.
        ``SWAKVPool``: a sliding-window layer lives in ``swa_kv_pool`` and every full-pool slot id -- the ones in
        ``req_to_token`` and ``out_cache_loc`` -- goes through ``full_to_swa_index_mapping`` first (the reference
        translates a materialised ``kv_indices`` list, flashinfer_backend.py:754, 1001; the decode kernel applies
        the map to the ids it reads in place instead)."""
        smap = None
        if not self.is_swa_pool:
            start = getattr(pool, "start_layer", 0) or 0
            kb, vb = pool.k_buffer[layer.layer_id - start], pool.v_buffer[layer.layer_id - start]
        else:
            pid, is_swa = pool.layers_mapping[layer.layer_id]
            sub = pool.swa_kv_pool if is_swa else pool.full_kv_pool
            start = getattr(sub, "start_layer", 0) or 0
            smap = pool.full_to_swa_index_mapping if is_swa else None
            kb, vb = sub.k_buffer[pid - start], sub.v_buffer[pid - start]
        if self.kv_dtype == capi.FP8_E5M2 and kb.dtype == torch.uint8:
            # fp8 pools are stored as uint8 (memory_pool.py:114-118): the e5m2 kind travels with the view
            kb, vb = kb.view(torch.float8_e5m2), vb.view(torch.float8_e5m2)
        return kb, vb, smap

    def last_lse(self, bs: int) -> torch.Tensor:
        """ln sum exp(scaled logits) of the most recent forward_decode, ``[bs, heads]``."""
        if self.lse is None:
            raise RuntimeError("construct the backend with return_lse=True")
        return self.lse[: bs * self.num_head].view(bs, self.num_head)

    def forward_extend(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True):
        """Native extend / prefill attention (``TritonAttnBackend.forward_extend``,
        triton_backend.py:632-685): KV store, then one ``extend`` launch over cached prefix + new
        tokens.  With a ``prefill_backend`` the call is delegated instead."""
        if self.prefill_backend is not None:
            return self.prefill_backend.forward_extend(q, k, v, layer, forward_batch, save_kv_cache=save_kv_cache)
        if getattr(layer, "is_cross_attention", False):
            raise NotImplementedError("B200AttnBackend: cross attention is not supported")
        if layer.qk_head_dim != layer.v_head_dim:
            raise NotImplementedError("B200AttnBackend: qk_head_dim != v_head_dim is not supported")
        fp8 = self.kv_dtype in (capi.FP8_E4M3, capi.FP8_E5M2)
        q = q.reshape(-1, layer.tp_q_head_num * layer.qk_head_dim)
        o = torch.empty_like(q)
        k_pool, v_pool, slot_map = self._layer_pools(forward_batch.token_to_kv_pool, layer)
        loc = forward_batch.out_cache_loc
        if loc.dtype != torch.int64:
            loc = loc.to(torch.int64)
        if slot_map is not None:
            loc = slot_map[loc]
        if save_kv_cache and k is not None:
            store_kv(k_pool, v_pool, loc, k.view(-1, layer.tp_k_head_num, layer.qk_head_dim),
                     v.view(-1, layer.tp_v_head_num, layer.v_head_dim),
                     getattr(layer, "k_scale", None), getattr(layer, "v_scale", None))
        md = self.extend_metadata
        tokens = q.shape[0]
        if tokens == 0:
            return o
        if md is None:
            raise RuntimeError("B200AttnBackend.forward_extend: init_forward_metadata was not run for this batch")
        causal = getattr(getattr(layer, "attn_type", None), "name", "DECODER") != "ENCODER_ONLY"
        sw = getattr(layer, "sliding_window_size", -1)
        windowed = sw is not None and sw > -1
        kv_indptr = md.kv_indptr
        if windowed:
            if md.window_kv_indptr is None:
                raise NotImplementedError("B200AttnBackend: sliding-window layers need model_runner.sliding_window_size "
                                          "and a plain EXTEND batch")
            kv_indptr = md.window_kv_indptr
        kv_indices = md.kv_indices
        if slot_map is not None:
            # SWAKVPool: the prefix list of a window layer as swa-pool slots (eager prefill only: one host sync for
            # the list length, as the reference's own extend metadata does, triton_backend.py:283-300)
            kv_indices = self._swa_prefix_indices(md, kv_indptr, slot_map, windowed)
        max_len = tokens if md.max_extend_len is None else min(md.max_extend_len, tokens)
        k_scale = getattr(layer, "k_scale_float", None)
        v_scale = getattr(layer, "v_scale_float", None)
        torch.ops.b200attn.extend(
            q.view(tokens, layer.tp_q_head_num, layer.qk_head_dim),
            o.view(tokens, layer.tp_q_head_num, layer.v_head_dim),
            self.lse_extend(tokens) if self.return_lse else None,
            k_pool, v_pool, md.qo_indptr, kv_indptr, kv_indices,
            self.req_to_token if kv_indices is None else None, md.req_pool_indices, md.prefix_lens,
            loc, md.custom_mask, md.mask_indptr, int(max_len), causal,
            int(sw) if windowed else -1, True, float(layer.scaling),
            float(getattr(layer, "logit_cap", 0.0) or 0.0),
            float(k_scale) if (fp8 and k_scale is not None) else 1.0,
            float(v_scale) if (fp8 and v_scale is not None) else 1.0, int(md.avg_kv_len))
        return o

    def _swa_prefix_indices(self, md, kv_indptr: torch.Tensor, slot_map: torch.Tensor, windowed: bool) -> torch.Tensor:
        """Flat prefix slot list of the batch (request b: entries kv_indptr[b] .. kv_indptr[b + 1]) translated to
        swa-pool slots.  Window layers keep the LAST kv_indptr[b+1] - kv_indptr[b] prefix tokens of a request."""
        if md.kv_indices is not None:
            return slot_map[md.kv_indices.long()].to(torch.int32)
        if md.req_pool_indices is None or md.prefix_lens is None:
            raise NotImplementedError("B200AttnBackend: SWAKVPool with this extend metadata")
        lens = (kv_indptr[1:] - kv_indptr[:-1]).long()
        total = int(kv_indptr[-1].item())
        if total == 0:
            return torch.zeros(1, dtype=torch.int32, device=self.device)
        first = md.prefix_lens.long() - lens if windowed else torch.zeros_like(lens)
        pos = torch.arange(total, device=self.device) - torch.repeat_interleave(kv_indptr[:-1].long() - first, lens)
        rows = torch.repeat_interleave(md.req_pool_indices.long(), lens)
        return slot_map[self.req_to_token[rows, pos].long()].to(torch.int32)

    def lse_extend(self, tokens: int) -> torch.Tensor:
        """fp32 ``[tokens, heads]`` buffer the extend kernel writes ln sum exp(scaled logits) into."""
        buf = getattr(self, "_lse_extend", None)
        if buf is None or buf.numel() < tokens * self.num_head:
            buf = torch.empty(tokens * self.num_head, dtype=torch.float32, device=self.device)
            self._lse_extend = buf
        return buf[: tokens * self.num_head]

    def support_triton(self):
        return True


class B200MultiStepDraftBackend:
    """``speculative_num_steps`` decode backends behind one object, for the consecutive draft-decode steps
    of EAGLE-style speculative decoding: the drop-in for ``TritonMultiStepDraftBackend``
    (triton_backend.py:735-872) -- same constructor, same four hooks, same ``attn_backends`` list the
    draft worker indexes per step.

    All steps share one ``kv_indptr`` ``[steps, max_bs * topk + 1]`` and one ``kv_indices`` ``[steps, n]``
    buffer, filled by ONE ``draft_kv_indices`` launch per draft forward (the reference launches the Triton
    kernel ``generate_draft_decode_kv_indices`` with the same grid, eagle_utils.py:844-920); step i's backend
    then plans over its row through the CSR input mode.
    """

    def __init__(self, model_runner: Any, topk: int, speculative_num_steps: int, **backend_kwargs):
        self.topk = int(topk)
        self.speculative_num_steps = int(speculative_num_steps)
        self.device = torch.device(model_runner.device)
        max_bs = model_runner.req_to_token_pool.size * self.topk
        self.kv_indptr = torch.zeros((self.speculative_num_steps, max_bs + 1), dtype=torch.int32, device=self.device)
        self.attn_backends = [
            B200AttnBackend(model_runner, skip_prefill=True, kv_indptr_buf=self.kv_indptr[i], **backend_kwargs)
            for i in range(self.speculative_num_steps)
        ]
        self.max_context_len = self.attn_backends[0].max_context_len
        self.num_head = self.attn_backends[0].num_head
        self.pool_len = model_runner.req_to_token_pool.req_to_token.shape[1]
        sa = getattr(model_runner, "server_args", None)
        self.page_size = int(getattr(sa, "page_size", None) or getattr(model_runner, "page_size", None) or 1)
        self.cuda_graph_kv_indices: Optional[torch.Tensor] = None

    def common_template(self, forward_batch, kv_indices_buffer: torch.Tensor, call_fn) -> None:
        num_seqs = forward_batch.batch_size
        bs = self.topk * num_seqs
        seq_lens_sum = forward_batch.seq_lens_sum
        torch.ops.b200attn.draft_kv_indices(
            forward_batch.req_pool_indices[:num_seqs], forward_batch.req_to_token_pool.req_to_token,
            forward_batch.seq_lens[:num_seqs], forward_batch.positions, kv_indices_buffer, self.kv_indptr,
            self.topk, self.page_size)
        for i in range(self.speculative_num_steps):
            forward_batch.spec_info.kv_indptr = self.kv_indptr[i, : bs + 1]
            # seq_lens_sum is -1 on graph replays (the reference slices with it all the same, :806-808, which
            # leaves a harmlessly short view); the whole row is handed on instead -- kv_indptr bounds what is read
            n = seq_lens_sum * self.topk + bs * (i + 1) if seq_lens_sum is not None and seq_lens_sum >= 0 else None
            forward_batch.spec_info.kv_indices = kv_indices_buffer[i] if n is None else kv_indices_buffer[i][:n]
            call_fn(i, forward_batch)

    def init_forward_metadata(self, forward_batch) -> None:
        kv_indices = torch.empty(
            (self.speculative_num_steps, forward_batch.batch_size * self.topk * self.max_context_len),
            dtype=torch.int32, device=self.device)

        def call_fn(i, fb):
            # no clone needed (the reference clones, :822-827): every step owns its row of both buffers
            self.attn_backends[i].init_forward_metadata(fb)

        self.common_template(forward_batch, kv_indices, call_fn)

    def init_cuda_graph_state(self, max_bs: int, max_num_tokens: int) -> None:
        self.cuda_graph_kv_indices = torch.zeros(
            (self.speculative_num_steps, max_num_tokens * self.max_context_len), dtype=torch.int32, device=self.device)
        for i in range(self.speculative_num_steps):
            self.attn_backends[i].init_cuda_graph_state(max_bs, max_num_tokens,
                                                        kv_indices_buf=self.cuda_graph_kv_indices[i])

    def init_forward_metadata_capture_cuda_graph(self, forward_batch) -> None:
        def call_fn(i, fb):
            self.attn_backends[i].init_forward_metadata_capture_cuda_graph(
                fb.batch_size, fb.batch_size * self.topk, fb.req_pool_indices, fb.seq_lens,
                encoder_lens=None, forward_mode=ForwardMode.DECODE, spec_info=fb.spec_info)

        self.common_template(forward_batch, self.cuda_graph_kv_indices, call_fn)

    def init_forward_metadata_replay_cuda_graph(self, forward_batch, bs: int) -> None:
        def call_fn(i, fb):
            self.attn_backends[i].init_forward_metadata_replay_cuda_graph(
                bs, fb.req_pool_indices, fb.seq_lens, seq_lens_sum=-1, encoder_lens=None,
                forward_mode=ForwardMode.DECODE, spec_info=fb.spec_info, seq_lens_cpu=None)

        self.common_template(forward_batch, self.cuda_graph_kv_indices, call_fn)
