"""Host-side mirror of the reference interfaces the decode path plugs into.

Only the *interface* is restated (names, argument meaning, layouts) so the
backend can run and be tested without sglang installed; inside a real sglang
process the backend receives sglang's own objects, which have the same
attributes.  References (paths relative to the reference tree):

* ``AttentionBackend``   python/sglang/srt/layers/attention/base_attn_backend.py:14-117
* ``ForwardMode``        python/sglang/srt/model_executor/forward_batch_info.py:67-137
* ``ForwardBatch``       python/sglang/srt/model_executor/forward_batch_info.py:163-311
* ``ReqToTokenPool``     python/sglang/srt/mem_cache/memory_pool.py:47-95
* ``MHATokenToKVPool``   python/sglang/srt/mem_cache/memory_pool.py:162-417
* ``RadixAttention``     python/sglang/srt/layers/radix_attention.py:39-110
"""
from __future__ import annotations

import enum
from abc import ABC, abstractmethod
from dataclasses import dataclass
from typing import Any, List, Optional, Union

import torch


class ForwardMode(enum.IntEnum):
    EXTEND = enum.auto()
    DECODE = enum.auto()
    MIXED = enum.auto()
    IDLE = enum.auto()
    TARGET_VERIFY = enum.auto()
    DRAFT_EXTEND = enum.auto()
    DUMMY_FIRST = enum.auto()

    def is_decode(self):
        return self == ForwardMode.DECODE

    def is_idle(self):
        return self == ForwardMode.IDLE

    def is_decode_or_idle(self):
        return self in (ForwardMode.DECODE, ForwardMode.IDLE)

    def is_extend(self):
        return self in (ForwardMode.EXTEND, ForwardMode.MIXED, ForwardMode.DRAFT_EXTEND,
                        ForwardMode.TARGET_VERIFY)

    def is_target_verify(self):
        return self == ForwardMode.TARGET_VERIFY

    def is_draft_extend(self):
        return self == ForwardMode.DRAFT_EXTEND


class AttentionType(enum.Enum):
    """layers/radix_attention.py:25-36."""
    DECODER = "decoder"
    ENCODER_ONLY = "encoder_only"


class ReqToTokenPool:
    """request -> token-slot map: int32 ``[size, max_context_len]``."""

    def __init__(self, size: int, max_context_len: int, device: Union[str, torch.device]):
        self.size = size
        self.max_context_len = max_context_len
        self.device = device
        self.req_to_token = torch.zeros((size, max_context_len), dtype=torch.int32, device=device)
        self.free_slots = list(range(size))

    def write(self, indices, values):
        self.req_to_token[indices] = values

    def available_size(self):
        return len(self.free_slots)

    def alloc(self, need_size: int) -> Optional[List[int]]:
        if need_size > len(self.free_slots):
            return None
        out, self.free_slots = self.free_slots[:need_size], self.free_slots[need_size:]
        return out

    def free(self, free_index):
        if isinstance(free_index, int):
            self.free_slots.append(free_index)
        else:
            self.free_slots.extend(free_index)

    def clear(self):
        self.free_slots = list(range(self.size))


class MHATokenToKVPool:
    """Per-layer K and V pools ``[size + page_size, head_num, head_dim]``.

    fp8 pools are stored as ``uint8`` and viewed as ``float8_e4m3fn`` by the
    getters; slots ``0 .. page_size-1`` are the dummy page padded tokens write
    to.  ``set_kv_buffer`` has the reference's signature but runs the sm_100a
    ``kv_store`` kernel (one launch for K and V, fused divide + fp8 cast).
    """

    def __init__(self, size: int, page_size: int, dtype: torch.dtype, head_num: int, head_dim: int,
                 layer_num: int, device: Union[str, torch.device], start_layer: Optional[int] = None,
                 v_head_dim: Optional[int] = None):
        self.size = size
        self.page_size = page_size
        self.dtype = dtype
        self.device = device
        self.store_dtype = torch.uint8 if dtype in (torch.float8_e5m2, torch.float8_e4m3fn) else dtype
        self.layer_num = layer_num
        self.start_layer = start_layer or 0
        self.head_num = head_num
        self.head_dim = head_dim
        self.v_head_dim = head_dim if v_head_dim is None else v_head_dim
        rows = size + page_size
        self.k_buffer = [torch.zeros((rows, head_num, head_dim), dtype=self.store_dtype, device=device)
                         for _ in range(layer_num)]
        self.v_buffer = [torch.zeros((rows, head_num, self.v_head_dim), dtype=self.store_dtype, device=device)
                         for _ in range(layer_num)]
        self.layer_transfer_counter = None

    def _view(self, t: torch.Tensor) -> torch.Tensor:
        return t.view(self.dtype) if self.store_dtype != self.dtype else t

    def get_key_buffer(self, layer_id: int):
        return self._view(self.k_buffer[layer_id - self.start_layer])

    def get_value_buffer(self, layer_id: int):
        return self._view(self.v_buffer[layer_id - self.start_layer])

    def get_kv_buffer(self, layer_id: int):
        return self.get_key_buffer(layer_id), self.get_value_buffer(layer_id)

    def set_kv_buffer(self, layer, loc: torch.Tensor, cache_k: torch.Tensor, cache_v: torch.Tensor,
                      k_scale=None, v_scale=None, layer_id_override: Optional[int] = None):
        from .backend import store_kv  # local import: the ops need the built library

        layer_id = layer_id_override if layer_id_override is not None else layer.layer_id
        kb, vb = self.k_buffer[layer_id - self.start_layer], self.v_buffer[layer_id - self.start_layer]
        if self.dtype == torch.float8_e5m2:
            kb, vb = kb.view(self.dtype), vb.view(self.dtype)
        store_kv(kb, vb, loc, cache_k, cache_v, k_scale, v_scale)


class SWAKVPool:
    """Hybrid pool of sliding-window models (``SWAKVPool``, mem_cache/memory_pool.py:420-553): full-attention
    layers live in ``full_kv_pool``, sliding-window layers in the smaller ``swa_kv_pool`` whose slot of a token
    is ``full_to_swa_index_mapping[full slot]`` (filled by the allocator, mem_cache/allocator.py).  Same
    attribute and method names as the reference class."""

    def __init__(self, size: int, size_swa: int, dtype: torch.dtype, head_num: int, head_dim: int,
                 swa_attention_layer_ids, full_attention_layer_ids, enable_kvcache_transpose: bool = False,
                 device: Union[str, torch.device] = "cuda"):
        assert not enable_kvcache_transpose
        self.size, self.size_swa, self.dtype, self.device = size, size_swa, dtype, device
        self.page_size = 1
        self.swa_layer_nums, self.full_layer_nums = len(swa_attention_layer_ids), len(full_attention_layer_ids)
        self.swa_kv_pool = MHATokenToKVPool(size_swa, 1, dtype, head_num, head_dim, self.swa_layer_nums, device)
        self.full_kv_pool = MHATokenToKVPool(size, 1, dtype, head_num, head_dim, self.full_layer_nums, device)
        self.layers_mapping = {}
        for i, gid in enumerate(full_attention_layer_ids):
            self.layers_mapping[gid] = (i, False)
        for i, gid in enumerate(swa_attention_layer_ids):
            self.layers_mapping[gid] = (i, True)
        self.full_to_swa_index_mapping: Optional[torch.Tensor] = None

    def _pool(self, layer_id: int):
        pid, is_swa = self.layers_mapping[layer_id]
        return (self.swa_kv_pool if is_swa else self.full_kv_pool), pid

    def get_key_buffer(self, layer_id: int):
        pool, pid = self._pool(layer_id)
        return pool.get_key_buffer(pid)

    def get_value_buffer(self, layer_id: int):
        pool, pid = self._pool(layer_id)
        return pool.get_value_buffer(pid)

    def get_kv_buffer(self, layer_id: int):
        return self.get_key_buffer(layer_id), self.get_value_buffer(layer_id)

    def translate_loc_from_full_to_swa(self, kv_indices: torch.Tensor):
        assert self.full_to_swa_index_mapping is not None
        return self.full_to_swa_index_mapping[kv_indices].to(torch.int32)

    def set_kv_buffer(self, layer, loc: torch.Tensor, cache_k: torch.Tensor, cache_v: torch.Tensor,
                      k_scale=None, v_scale=None):
        pool, pid = self._pool(layer.layer_id)
        if self.layers_mapping[layer.layer_id][1] and self.full_to_swa_index_mapping is not None:
            loc = self.translate_loc_from_full_to_swa(loc)
        pool.set_kv_buffer(None, loc, cache_k, cache_v, k_scale, v_scale, layer_id_override=pid)


class AttnLayer:
    """The attributes of ``RadixAttention`` the attention backends read."""

    def __init__(self, num_heads: int, head_dim: int, scaling: float, num_kv_heads: int, layer_id: int,
                 logit_cap: float = 0.0, v_head_dim: int = -1, sliding_window_size: int = -1,
                 is_cross_attention: bool = False, attn_type: AttentionType = AttentionType.DECODER):
        self.attn_type = attn_type
        self.tp_q_head_num = num_heads
        self.tp_k_head_num = num_kv_heads
        self.tp_v_head_num = num_kv_heads
        self.head_dim = head_dim
        self.qk_head_dim = head_dim
        self.v_head_dim = v_head_dim if v_head_dim != -1 else head_dim
        self.scaling = scaling
        self.layer_id = layer_id
        self.logit_cap = logit_cap
        self.sliding_window_size = sliding_window_size or -1
        self.is_cross_attention = is_cross_attention
        self.k_scale = None
        self.v_scale = None
        self.k_scale_float = None
        self.v_scale_float = None

    def forward(self, q, k, v, forward_batch, save_kv_cache: bool = True, **kwargs):
        if k is not None:
            assert v is not None
            k = k.view(-1, self.tp_k_head_num, self.qk_head_dim)
            v = v.view(-1, self.tp_v_head_num, self.v_head_dim)
        return forward_batch.attn_backend.forward(q, k, v, self, forward_batch, save_kv_cache, **kwargs)

    __call__ = forward


@dataclass
class ForwardBatch:
    """The fields of the reference ``ForwardBatch`` that attention backends use."""

    forward_mode: ForwardMode
    batch_size: int
    req_pool_indices: torch.Tensor
    seq_lens: torch.Tensor
    out_cache_loc: torch.Tensor
    seq_lens_sum: int
    req_to_token_pool: Any = None
    token_to_kv_pool: Any = None
    attn_backend: Any = None
    seq_lens_cpu: Optional[torch.Tensor] = None
    spec_info: Any = None
    encoder_lens: Optional[torch.Tensor] = None
    input_ids: Optional[torch.Tensor] = None
    positions: Optional[torch.Tensor] = None
    extend_prefix_lens: Optional[torch.Tensor] = None
    extend_seq_lens: Optional[torch.Tensor] = None
    extend_seq_lens_cpu: Optional[List[int]] = None
    extend_prefix_lens_cpu: Optional[List[int]] = None


class AttentionBackend(ABC):
    """Same method set and dispatch as the reference base class."""

    @abstractmethod
    def init_forward_metadata(self, forward_batch: ForwardBatch):
        raise NotImplementedError()

    def init_cuda_graph_state(self, max_bs: int, max_num_tokens: int):
        raise NotImplementedError()

    def init_forward_metadata_capture_cuda_graph(self, bs, num_tokens, req_pool_indices, seq_lens,
                                                 encoder_lens, forward_mode, spec_info):
        raise NotImplementedError()

    def init_forward_metadata_replay_cuda_graph(self, bs, req_pool_indices, seq_lens, seq_lens_sum,
                                                encoder_lens, forward_mode, spec_info, seq_lens_cpu):
        raise NotImplementedError()

    def get_cuda_graph_seq_len_fill_value(self):
        raise NotImplementedError()

    def forward(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
        if forward_batch.forward_mode.is_idle():
            return q.new_empty(q.shape[0], layer.tp_q_head_num * layer.v_head_dim)
        elif forward_batch.forward_mode.is_decode():
            return self.forward_decode(q, k, v, layer, forward_batch, save_kv_cache=save_kv_cache, **kwargs)
        else:
            return self.forward_extend(q, k, v, layer, forward_batch, save_kv_cache=save_kv_cache, **kwargs)

    def forward_decode(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True):
        raise NotImplementedError()

    def forward_extend(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True):
        raise NotImplementedError()

    def support_triton(self):
        return True
