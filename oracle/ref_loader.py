"""Loader for the reference's OWN Python classes (oracle tooling, container only).

Imports ``TorchNativeAttnBackend``, ``MHATokenToKVPool``, ``ReqToTokenPool``,
``ForwardBatch``, ``ForwardMode`` and ``RadixAttention`` from the read-only reference tree
so golden vectors can be generated from the reference itself.
``/root/reference`` does not exist on the GPU box: nothing that runs there may call this.

Three shims are needed because the container is not the reference's pinned environment:
1. ``IPython`` is absent (python/sglang/utils.py:25 imports it);
2. ``torch._C._cpu._is_amx_tile_supported`` is missing in torch 2.11
   (python/sglang/srt/utils.py:217);
3. ``sglang.srt.model_executor.cuda_graph_runner`` (lazily imported by set_kv_buffer,
   mem_cache/memory_pool.py:379) drags in configs that fail under transformers 5.x, so a
   stub exposing ``get_is_capture_mode() -> False`` is registered instead.
"""
from __future__ import annotations

import os
import sys
import types

REFERENCE_ROOT = os.environ.get("B200ATTN_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "python", "sglang"))


_cache = None


def load():
    """Returns a namespace with the reference classes; raises if the tree is not mounted."""
    global _cache
    if _cache is not None:
        return _cache
    if not available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    import torch

    if "IPython" not in sys.modules:
        ip = types.ModuleType("IPython")
        disp = types.ModuleType("IPython.display")
        disp.HTML = object
        disp.display = lambda *a, **k: None
        ip.display = disp
        sys.modules["IPython"] = ip
        sys.modules["IPython.display"] = disp
    if not hasattr(torch._C._cpu, "_is_amx_tile_supported"):
        torch._C._cpu._is_amx_tile_supported = lambda: False
    sys.path.insert(0, os.path.join(REFERENCE_ROOT, "python"))
    try:
        from sglang.srt.layers.attention.torch_native_backend import TorchNativeAttnBackend
        from sglang.srt.layers.radix_attention import RadixAttention
        from sglang.srt.mem_cache.memory_pool import MHATokenToKVPool, ReqToTokenPool, SWAKVPool
        from sglang.srt.model_executor.forward_batch_info import ForwardBatch, ForwardMode
    finally:
        sys.path.pop(0)
    name = "sglang.srt.model_executor.cuda_graph_runner"
    if name not in sys.modules:
        cg = types.ModuleType(name)
        cg.get_is_capture_mode = lambda: False
        sys.modules[name] = cg
    _cache = types.SimpleNamespace(
        TorchNativeAttnBackend=TorchNativeAttnBackend, RadixAttention=RadixAttention,
        MHATokenToKVPool=MHATokenToKVPool, ReqToTokenPool=ReqToTokenPool, SWAKVPool=SWAKVPool,
        ForwardBatch=ForwardBatch, ForwardMode=ForwardMode)
    return _cache
