"""ctypes binding of the C-ABI in ``include/b200attn.h``.

This is the only place that touches ``libb200attn.so``.  There is no CPU or
PyTorch fallback: if the library is missing, importing the ops raises.
Structure layouts mirror the header field by field.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_HERE = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("B200ATTN_LIB", _HERE / "libb200attn.so"))

# enums (include/b200attn.h)
OK = 0
E_INVALID, E_UNSUPPORTED, E_ALIGN, E_CUDA, E_WORKSPACE, E_STATIC_LIMIT = -1, -2, -3, -4, -5, -6
STATIC_WS_BYTES = 64 << 20  # B200ATTN_STATIC_WS_BYTES
BF16, F16, FP8_E4M3, F32, FP8_E5M2 = 0, 1, 2, 3, 4
I32, I64 = 0, 1
SCALE_NONE, SCALE_TRUE_DIV, SCALE_RECIP_MUL = 0, 1, 2
ABI_VERSION = 6


class KvStoreArgs(C.Structure):
    _fields_ = [
        ("k", C.c_void_p),
        ("v", C.c_void_p),
        ("k_stride_token", C.c_int64),
        ("k_stride_head", C.c_int64),
        ("v_stride_token", C.c_int64),
        ("v_stride_head", C.c_int64),
        ("loc", C.c_void_p),
        ("k_pool", C.c_void_p),
        ("v_pool", C.c_void_p),
        ("k_pool_stride_slot", C.c_int64),
        ("v_pool_stride_slot", C.c_int64),
        ("pool_slots", C.c_int64),
        ("n_tokens", C.c_int32),
        ("num_kv_heads", C.c_int32),
        ("head_dim", C.c_int32),
        ("v_head_dim", C.c_int32),
        ("src_dtype", C.c_int32),
        ("pool_dtype", C.c_int32),
        ("scale_mode", C.c_int32),
        ("mutate_src", C.c_int32),
        ("k_scale", C.c_float),
        ("v_scale", C.c_float),
        ("k_scale_ptr", C.c_void_p),
        ("v_scale_ptr", C.c_void_p),
    ]


class RopeKvStoreArgs(C.Structure):
    _fields_ = [
        ("q", C.c_void_p),
        ("q_stride_token", C.c_int64),
        ("q_stride_head", C.c_int64),
        ("num_q_heads", C.c_int32),
        ("rotary_dim", C.c_int32),
        ("is_neox", C.c_int32),
        ("pos_dtype", C.c_int32),
        ("positions", C.c_void_p),
        ("cos_sin_cache", C.c_void_p),
        ("cache_stride", C.c_int64),
        ("max_position", C.c_int64),
        ("save_kv", C.c_int32),
        ("pad_", C.c_int32),
        ("kv", KvStoreArgs),
    ]


class PlanArgs(C.Structure):
    _fields_ = [
        ("req_to_token", C.c_void_p),
        ("req_to_token_stride", C.c_int64),
        ("req_pool_indices", C.c_void_p),
        ("seq_lens", C.c_void_p),
        ("index_dtype", C.c_int32),
        ("batch", C.c_int32),
        ("num_q_heads", C.c_int32),
        ("num_kv_heads", C.c_int32),
        ("head_dim", C.c_int32),
        ("kv_dtype", C.c_int32),
        ("csr_indptr", C.c_void_p),
        ("kv_indptr", C.c_void_p),
        ("kv_indices", C.c_void_p),
        ("plan", C.c_void_p),
        ("plan_bytes", C.c_size_t),
        ("max_ctas", C.c_int32),
        ("min_chunks_per_cta", C.c_int32),
        ("dynamic_percent", C.c_int32),
        ("dynamic_range_chunks", C.c_int32),
        ("sliding_window", C.c_int32),
        ("pad_", C.c_int32),
    ]


class DecodeArgs(C.Structure):
    _fields_ = [
        ("q", C.c_void_p),
        ("q_stride_batch", C.c_int64),
        ("q_stride_head", C.c_int64),
        ("k_pool", C.c_void_p),
        ("v_pool", C.c_void_p),
        ("k_pool_stride_slot", C.c_int64),
        ("v_pool_stride_slot", C.c_int64),
        ("o", C.c_void_p),
        ("o_stride_batch", C.c_int64),
        ("o_stride_head", C.c_int64),
        ("lse", C.c_void_p),
        ("plan", C.c_void_p),
        ("indices", C.c_void_p),
        ("ws_o", C.c_void_p),
        ("ws_bytes", C.c_size_t),
        ("batch", C.c_int32),
        ("num_q_heads", C.c_int32),
        ("num_kv_heads", C.c_int32),
        ("head_dim", C.c_int32),
        ("v_head_dim", C.c_int32),
        ("q_dtype", C.c_int32),
        ("kv_dtype", C.c_int32),
        ("sm_scale", C.c_float),
        ("logit_cap", C.c_float),
        ("k_scale", C.c_float),
        ("v_scale", C.c_float),
        ("phases", C.c_int32),
        ("slot_map", C.c_void_p),
        ("pool_slots", C.c_int64),
    ]


class ExtendArgs(C.Structure):
    _fields_ = [
        ("q", C.c_void_p),
        ("q_stride_token", C.c_int64),
        ("q_stride_head", C.c_int64),
        ("o", C.c_void_p),
        ("o_stride_token", C.c_int64),
        ("o_stride_head", C.c_int64),
        ("lse", C.c_void_p),
        ("k_pool", C.c_void_p),
        ("v_pool", C.c_void_p),
        ("k_pool_stride_slot", C.c_int64),
        ("v_pool_stride_slot", C.c_int64),
        ("qo_indptr", C.c_void_p),
        ("kv_indptr", C.c_void_p),
        ("kv_indices", C.c_void_p),
        ("req_to_token", C.c_void_p),
        ("req_to_token_stride", C.c_int64),
        ("req_pool_indices", C.c_void_p),
        ("prefix_lens", C.c_void_p),
        ("index_dtype", C.c_int32),
        ("mask_indptr_dtype", C.c_int32),
        ("extend_loc", C.c_void_p),
        ("custom_mask", C.c_void_p),
        ("mask_indptr", C.c_void_p),
        ("batch", C.c_int32),
        ("total_tokens", C.c_int32),
        ("max_extend_len", C.c_int32),
        ("num_q_heads", C.c_int32),
        ("num_kv_heads", C.c_int32),
        ("head_dim", C.c_int32),
        ("v_head_dim", C.c_int32),
        ("q_dtype", C.c_int32),
        ("kv_dtype", C.c_int32),
        ("causal", C.c_int32),
        ("sliding_window", C.c_int32),
        ("skip_prefix_custom_mask", C.c_int32),
        ("sm_scale", C.c_float),
        ("logit_cap", C.c_float),
        ("k_scale", C.c_float),
        ("v_scale", C.c_float),
        ("avg_kv_len_hint", C.c_int32),
    ]


class ExtendPlanArgs(C.Structure):
    _fields_ = [
        ("extend_seq_lens", C.c_void_p),
        ("prefix_lens", C.c_void_p),
        ("index_dtype", C.c_int32),
        ("batch", C.c_int32),
        ("uniform_extend_len", C.c_int32),
        ("sliding_window", C.c_int32),
        ("qo_indptr", C.c_void_p),
        ("kv_indptr", C.c_void_p),
        ("mask_indptr", C.c_void_p),
    ]


class MergeStatesArgs(C.Structure):
    _fields_ = [
        ("o_a", C.c_void_p),
        ("lse_a", C.c_void_p),
        ("o_b", C.c_void_p),
        ("lse_b", C.c_void_p),
        ("o_out", C.c_void_p),
        ("lse_out", C.c_void_p),
        ("tokens", C.c_int32),
        ("heads", C.c_int32),
        ("head_dim", C.c_int32),
        ("dtype", C.c_int32),
    ]


class DraftKvIndicesArgs(C.Structure):
    _fields_ = [
        ("req_pool_indices", C.c_void_p),
        ("req_to_token", C.c_void_p),
        ("req_to_token_stride", C.c_int64),
        ("seq_lens", C.c_void_p),
        ("index_dtype", C.c_int32),
        ("positions_dtype", C.c_int32),
        ("positions", C.c_void_p),
        ("kv_indices", C.c_void_p),
        ("kv_indices_stride", C.c_int64),
        ("kv_indptr", C.c_void_p),
        ("kv_indptr_stride", C.c_int64),
        ("num_steps", C.c_int32),
        ("num_seqs", C.c_int32),
        ("topk", C.c_int32),
        ("page_size", C.c_int32),
    ]


# name -> (restype, argtypes); every symbol include/b200attn.h declares.
SYMBOLS = {
    "b200attn_kv_store": (C.c_int, [C.POINTER(KvStoreArgs), C.c_void_p]),
    "b200attn_rope_kv_store": (C.c_int, [C.POINTER(RopeKvStoreArgs), C.c_void_p]),
    "b200attn_plan_bytes": (C.c_size_t, [C.c_int32]),
    "b200attn_plan": (C.c_int, [C.POINTER(PlanArgs), C.c_void_p]),
    "b200attn_workspace_bytes": (
        C.c_size_t,
        [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32],
    ),
    "b200attn_decode": (C.c_int, [C.POINTER(DecodeArgs), C.c_void_p]),
    "b200attn_decode_static_ws": (C.c_int, [C.POINTER(DecodeArgs), C.c_void_p]),
    "b200attn_extend": (C.c_int, [C.POINTER(ExtendArgs), C.c_void_p]),
    "b200attn_extend_plan": (C.c_int, [C.POINTER(ExtendPlanArgs), C.c_void_p]),
    "b200attn_draft_kv_indices": (C.c_int, [C.POINTER(DraftKvIndicesArgs), C.c_void_p]),
    "b200attn_merge_states": (C.c_int, [C.POINTER(MergeStatesArgs), C.c_void_p]),
    "b200attn_abi_version": (C.c_int, []),
    "b200attn_strerror": (C.c_char_p, [C.c_int]),
    "b200attn_launch_count": (C.c_uint64, []),
    "b200attn_decode_geometry": (
        C.c_int,
        [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32 * 6)],
    ),
    "b200attn_decode_geometry_batch": (
        C.c_int,
        [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32 * 6)],
    ),
}


class B200AttnError(RuntimeError):
    def __init__(self, fn: str, code: int, msg: str):
        super().__init__(f"{fn} failed: {msg} (code {code})")
        self.code = code


_lib = None


def lib() -> C.CDLL:
    """Load libb200attn.so once.  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise ImportError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (or `make -C ltp-sglang_b200/csrc`). There is no CPU fallback."
            )
        handle = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(handle, name)  # AttributeError if the symbol is missing
            fn.restype = res
            fn.argtypes = args
        got = handle.b200attn_abi_version()
        if got != ABI_VERSION:
            raise ImportError(f"libb200attn ABI version {got}, binding expects {ABI_VERSION}")
        _lib = handle
    return _lib


def strerror(code: int) -> str:
    return lib().b200attn_strerror(code).decode()


def check(fn: str, code: int) -> None:
    if code != OK:
        raise B200AttnError(fn, code, strerror(code))


def launch_count() -> int:
    return int(lib().b200attn_launch_count())


def decode_geometry(num_q_heads: int, num_kv_heads: int, head_dim: int, kv_dtype: int, batch: int = 0) -> dict:
    """Geometry of the decode kernel for a shape; ``batch`` > 0: the one plan / decode use for that batch size."""
    out = (C.c_int32 * 6)()
    check(
        "b200attn_decode_geometry_batch",
        lib().b200attn_decode_geometry_batch(batch, num_q_heads, num_kv_heads, head_dim, kv_dtype, C.byref(out)),
    )
    keys = ("heads_per_cta", "token_slices", "chunk_tokens", "stages", "smem_bytes", "head_groups")
    return dict(zip(keys, list(out)))
