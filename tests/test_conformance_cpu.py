"""Boundary conformance: everything ``B200AttnBackend`` reads from sglang's objects exists on
the REFERENCE's own classes under the same names, and the hooks it implements have the reference's parameter
lists -- so the mirror classes in ``b200attn/interfaces.py`` (which every GPU test goes through) cannot drift
away from what a real sglang process hands the backend.  Imports the reference through ``oracle/ref_loader.py``
(container only: skipped where the read-only tree is not mounted)."""
from __future__ import annotations

import ast
import inspect
import os
import sys

import pytest

from oracle import ref_loader

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ltp-sglang_b200"))

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference tree not mounted")


@pytest.fixture(scope="module")
def ref():
    return ref_loader.load()


def _params(fn):
    return [p for p in inspect.signature(fn).parameters if p != "self"]


def test_attention_backend_hooks_match_reference(ref):
    """Same hook names and parameter lists as AttentionBackend / TritonAttnBackend
    (base_attn_backend.py:14-117, triton_backend.py:338-630)."""
    import b200attn

    sys.path.insert(0, os.path.join(ref_loader.REFERENCE_ROOT, "python"))
    try:
        from sglang.srt.layers.attention.base_attn_backend import AttentionBackend as RefBase
    finally:
        sys.path.pop(0)
    ours = b200attn.B200AttnBackend
    for name in ("init_forward_metadata", "init_cuda_graph_state", "init_forward_metadata_capture_cuda_graph",
                 "init_forward_metadata_replay_cuda_graph", "get_cuda_graph_seq_len_fill_value", "forward",
                 "forward_decode", "forward_extend", "support_triton"):
        assert hasattr(RefBase, name), f"reference AttentionBackend has no {name}"
        want, got = _params(getattr(RefBase, name)), _params(getattr(ours, name))
        assert got[: len(want)] == want, f"{name}: reference takes {want}, backend takes {got}"
    # the mirror base class dispatches exactly like the reference's forward()
    assert _params(b200attn.AttentionBackend.forward) == _params(RefBase.forward)


def _attrs_read(src_file, var_names):
    """Attribute names read off the given variables in a source file (``var.attr`` and getattr(var, "attr"))."""
    tree = ast.parse(open(src_file).read())
    out = {v: set() for v in var_names}
    for node in ast.walk(tree):
        if isinstance(node, ast.Attribute) and isinstance(node.value, ast.Name) and node.value.id in out:
            out[node.value.id].add(node.attr)
        if isinstance(node, ast.Call) and isinstance(node.func, ast.Name) and node.func.id == "getattr" \
                and len(node.args) >= 2 and isinstance(node.args[0], ast.Name) and node.args[0].id in out \
                and isinstance(node.args[1], ast.Constant):
            out[node.args[0].id].add(node.args[1].value)
    return out


def test_every_attribute_the_backend_reads_exists_in_the_reference(ref):
    src = os.path.join(ROOT, "ltp-sglang_b200", "b200attn", "backend.py")
    reads = _attrs_read(src, ["forward_batch", "layer", "pool", "fb"])
    # ForwardBatch: dataclass fields (forward_batch_info.py:163-311)
    fb_fields = set(ref.ForwardBatch.__dataclass_fields__)
    missing = {a for a in reads["forward_batch"] | reads["fb"] if a not in fb_fields and not hasattr(ref.ForwardBatch, a)}
    assert not missing, f"ForwardBatch has no {sorted(missing)}"
    # RadixAttention: attributes set in __init__ (radix_attention.py:39-91)
    radix_src = inspect.getsource(ref.RadixAttention.__init__)
    optional = {"attn_type"} if "attn_type" not in radix_src else set()
    missing = {a for a in reads["layer"] if f"self.{a}" not in radix_src and a not in optional}
    assert not missing, f"RadixAttention.__init__ sets no {sorted(missing)}"
    # MHATokenToKVPool: k_buffer / v_buffer / start_layer (memory_pool.py:164-258)
    # ... and, for the hybrid pool, SWAKVPool's members (memory_pool.py:420-553)
    pool_src = inspect.getsource(ref.MHATokenToKVPool) + inspect.getsource(ref.MHATokenToKVPool.__mro__[1])
    swa_src = inspect.getsource(ref.SWAKVPool)
    missing = {a for a in reads["pool"] if f"self.{a}" not in pool_src and not hasattr(ref.MHATokenToKVPool, a)
               and f"self.{a}" not in swa_src and not hasattr(ref.SWAKVPool, a)}
    assert not missing, f"neither MHATokenToKVPool nor SWAKVPool has {sorted(missing)}"
    import b200attn
    for name in ("get_key_buffer", "get_value_buffer", "get_kv_buffer", "translate_loc_from_full_to_swa", "set_kv_buffer"):
        assert _params(getattr(ref.SWAKVPool, name))[:2] == _params(getattr(b200attn.SWAKVPool, name))[:2], name
    assert _params(ref.SWAKVPool.__init__) == _params(b200attn.SWAKVPool.__init__)


def test_mirror_classes_cover_the_reference_surface(ref):
    """interfaces.py mirrors: every field / method the mirror exposes exists in the reference class, and
    ForwardMode has the reference's members with the same predicates."""
    import b200attn

    for f in b200attn.ForwardBatch.__dataclass_fields__:
        assert f in ref.ForwardBatch.__dataclass_fields__, f"mirror ForwardBatch field {f} is not in the reference"
    for m in ("EXTEND", "DECODE", "MIXED", "IDLE", "TARGET_VERIFY", "DRAFT_EXTEND"):
        assert hasattr(ref.ForwardMode, m) and hasattr(b200attn.ForwardMode, m)
        rm, om = getattr(ref.ForwardMode, m), getattr(b200attn.ForwardMode, m)
        for pred in ("is_decode", "is_idle", "is_decode_or_idle", "is_extend", "is_target_verify", "is_draft_extend"):
            assert bool(getattr(rm, pred)()) == bool(getattr(om, pred)()), f"ForwardMode.{m}.{pred}() differs"
    for name in ("get_key_buffer", "get_value_buffer", "get_kv_buffer", "set_kv_buffer"):
        assert _params(getattr(ref.MHATokenToKVPool, name))[:2] == _params(getattr(b200attn.MHATokenToKVPool, name))[:2]
    want = _params(ref.MHATokenToKVPool.set_kv_buffer)
    got = _params(b200attn.MHATokenToKVPool.set_kv_buffer)
    assert got[: len(want)] == want, (want, got)
    for name in ("write", "available_size", "alloc", "free", "clear"):
        assert hasattr(ref.ReqToTokenPool, name) and hasattr(b200attn.ReqToTokenPool, name)
    # RadixAttention attributes the mirror layer carries
    layer = b200attn.AttnLayer(8, 64, 0.125, 2, 0)
    radix_src = inspect.getsource(ref.RadixAttention.__init__)
    for a in vars(layer):
        assert f"self.{a}" in radix_src or a == "attn_type", f"mirror AttnLayer.{a} is not a RadixAttention attribute"


def test_multi_step_draft_backend_matches_reference_surface():
    """TritonMultiStepDraftBackend's constructor and hooks (triton_backend.py:735-872), read from its source (the
    module needs triton + CUDA to import)."""
    import b200attn

    path = os.path.join(ref_loader.REFERENCE_ROOT, "python", "sglang", "srt", "layers", "attention", "triton_backend.py")
    tree = ast.parse(open(path).read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "TritonMultiStepDraftBackend")
    ref_methods = {f.name: [a.arg for a in f.args.args if a.arg != "self"] for f in cls.body if isinstance(f, ast.FunctionDef)}
    for name, want in ref_methods.items():
        got = _params(getattr(b200attn.B200MultiStepDraftBackend, name))
        got = [g for g in got if g != "backend_kwargs"]
        assert got == want, f"{name}: reference takes {want}, ours {got}"
    tcls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "TritonAttnBackend")
    init = next(f for f in tcls.body if isinstance(f, ast.FunctionDef) and f.name == "__init__")
    want = [a.arg for a in init.args.args if a.arg != "self"]
    assert _params(b200attn.B200AttnBackend.__init__)[: len(want)] == want
    cgs = next(f for f in tcls.body if isinstance(f, ast.FunctionDef) and f.name == "init_cuda_graph_state")
    assert _params(b200attn.B200AttnBackend.init_cuda_graph_state) == [a.arg for a in cgs.args.args if a.arg != "self"]
