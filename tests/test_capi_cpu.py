"""CPU tests of the drop-in boundary: libb200attn.so loads, exports every symbol include/b200attn.h
declares, the ctypes structures match the C layouts, and argument errors are reported without
touching a GPU (no compute calls here)."""
from __future__ import annotations

import ctypes as C
import os
import re
import subprocess
import sys
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200attn.h")


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200attn_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from b200attn import capi

    lib = capi.lib()
    declared = _declared_functions()
    assert len(declared) >= 10
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/b200attn.h but not exported"
        assert name in capi.SYMBOLS, f"{name} has no ctypes prototype in capi.SYMBOLS"
    out = subprocess.run(["nm", "-D", "--defined-only", str(capi.LIB_PATH)], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (\w+)", out))
    assert set(declared) <= exported
    # nothing but the C-ABI leaks out (hidden visibility for everything else)
    assert all(s.startswith("b200attn_") for s in exported if not s.startswith("_")), exported


def test_abi_version_and_strerror():
    from b200attn import capi

    assert capi.lib().b200attn_abi_version() == capi.ABI_VERSION == 6
    for code in (0, -1, -2, -3, -4, -5, -99):
        assert isinstance(capi.strerror(code), str) and capi.strerror(code)
    assert "unsupported" in capi.strerror(capi.E_UNSUPPORTED)


def test_ctypes_layout_matches_c_header():
    """Compile a C program against the header and compare sizeof / offsetof with ctypes."""
    from b200attn import capi

    structs = {"b200attn_kv_store_args": capi.KvStoreArgs, "b200attn_plan_args": capi.PlanArgs,
               "b200attn_decode_args": capi.DecodeArgs, "b200attn_merge_states_args": capi.MergeStatesArgs,
               "b200attn_rope_kv_store_args": capi.RopeKvStoreArgs, "b200attn_extend_args": capi.ExtendArgs,
               "b200attn_extend_plan_args": capi.ExtendPlanArgs}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "b200attn.h"', 'int main(void){']
    for cname, cls in structs.items():
        lines.append(f'printf("{cname} %zu\\n", sizeof({cname}));')
        for fname, _ in cls._fields_:
            lines.append(f'printf("{cname}.{fname} %zu\\n", offsetof({cname}, {fname}));')
    lines.append("return 0;}")
    with tempfile.TemporaryDirectory() as d:
        src = os.path.join(d, "layout.c")
        open(src, "w").write("\n".join(lines))
        exe = os.path.join(d, "layout")
        subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.dirname(HEADER), src, "-o", exe], check=True)
        got = dict(l.split() for l in subprocess.run([exe], capture_output=True, text=True, check=True).stdout.splitlines())
    for cname, cls in structs.items():
        assert int(got[cname]) == C.sizeof(cls), cname
        for fname, _ in cls._fields_:
            assert int(got[f"{cname}.{fname}"]) == getattr(cls, fname).offset, f"{cname}.{fname}"


def test_header_is_plain_c():
    """The boundary is a C ABI: the header compiles as C99 and as C++, with no CUDA / torch includes."""
    src = open(HEADER).read()
    assert "#include <cuda" not in src and "#include <torch" not in src and "#include <ATen" not in src
    with tempfile.TemporaryDirectory() as d:
        for ext, cc in ((".c", ["gcc", "-std=c99"]), (".cpp", ["g++", "-std=c++17"])):
            f = os.path.join(d, "t" + ext)
            open(f, "w").write('#include "b200attn.h"\nint main(void){b200attn_decode_args a; a.batch = 0; return a.batch;}\n')
            subprocess.run(cc + ["-Wall", "-Werror", "-fsyntax-only", "-I", os.path.dirname(HEADER), f], check=True)


def test_geometry_table():
    from b200attn import capi

    g = capi.decode_geometry(32, 8, 128, capi.BF16)       # C2 / C5
    assert g == {"heads_per_cta": 8, "token_slices": 1, "chunk_tokens": 16, "stages": 3,
                 "smem_bytes": 198400, "head_groups": 1}
    g = capi.decode_geometry(8, 1, 128, capi.BF16)        # C4 at TP8: one kv head per rank
    assert (g["heads_per_cta"], g["token_slices"], g["chunk_tokens"]) == (1, 8, 128)
    g = capi.decode_geometry(32, 8, 128, capi.FP8_E4M3)   # C3
    # fp8: two 16-token sub-chunks per warp and chunk, so a slot carries the bytes of a 16-bit one
    assert g["chunk_tokens"] == 32 and g["stages"] == 3 and g["smem_bytes"] <= 227 * 1024
    g = capi.decode_geometry(32, 8, 64, capi.BF16, batch=16)   # C1: few requests -> 2 heads per CTA, 256-byte rows
    assert (g["heads_per_cta"], g["token_slices"], g["head_groups"], g["chunk_tokens"]) == (2, 4, 4, 64)
    g = capi.decode_geometry(32, 8, 64, capi.BF16, batch=64)   # enough requests: the wide rows stay
    assert g["heads_per_cta"] == 8
    g = capi.decode_geometry(32, 32, 128, capi.BF16)      # MHA: four head groups of eight
    assert g["head_groups"] == 4
    g = capi.decode_geometry(8, 4, 256, capi.BF16)        # Gemma-shaped: K and V of a chunk in separate pipeline slots
    assert (g["heads_per_cta"], g["token_slices"], g["stages"]) == (4, 2, 3)
    assert capi.decode_geometry(32, 8, 80, capi.BF16)["heads_per_cta"] == 8 and capi.decode_geometry(32, 8, 96, capi.BF16)["stages"] == 4
    for bad in ((32, 8, 72), (32, 8, 512), (33, 8, 128), (64, 2, 128)):  # head dims without a kernel, ragged, group 32
        with pytest.raises(capi.B200AttnError) as e:
            capi.decode_geometry(*bad, capi.BF16)
        assert e.value.code == capi.E_UNSUPPORTED


def test_argument_errors_need_no_gpu():
    """Validation happens before any CUDA call: bad arguments return codes, never launch."""
    from b200attn import capi

    lib = capi.lib()
    n0 = capi.launch_count()
    assert lib.b200attn_kv_store(None, None) == capi.E_INVALID
    assert lib.b200attn_plan(None, None) == capi.E_INVALID
    assert lib.b200attn_decode(None, None) == capi.E_INVALID
    assert lib.b200attn_merge_states(None, None) == capi.E_INVALID
    assert lib.b200attn_extend(None, None) == capi.E_INVALID
    assert lib.b200attn_extend_plan(None, None) == capi.E_INVALID
    e = capi.ExtendArgs()
    e.batch, e.total_tokens, e.max_extend_len = 2, 16, 8
    assert lib.b200attn_extend(C.byref(e), None) == capi.E_INVALID            # null pointers
    e.batch = 0
    assert lib.b200attn_extend(C.byref(e), None) == capi.OK                   # empty batch is a no-op
    a = capi.KvStoreArgs()
    a.n_tokens, a.num_kv_heads, a.head_dim, a.v_head_dim = 4, 2, 64, 64
    assert lib.b200attn_kv_store(C.byref(a), None) == capi.E_INVALID          # null pointers
    a.n_tokens = 0
    assert lib.b200attn_kv_store(C.byref(a), None) == capi.OK                 # empty batch is a no-op
    p = capi.PlanArgs()
    p.batch = 4
    assert lib.b200attn_plan(C.byref(p), None) == capi.E_INVALID
    d = capi.DecodeArgs()
    d.batch = -1
    assert lib.b200attn_decode(C.byref(d), None) == capi.E_INVALID
    # the allocation-free decode entry point: its own checks come before any CUDA call too
    assert lib.b200attn_decode_static_ws(None, None) == capi.E_INVALID
    s = capi.DecodeArgs()
    s.batch, s.num_q_heads, s.num_kv_heads, s.head_dim, s.v_head_dim = 4096, 64, 8, 128, 128
    s.q_dtype = s.kv_dtype = capi.BF16
    assert lib.b200attn_workspace_bytes(4096, 64, 8, 128, 128, capi.BF16) > capi.STATIC_WS_BYTES
    assert lib.b200attn_decode_static_ws(C.byref(s), None) == capi.E_STATIC_LIMIT   # beyond the static array
    assert "static workspace" in capi.strerror(capi.E_STATIC_LIMIT)
    s.batch = 0
    assert lib.b200attn_decode_static_ws(C.byref(s), None) == capi.OK               # empty batch is a no-op
    s.batch, s.ws_bytes = 8, 64
    assert lib.b200attn_decode_static_ws(C.byref(s), None) == capi.E_INVALID        # the library owns the workspace
    s.ws_bytes, s.head_dim, s.v_head_dim = 0, 72, 72
    assert lib.b200attn_decode_static_ws(C.byref(s), None) == capi.E_UNSUPPORTED
    assert capi.launch_count() == n0
    assert lib.b200attn_plan_bytes(128) > 0 and lib.b200attn_plan_bytes(-1) == 0
    assert lib.b200attn_workspace_bytes(128, 32, 8, 128, 128, capi.BF16) > 0
    assert lib.b200attn_workspace_bytes(128, 32, 8, 72, 72, capi.BF16) == 0


def test_ops_refuse_cpu_tensors():
    """torch.ops.b200attn.* exist and have no CPU kernel: the product path cannot fall back."""
    import torch

    import b200attn  # noqa: F401

    for name in ("kv_store", "plan", "plan_csr", "decode", "merge_states"):
        assert hasattr(torch.ops.b200attn, name)
    z = torch.zeros(1, 1, 8, dtype=torch.bfloat16)
    with pytest.raises((NotImplementedError, RuntimeError)):
        torch.ops.b200attn.merge_states(z, torch.zeros(1, 1), z, torch.zeros(1, 1), z.clone(), None)


def test_backend_refuses_cpu_device():
    import b200attn
    import synth
    from tests.harness import make_runner

    w = synth.make_workload(synth.Config("t", 2, 0, 8, 2, 64, ragged=(3, 20), seed=1), "cpu")
    # the mirror pool itself is plain torch and can live on CPU; the backend cannot
    runner = make_runner(w, "cpu")
    with pytest.raises(RuntimeError, match="CUDA"):
        b200attn.B200AttnBackend(runner)


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under ltp-sglang_b200/ may reference it."""
    pkg = os.path.join(ROOT, "ltp-sglang_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                bad = re.search(r"(import\s+oracle|from\s+oracle|oracle[/.](_ref|decode_oracle|ref_|c/)|#include\s+\"[^\"]*oracle)",
                                text)
                assert bad is None, (os.path.join(dirpath, f), bad.group(0))
