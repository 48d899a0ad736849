"""Loader for the REFERENCE's compiled CPU decode kernel (oracle/_ref/, built by
oracle/ref_build/Makefile from /root/reference/sgl-kernel/csrc/cpu/decode.cpp).

TEST / BASELINE INFRASTRUCTURE ONLY: used to validate the restated oracle and as the
``"kind": "reference"`` CPU baseline of bench.py.  Call pattern follows the reference's
``IntelAMXAttnBackend.forward_decode`` (layers/attention/intel_amx_backend.py:91-125): the op
fuses "store new K/V at loc" + split-KV decode (8 fixed splits) + merge.
"""
from __future__ import annotations

import os
import subprocess
import sys
from typing import Optional

import torch

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
NUM_KV_SPLITS = 8  # the reference CPU backend's fixed split count (intel_amx_backend.py:27-33)
_loaded: Optional[str] = None


def _cpu_flags() -> set:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def variant() -> Optional[str]:
    """Which prebuilt variant this host can run (None if nothing was built)."""
    flags = _cpu_flags()
    native = os.path.join(_DIR, "libsgl_cpu_decode_native.so")
    portable = os.path.join(_DIR, "libsgl_cpu_decode_portable.so")
    need = {"avx512f", "avx512_bf16", "amx_bf16", "amx_tile", "avx512vl", "avx512bw", "avx512dq"}
    if os.path.exists(native) and need <= flags:
        return native
    if os.path.exists(portable) and {"avx2", "fma"} <= flags:
        return portable
    return None


def load() -> str:
    """torch.ops.load_library on the variant; probes it in a subprocess first so an illegal
    instruction on a foreign CPU cannot take the caller down."""
    global _loaded
    if _loaded:
        return _loaded
    path = variant()
    if path is None:
        raise RuntimeError("oracle/_ref is not built (or this CPU cannot run it)")
    probe = ("import torch,sys; torch.ops.load_library(sys.argv[1]); import torch as t;"
             "q=t.randn(1,2,64).bfloat16(); kb=t.randn(8,1,64).bfloat16(); vb=kb.clone(); o=t.zeros(1,2,64).bfloat16();"
             "k=t.randn(1,1,64).bfloat16(); al=t.empty(1,2,8,65);"
             "t.ops.b200attn_ref.decode_attention_cpu(q,kb,vb,o,k,k.clone(),t.tensor([3]),al,"
             "t.arange(8,dtype=t.int32).reshape(1,8),t.tensor([0]),t.tensor([4]),0.125,0.0); print('ok')")
    r = subprocess.run([sys.executable, "-c", probe, path], capture_output=True, text=True, timeout=300)
    if r.returncode != 0 or "ok" not in r.stdout:
        raise RuntimeError(f"oracle/_ref probe failed (rc={r.returncode}): {r.stderr[-300:]}")
    torch.ops.load_library(path)
    _loaded = path
    return path


def decode_attention_cpu(q, k_pool, v_pool, k_new, v_new, loc, req_to_token, req_pool_indices, seq_lens,
                         sm_scale: float, logit_cap: float = 0.0, attn_logits=None) -> torch.Tensor:
    """Runs the reference kernel: stores k_new/v_new at loc (in place on the pools) and returns
    o [B, Hq, Dv].  q [B, Hq, D]; pools [slots, Hkv, D]; int64 req_pool_indices / seq_lens."""
    load()
    B, Hq, _ = q.shape
    Dv = v_pool.shape[-1]
    if attn_logits is None:
        attn_logits = torch.empty((B, Hq, NUM_KV_SPLITS, Dv + 1), dtype=torch.float32)
    o = torch.empty((B, Hq, Dv), dtype=q.dtype)
    torch.ops.b200attn_ref.decode_attention_cpu(
        q, k_pool, v_pool, o, k_new, v_new, loc.to(torch.int64), attn_logits, req_to_token,
        req_pool_indices.to(torch.int64), seq_lens.to(torch.int64), float(sm_scale), float(logit_cap))
    return o
