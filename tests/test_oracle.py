"""CPU tests: the oracle against the golden vectors produced by the reference itself
(tests/golden/make_golden.py), and against the reference's live code where the read-only tree is
mounted.  These pin the oracle; the GPU tests then compare the CUDA path with the oracle."""
from __future__ import annotations

import os
import types

import numpy as np
import pytest
import torch

import synth
from oracle import decode_oracle as oracle
from oracle import ref_loader
from tests.golden.make_golden import CASES, EXTEND_CASES, KAT, digest, digest_extend
from tests.harness import GOLDEN_DIR, assert_close_out, canon_e4m3_nan, oracle_extend, oracle_step


def _from_bits(a: np.ndarray, dt: torch.dtype) -> torch.Tensor:
    return torch.from_numpy(a.view(np.int16).copy()).view(dt)


@pytest.mark.parametrize("name", sorted(CASES))
def test_decode_oracle_vs_reference_golden(name):
    cfg = CASES[name]
    g = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    w = synth.make_workload(cfg, "cpu")
    assert digest(w) == g["digest"].tobytes().decode(), "synthetic inputs drifted: regenerate goldens"
    fp8 = cfg.kv_dtype == "fp8_e4m3"
    if fp8:  # golden: dequantised copy, scales 1, no store (see make_golden.py)
        out, lse = oracle.decode_attention(w.q.reshape(cfg.batch, cfg.num_q_heads, cfg.head_dim), w.k_pool,
                                           w.v_pool, w.req_to_token, w.req_pool_indices, w.seq_lens,
                                           w.sm_scale)
    else:
        out, lse, kp, vp, _, _ = oracle_step(w)
        loc = w.out_cache_loc
        dt = w.k_pool.dtype
        assert torch.equal(kp[loc], _from_bits(g["k_rows_after"], dt))
        assert torch.equal(vp[loc], _from_bits(g["v_rows_after"], dt))
    ref_dt = torch.float16 if cfg.q_dtype == "f16" else torch.bfloat16
    ref = _from_bits(g["out_bits"], ref_dt).float().numpy()
    # the reference output is a bf16/f16-rounded SDPA; the oracle is float64
    assert_close_out(out, ref, atol=4e-3, rtol=8e-3, what=f"{name}: oracle vs reference")
    indptr, indices = oracle.kv_indices(w.req_to_token, w.req_pool_indices, w.seq_lens)
    assert np.array_equal(indptr, g["kv_indptr"]) and np.array_equal(indices, g["kv_indices"])
    # lse consistency: lse = ln sum exp(scaled logits) => exp(logit - lse) sums to 1 (spot check)
    assert np.isfinite(lse).all()


def test_torch_native_port_matches_golden():
    """The timed CPU port (per-request SDPA in the model dtype) reproduces the reference bit for bit
    on the non-fp8 goldens: it is the same arithmetic, restated."""
    for name, cfg in CASES.items():
        if cfg.kv_dtype == "fp8_e4m3":
            continue
        g = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
        w = synth.make_workload(cfg, "cpu")
        _, _, kp, vp, _, _ = oracle_step(w)
        out = oracle.torch_native_decode(w.q.reshape(cfg.batch, cfg.num_q_heads, cfg.head_dim), kp, vp,
                                         w.req_to_token, w.req_pool_indices, w.seq_lens, w.sm_scale)
        ref_dt = torch.float16 if cfg.q_dtype == "f16" else torch.bfloat16
        assert torch.equal(out, _from_bits(g["out_bits"], ref_dt)), name


def test_fp8_cast_table_exhaustive():
    table = np.load(os.path.join(GOLDEN_DIR, "g6_cast_table_bf16_e4m3fn.npz"))["table"]
    allbf = np.arange(65536, dtype=np.uint32).astype(np.uint16)
    got = oracle.f32_to_e4m3fn_bits(oracle.bf16_bits_to_f32(allbf))
    assert np.array_equal(got, table)
    # and against torch in this process (not only the stored table)
    t = torch.from_numpy(allbf.view(np.int16)).view(torch.bfloat16).to(torch.float8_e4m3fn).view(torch.uint8)
    assert np.array_equal(got, t.numpy())


def test_fp8_cast_f32_inputs_and_kat():
    x = np.array(KAT + [-v for v in KAT], dtype=np.float32)
    want = torch.from_numpy(x).to(torch.float8_e4m3fn).view(torch.uint8).numpy()
    assert np.array_equal(oracle.f32_to_e4m3fn_bits(x), want)
    rng = np.random.default_rng(0)
    y = (rng.standard_normal(200000) * np.exp(rng.uniform(-12, 7, 200000))).astype(np.float32)
    want = torch.from_numpy(y).to(torch.float8_e4m3fn).view(torch.uint8).numpy()
    assert np.array_equal(oracle.f32_to_e4m3fn_bits(y), want)
    # survey table (Appendix A.7)
    tab = [(448.0, 0x7E), (456.0, 0x7E), (464.0, 0x7E), (472.0, 0x7F), (480.0, 0x7F), (1e4, 0x7F),
           (-1e4, 0xFF), (2.0 ** -9, 0x01), (2.0 ** -10, 0x00), (0.0, 0x00), (-0.0, 0x80)]
    for k, v in tab:
        assert int(oracle.f32_to_e4m3fn_bits(np.array([k], dtype=np.float32))[0]) == v, k


def test_e4m3_decode_table():
    b = np.arange(256, dtype=np.uint8)
    want = torch.from_numpy(b).view(torch.float8_e4m3fn).float().numpy()
    got = oracle.e4m3fn_bits_to_f32(b)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.array_equal(got[~np.isnan(got)], want[~np.isnan(want)])


def test_bf16_rounding():
    rng = np.random.default_rng(1)
    x = (rng.standard_normal(100000) * np.exp(rng.uniform(-20, 20, 100000))).astype(np.float32)
    want = torch.from_numpy(x).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(oracle.f32_to_bf16_bits(x), want)


@pytest.mark.parametrize("dname", ["bf16", "f16"])
@pytest.mark.parametrize("tag", ["none", "one", "half_two", "odd"])
def test_fp8_store_oracle_vs_reference_golden(dname, tag):
    g = np.load(os.path.join(GOLDEN_DIR, f"g6_fp8_store_{dname}_{tag}.npz"))
    dt = torch.bfloat16 if dname == "bf16" else torch.float16
    k, v = _from_bits(g["k_in"], dt), _from_bits(g["v_in"], dt)
    loc = torch.from_numpy(g["loc"])
    kp = torch.zeros(g["k_pool"].shape, dtype=torch.uint8)
    vp = torch.zeros(g["v_pool"].shape, dtype=torch.uint8)
    ks, vs = float(g["k_scale"]), float(g["v_scale"])
    oracle.kv_store(kp, vp, loc, k, v, None if ks < 0 else ks, None if vs < 0 else vs, oracle.SCALE_TRUE_DIV)
    assert np.array_equal(canon_e4m3_nan(kp.numpy()), canon_e4m3_nan(g["k_pool"]))
    assert np.array_equal(canon_e4m3_nan(vp.numpy()), canon_e4m3_nan(g["v_pool"]))
    for got, want in ((k, g["k_after"]), (v, g["v_after"])):
        wf = _from_bits(want, dt).float().numpy()
        gf = got.float().numpy()
        nan = np.isnan(wf)
        assert np.array_equal(gf[~nan], wf[~nan]) and np.isnan(gf[nan]).all()


def test_merge_oracle_vs_reference_golden():
    for dname, dt in (("bf16", torch.bfloat16), ("f16", torch.float16)):
        g = np.load(os.path.join(GOLDEN_DIR, f"g8_merge_{dname}.npz"))
        oa = _from_bits(g["o_a"], dt).float().numpy()
        ob = _from_bits(g["o_b"], dt).float().numpy()
        out, lse = oracle.merge_states(oa, g["lse_a"], ob, g["lse_b"])
        assert np.abs(out - g["out"]).max() <= 1e-2  # reference output is rounded to bf16/f16
        fin = np.isfinite(g["lse"])
        assert np.abs(lse[fin] - g["lse"][fin]).max() <= 1e-5


def test_merge_n_equals_full_softmax():
    """Splitting a request into pieces and merging (stage-2 semantics) is exact."""
    cfg = synth.Config("m", 2, 0, 8, 2, 64, ragged=(90, 200), seed=3)
    w = synth.make_workload(cfg, "cpu")
    full, full_lse, kp, vp, _, _ = oracle_step(w, save_kv=False)
    outs, lses = [], []
    cuts = [0, 17, 64, 10 ** 9]
    q = w.q.reshape(cfg.batch, cfg.num_q_heads, cfg.head_dim)
    for a, b in zip(cuts[:-1], cuts[1:]):
        r2t = w.req_to_token[:, a:].contiguous()
        sl = torch.clamp(torch.clamp(w.seq_lens, max=b) - a, min=0)
        o, l = oracle.decode_attention(q, kp, vp, r2t, w.req_pool_indices, sl, w.sm_scale)
        outs.append(o)
        lses.append(l)
    out, lse = oracle.merge_n(np.stack(outs), np.stack(lses))
    assert np.abs(out - full).max() <= 1e-6 and np.abs(lse - full_lse).max() <= 1e-5


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not mounted (GPU box)")
def test_oracle_vs_live_reference():
    """Where the reference tree is mounted, run its set_kv_buffer + torch_native forward_decode live
    on a fresh seed (not one of the stored goldens)."""
    from tests.golden.make_golden import run_reference_decode

    ref = ref_loader.load()
    cfg = synth.Config("live", 4, 0, 16, 4, 128, page_size=4, ragged=(5, 333), share_groups=2, seed=77)
    w = synth.make_workload(cfg, "cpu")
    out_ref, pool = run_reference_decode(ref, w, w.k_pool, w.v_pool, torch.bfloat16)
    out, _, kp, vp, _, _ = oracle_step(w)
    assert torch.equal(pool.k_buffer[0], kp) and torch.equal(pool.v_buffer[0], vp)
    assert_close_out(out, out_ref.float().numpy().reshape(out.shape), atol=4e-3, rtol=8e-3)


# ----------------------------------------------------------------------------- #
# rotary embedding (producer side of the KV store)
# ----------------------------------------------------------------------------- #
from tests.golden.make_golden import ROPE_CASES  # noqa: E402


@pytest.mark.parametrize("name", sorted(ROPE_CASES))
def test_rope_oracle_vs_reference_golden(name):
    """oracle.rope / oracle.cos_sin_cache against the reference's RotaryEmbedding.forward_native:
    fp32 run (exact arithmetic) within fp32 rounding; bf16 run (cos/sin and every product rounded to
    bf16: three roundings of 2^-9 on |x| up to ~4) within 3 bf16 ulps of the operand magnitude -- the
    reference's own test allows atol = rtol = 1e-2 (sgl-kernel/tests/test_rotary_embedding.py:190-193)."""
    g = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    hs, rot, maxpos, base, neox, n, hq, hkv = (int(x) for x in g["meta"])
    pos = torch.from_numpy(g["positions"])
    cache = oracle.cos_sin_cache(rot, maxpos, float(base))
    assert np.allclose(cache[pos].numpy(), g["cache_rows"], rtol=0, atol=2e-6)
    # the golden rows stand in for the cache so the arithmetic check is independent of libm
    rows = torch.from_numpy(g["cache_rows"])
    idx = torch.arange(n)
    for key, heads in (("q", hq), ("k", hkv)):
        x = _from_bits(g[key + "_in"], torch.bfloat16).view(n, heads, hs)
        got = oracle.rope(idx, x, rows, bool(neox)).reshape(n, heads * hs)
        ref32 = g[key + "_out_f32"].astype(np.float64)
        assert np.abs(got - ref32).max() <= 1e-6 * max(1.0, np.abs(ref32).max())
        ref16 = _from_bits(g[key + "_out_bf16"], torch.bfloat16).float().numpy()
        mag = np.abs(x.float().numpy()).reshape(n, heads, hs).max(axis=-1, keepdims=True)
        bound = np.broadcast_to(3 * 2.0 ** -8 * np.maximum(mag, 1.0), (n, heads, hs)).reshape(n, heads * hs)
        assert np.all(np.abs(got - ref16) <= bound)


# --------------------------------------------------------------------------- #
# extend (prefill with cached prefix)
# --------------------------------------------------------------------------- #
@pytest.mark.parametrize("name", sorted(EXTEND_CASES))
def test_extend_oracle_vs_reference_golden(name):
    """``extend_oracle.extend_attention`` against the reference's own
    ``TorchNativeAttnBackend.forward_extend`` output (bf16/f16-rounded SDPA), and the multi-token KV
    store against the reference's pool rows, bit for bit."""
    cfg, pl, el, causal = EXTEND_CASES[name]
    g = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    w = synth.make_extend_workload(cfg, pl, el, "cpu")
    assert digest_extend(w) == g["digest"].tobytes().decode(), "synthetic inputs drifted: regenerate goldens"
    out, lse, kp, vp = oracle_extend(w, causal=causal)
    dt = w.k_pool.dtype
    assert torch.equal(kp[w.out_cache_loc], _from_bits(g["k_rows_after"], dt))
    assert torch.equal(vp[w.out_cache_loc], _from_bits(g["v_rows_after"], dt))
    ref = _from_bits(g["out_bits"], dt).float().numpy()
    assert_close_out(out, ref, atol=4e-3, rtol=8e-3, what=f"{name}: extend oracle vs reference")
    assert np.isfinite(lse).all()


def test_extend_oracle_mask_properties():
    """Properties the reference's own Triton test asserts (test_triton_attention_kernels.py:189-224):
    a custom mask equal to the causal mask reproduces the causal result; plus: a window wider than
    every prefix changes nothing, and logit_cap -> large approaches the uncapped result."""
    from oracle import extend_oracle

    cfg = synth.Config("m", 3, 0, 8, 2, 64, seed=31)
    pl, el = (9, 0, 40), (17, 30, 5)
    w = synth.make_extend_workload(cfg, pl, el, "cpu")
    base, base_lse, _, _ = oracle_extend(w, causal=True)
    _, _, mi = extend_oracle.extend_indptrs(w.prefix_lens_np, w.extend_lens_np)
    mask = np.zeros(int(mi[-1]), dtype=np.uint8)
    for b in range(3):
        m = np.ones((el[b], pl[b] + el[b]), dtype=np.uint8)
        m[:, pl[b]:] = np.tril(np.ones((el[b], el[b]), dtype=np.uint8))
        mask[mi[b]:mi[b + 1]] = m.reshape(-1)
    got, got_lse, _, _ = oracle_extend(w, causal=True, custom_mask=mask)
    assert np.array_equal(got, base) and np.array_equal(got_lse, base_lse)
    wide, _, _, _ = oracle_extend(w, causal=True, sliding_window=1000)
    assert np.array_equal(wide, base)
    capped, _, _, _ = oracle_extend(w, causal=True, logit_cap=1e4)
    assert np.abs(capped - base).max() < 1e-5
    # window semantics of the reference kernel: query i sees window-list entry j iff i <= j + W
    W = 4
    win, _, _, _ = oracle_extend(w, causal=True, sliding_window=W)
    assert np.abs(win - base).max() > 1e-3   # request 2 (prefix 40 > W + 1) changes
    qo, kv, _ = extend_oracle.extend_indptrs(w.prefix_lens_np, w.extend_lens_np, W)
    assert list(np.diff(kv)) == [min(p, W + 1) for p in pl] and list(np.diff(qo)) == list(el)


def test_e5m2_cast_matches_torch_exhaustively():
    """The oracle's float8_e5m2 cast and dequantisation against torch on all 65 536 bf16 and f16 bit patterns."""
    b = np.arange(65536, dtype=np.uint16)
    for x in (oracle.bf16_bits_to_f32(b), b.view(np.float16).astype(np.float32)):
        ok = ~np.isnan(x)
        ref = torch.from_numpy(x.copy()).to(torch.float8_e5m2).view(torch.uint8).numpy()
        got = oracle.f32_to_e5m2_bits(x)
        assert np.array_equal(ref[ok], got[ok])
        assert ((got[~ok] & 0x7F) == 0x7F).all()
    bb = np.arange(256, dtype=np.uint8)
    ref = torch.from_numpy(bb.copy()).view(torch.float8_e5m2).float().numpy()
    assert np.array_equal(np.nan_to_num(oracle.e5m2_bits_to_f32(bb), nan=7.0), np.nan_to_num(ref, nan=7.0))
