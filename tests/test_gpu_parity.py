"""GPU parity tests: the sm_100a path (through the C-ABI) against the CPU oracle and the
golden vectors produced by the reference itself.

Bars (BASELINE.json north_star): KV store and page indexing bit-exact; attention output within
max-abs 2e-2 / rel 1e-2 of the float64 oracle; LSE within 1e-3.
"""
from __future__ import annotations

import os

import numpy as np
import pytest
import torch

import synth
from tests import harness
from tests.harness import (GOLDEN_DIR, assert_close_out, canon_e4m3_nan, make_layer, make_runner,
                           oracle_step, run_decode)

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _bits(t: torch.Tensor) -> np.ndarray:
    t = t.detach().cpu().contiguous()
    if t.dtype in (torch.bfloat16, torch.float16):
        return t.view(torch.int16).numpy().view(np.uint16)
    if t.dtype in (torch.uint8, torch.float8_e4m3fn, torch.float8_e5m2):
        return t.view(torch.uint8).numpy()
    return t.numpy()


def _pair(cfg):
    return synth.make_workload(cfg, "cpu"), synth.make_workload(cfg, DEV)


def _check_same_inputs(w_cpu, w_gpu):
    # CPU and CUDA generators differ: the GPU workload must carry the CPU data for parity runs
    w_gpu.qkv.copy_(w_cpu.qkv)
    w_gpu.k_pool.copy_(w_cpu.k_pool)
    w_gpu.v_pool.copy_(w_cpu.v_pool)


def _run_case(cfg, logit_cap=0.0, max_ctas=0, min_chunks=0, index_dtype=torch.int64,
              materialize=False, lse_tol=1e-3, q_gain=1.0):
    import b200attn

    w_cpu, w = _pair(cfg)
    if q_gain != 1.0:   # peaked softmax: a few keys carry the row, so P.V errors are not averaged away
        w_cpu.qkv[:, : cfg.num_q_heads * cfg.head_dim] *= q_gain
    _check_same_inputs(w_cpu, w)
    runner = make_runner(w, DEV)
    backend = b200attn.B200AttnBackend(runner, return_lse=True, max_ctas=max_ctas, min_chunks=min_chunks,
                                       materialize_kv_indices=materialize)
    layer = make_layer(w, logit_cap=logit_cap)
    out = run_decode(backend, runner, w, layer=layer, index_dtype=index_dtype)
    torch.cuda.synchronize()
    ref, ref_lse, kp, vp, k_after, v_after = oracle_step(w_cpu, logit_cap=logit_cap)
    c = cfg
    pool = runner.token_to_kv_pool
    # KV store: raw pool bytes, whole pool
    assert np.array_equal(_bits(pool.k_buffer[0]), _bits(kp)), "k pool bytes differ"
    assert np.array_equal(_bits(pool.v_buffer[0]), _bits(vp)), "v pool bytes differ"
    got = out.float().cpu().numpy().reshape(c.batch, c.num_q_heads, c.head_dim)
    err = assert_close_out(got, ref)
    lse = backend.last_lse(c.batch).cpu().numpy()
    lerr = float(np.abs(lse - ref_lse).max())
    assert lerr <= lse_tol, f"lse err {lerr}"
    if materialize:
        from oracle import decode_oracle as oracle

        indptr, indices = oracle.kv_indices(w_cpu.req_to_token, w_cpu.req_pool_indices, w_cpu.seq_lens)
        md = backend.forward_metadata
        assert np.array_equal(md.kv_indptr.cpu().numpy(), indptr)
        assert np.array_equal(md.kv_indices.cpu().numpy(), indices)
    return err, lerr


# ----------------------------------------------------------------------------- #
# golden vectors from the reference
# ----------------------------------------------------------------------------- #
GOLDEN_CASES = ["g1_gqa4_d64_p1", "g2_gqa4_d128_p16_shared", "g3_mha_d128", "g4_gqa8_d128_p4",
                "g5_f16_gqa4_d64", "g7_fp8_gqa4_d128"]


@pytest.mark.parametrize("name", GOLDEN_CASES)
@pytest.mark.parametrize("forced_split", [False, True])
def test_golden_decode(name, forced_split):
    """Same inputs as tests/golden/make_golden.py fed to the reference; compare with its outputs."""
    import b200attn
    from tests.golden.make_golden import CASES, digest

    cfg = CASES[name]
    g = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    w_cpu, w = _pair(cfg)
    assert digest(w_cpu) == g["digest"].tobytes().decode(), "synthetic inputs drifted: regenerate goldens"
    _check_same_inputs(w_cpu, w)
    runner = make_runner(w, DEV)
    kw = dict(max_ctas=3, min_chunks=1) if forced_split else {}
    backend = b200attn.B200AttnBackend(runner, materialize_kv_indices=True, **kw)
    # the reference golden for fp8 is scale-free (dequantised copy, scales 1)
    layer = make_layer(w, with_scales=False)
    fp8 = cfg.kv_dtype.startswith("fp8")
    out = run_decode(backend, runner, w, layer=layer, save_kv_cache=not fp8)
    torch.cuda.synchronize()
    c = cfg
    ref_dt = torch.float16 if c.q_dtype == "f16" else torch.bfloat16
    ref = torch.from_numpy(g["out_bits"].view(np.int16)).view(ref_dt).float().numpy()
    got = out.float().cpu().numpy().reshape(ref.shape)
    # reference output is itself bf16/f16-rounded SDPA: same 2e-2 / 1e-2 bar
    assert_close_out(got, ref, what=f"{name} vs reference torch_native")
    md = backend.forward_metadata
    assert np.array_equal(md.kv_indptr.cpu().numpy(), g["kv_indptr"])
    assert np.array_equal(md.kv_indices.cpu().numpy(), g["kv_indices"])
    if not fp8:
        pool = runner.token_to_kv_pool
        loc = w.out_cache_loc
        assert np.array_equal(_bits(pool.k_buffer[0][loc]), g["k_rows_after"])
        assert np.array_equal(_bits(pool.v_buffer[0][loc]), g["v_rows_after"])


@pytest.mark.parametrize("dname", ["bf16", "f16"])
@pytest.mark.parametrize("tag", ["none", "one", "half_two", "odd"])
def test_golden_fp8_store(dname, tag):
    """fp8 KV store against the reference's set_kv_buffer (CPU => true-division semantics)."""
    import b200attn
    from b200attn import capi

    g = np.load(os.path.join(GOLDEN_DIR, f"g6_fp8_store_{dname}_{tag}.npz"))
    dt = torch.bfloat16 if dname == "bf16" else torch.float16
    k = torch.from_numpy(g["k_in"].view(np.int16)).view(dt).to(DEV)
    v = torch.from_numpy(g["v_in"].view(np.int16)).view(dt).to(DEV)
    loc = torch.from_numpy(g["loc"]).to(DEV)
    kp = torch.zeros(g["k_pool"].shape, dtype=torch.uint8, device=DEV)
    vp = torch.zeros(g["v_pool"].shape, dtype=torch.uint8, device=DEV)
    ks, vs = float(g["k_scale"]), float(g["v_scale"])
    if ks < 0:
        b200attn.store_kv(kp, vp, loc, k, v)
    else:
        # the CPU reference divided by an unrounded fp32 scalar: host floats + true division
        b200attn.store_kv(kp, vp, loc, k, v, ks, vs, scale_mode=capi.SCALE_TRUE_DIV)
    torch.cuda.synchronize()
    assert np.array_equal(canon_e4m3_nan(kp.cpu().numpy()), canon_e4m3_nan(g["k_pool"]))
    assert np.array_equal(canon_e4m3_nan(vp.cpu().numpy()), canon_e4m3_nan(g["v_pool"]))
    # div_ mutates the caller's tensors; NaN payloads may differ between CPU and CUDA rounding
    for got, want in ((k, g["k_after"]), (v, g["v_after"])):
        gb = _bits(got)
        nan = np.isnan(torch.from_numpy(want.view(np.int16)).view(dt).float().numpy())
        assert np.array_equal(gb[~nan], want[~nan])
        assert np.isnan(got.float().cpu().numpy()[nan]).all()


def test_fp8_cast_table_exhaustive():
    """All 65536 bf16 bit patterns through the kernel's cast == torch's own cast (CPU golden and
    torch CUDA on this box)."""
    import b200attn

    table = np.load(os.path.join(GOLDEN_DIR, "g6_cast_table_bf16_e4m3fn.npz"))["table"]
    allbf = torch.arange(65536, dtype=torch.int32).to(torch.int16).view(torch.bfloat16).to(DEV)
    k = allbf.view(512, 1, 128).clone()
    v = allbf.view(512, 1, 128).clone()
    kp = torch.zeros((512, 1, 128), dtype=torch.uint8, device=DEV)
    vp = torch.zeros_like(kp)
    loc = torch.arange(512, dtype=torch.int64, device=DEV)
    b200attn.store_kv(kp, vp, loc, k, v)
    torch.cuda.synchronize()
    assert np.array_equal(kp.cpu().numpy().reshape(-1), table)
    cuda_cast = allbf.to(torch.float8_e4m3fn).view(torch.uint8).cpu().numpy()
    assert np.array_equal(kp.cpu().numpy().reshape(-1), cuda_cast)


@pytest.mark.parametrize("scale", [0.5, 0.3, 1.7, 3.0])
def test_fp8_store_matches_torch_cuda_div(scale):
    """With a Python-float scale the reference's CUDA path is torch's `div_` by a CPU scalar
    (a * (1/b)); with a 0-dim tensor it is a true division.  Match both, bit for bit."""
    import b200attn

    g = torch.Generator(device=DEV).manual_seed(3)
    n, H, D = 64, 4, 128
    for as_tensor in (False, True):
        k = (torch.randn((n, H, D), generator=g, device=DEV) * 30).to(torch.bfloat16)
        v = (torch.randn((n, H, D), generator=g, device=DEV) * 5).to(torch.bfloat16)
        s = torch.tensor(scale, device=DEV) if as_tensor else scale
        k_ref, v_ref = k.clone(), v.clone()
        k_ref.div_(s)
        v_ref.div_(s)
        kp = torch.zeros((n + 1, H, D), dtype=torch.uint8, device=DEV)
        vp = torch.zeros_like(kp)
        loc = torch.randperm(n, device=DEV).to(torch.int64) + 1
        b200attn.store_kv(kp, vp, loc, k, v, s, s)
        torch.cuda.synchronize()
        assert torch.equal(k.view(torch.int16), k_ref.view(torch.int16)), f"div_ mismatch tensor={as_tensor}"
        assert torch.equal(v.view(torch.int16), v_ref.view(torch.int16))
        assert torch.equal(kp[loc], k_ref.to(torch.float8_e4m3fn).view(torch.uint8))
        assert torch.equal(vp[loc], v_ref.to(torch.float8_e4m3fn).view(torch.uint8))


# ----------------------------------------------------------------------------- #
# oracle parity over shapes / layouts
# ----------------------------------------------------------------------------- #
SHAPES = [
    # name, B, Hq, Hkv, D, page, ragged
    ("mha_d64", 4, 8, 8, 64, 1, (1, 300)),
    ("gqa2_d128", 5, 8, 4, 128, 1, (20, 500)),
    ("gqa4_d128", 7, 32, 8, 128, 1, (1, 900)),
    ("gqa8_d128", 3, 64, 8, 128, 16, (100, 1500)),
    ("gqa8_tp8", 6, 8, 1, 128, 1, (1, 2000)),
    ("gqa4_tp4", 6, 8, 2, 128, 64, (60, 1200)),
    ("gqa16_d64", 3, 32, 2, 64, 1, (10, 700)),
    ("gqa12_d128", 2, 12, 1, 128, 1, (300, 600)),
    ("mha32_d128", 2, 32, 32, 128, 1, (50, 400)),
    ("hkv3", 3, 12, 3, 64, 1, (30, 500)),
    ("gqa4_d64_tp8", 5, 4, 1, 64, 1, (1, 900)),      # 128-byte rows: swizzled cp.async stage, 4 rows per copy
    ("gqa8_d64_hkv2", 4, 16, 2, 64, 16, (100, 777)),  # two 128-byte heads per 256-byte row
    # head dims beyond 64 / 128 (the reference kernel takes any Lk: triton_ops/decode_attention.py:205, 425-430;
    # its own shape list has D = 80: test/srt/test_triton_attention_kernels.py:375-388)
    ("gemma_d256", 5, 8, 4, 256, 1, (1, 900)),        # K and V of a chunk in separate pipeline slots
    ("gemma_d256_b70", 70, 16, 8, 256, 16, (1, 300)), # enough requests for the widest rows (4 heads per CTA)
    ("mha_d256", 3, 8, 8, 256, 1, (50, 400)),
    ("gqa8_d256_tp", 4, 8, 1, 256, 1, (30, 1200)),
    ("gqa16_d256", 2, 16, 1, 256, 1, (30, 500)),      # two n8 tiles of heads
    ("phi_d96", 6, 32, 8, 96, 1, (1, 800)),
    ("phi_d96_b80", 80, 32, 8, 96, 4, (1, 200)),
    ("d96_tp8", 5, 4, 1, 96, 1, (10, 700)),           # 192-byte rows: bulk copies (not a power of two)
    ("ref_d80", 3, 32, 8, 80, 1, (1, 600)),
    ("ref_d80_b70", 70, 32, 4, 80, 1, (1, 150)),
    ("d80_hkv1", 4, 6, 1, 80, 16, (17, 500)),
]


@pytest.mark.parametrize("shape", SHAPES, ids=[s[0] for s in SHAPES])
def test_decode_shapes(shape):
    name, B, Hq, Hkv, D, P, rg = shape
    cfg = synth.Config(name, B, 0, Hq, Hkv, D, page_size=P, ragged=rg, seed=100 + len(name))
    _run_case(cfg)


@pytest.mark.parametrize("max_ctas,min_chunks", [(1, 1), (2, 1), (7, 1), (148, 1), (0, 0), (5, 3)])
def test_decode_split_plans(max_ctas, min_chunks):
    """Any split plan gives the same math: force 1..many pieces per request."""
    cfg = synth.Config("split", 6, 0, 32, 8, 128, page_size=16, ragged=(40, 1400), share_groups=2, seed=21)
    _run_case(cfg, max_ctas=max_ctas, min_chunks=min_chunks, materialize=True)


@pytest.mark.parametrize("idx", [torch.int32, torch.int64])
def test_decode_index_dtypes(idx):
    cfg = synth.Config("idx", 5, 0, 16, 4, 64, ragged=(1, 333), seed=22)
    _run_case(cfg, index_dtype=idx, materialize=True)


def test_decode_f16():
    _run_case(synth.Config("f16", 4, 0, 32, 8, 128, kv_dtype="f16", q_dtype="f16", ragged=(5, 800), seed=23))


@pytest.mark.parametrize("q_dtype", ["bf16", "f16"])
@pytest.mark.parametrize("D", [64, 128, 256])
def test_decode_fp8(q_dtype, D):
    cfg = synth.Config("fp8", 5, 0, 32, 8, D, kv_dtype="fp8_e4m3", q_dtype=q_dtype, page_size=16,
                       ragged=(10, 1100), k_scale=0.5, v_scale=2.0, seed=24)
    _run_case(cfg, max_ctas=9, min_chunks=2)


@pytest.mark.parametrize("D,heads,ragged", [(128, (32, 8), (10, 1100)), (64, (32, 8), (1, 600)), (256, (8, 4), (10, 700)),
                                            (128, (8, 1), (10, 1500)), (128, (32, 2), (10, 500))])
def test_decode_fp8_e5m2(D, heads, ragged):
    """``--kv-cache-dtype fp8_e5m2`` (server_args.py:828-833): the store casts like torch's ``.to(float8_e5m2)``
    (pool bytes bit-exact, RNE, overflow to infinity), decode reads the byte as the upper half of an f16."""
    cfg = synth.Config("e5m2", 5, 0, heads[0], heads[1], D, kv_dtype="fp8_e5m2", page_size=4, ragged=ragged,
                       k_scale=0.5, v_scale=2.0, seed=29)
    _run_case(cfg, max_ctas=9, min_chunks=2)


def test_fp8_e5m2_store_matches_torch_cuda():
    """The e5m2 cast against torch-CUDA's own ``.to(torch.float8_e5m2)`` on every bf16 bit pattern (the arithmetic
    ``set_kv_buffer`` runs on the GPU, memory_pool.py:385-391)."""
    import b200attn

    bits = torch.arange(65536, dtype=torch.int32, device=DEV).to(torch.int16)
    x = bits.view(torch.bfloat16).reshape(1024, 1, 64)
    ok = ~torch.isnan(x.float())
    kp = torch.zeros((1024, 1, 64), dtype=torch.uint8, device=DEV).view(torch.float8_e5m2)
    vp = torch.zeros_like(kp)
    loc = torch.arange(1024, device=DEV, dtype=torch.int64)
    b200attn.store_kv(kp, vp, loc, x.clone(), x.clone(), None, None)
    torch.cuda.synchronize()
    ref = x.to(torch.float8_e5m2).view(torch.uint8)
    assert torch.equal(kp.view(torch.uint8)[ok], ref[ok]) and torch.equal(vp.view(torch.uint8)[ok], ref[ok])
    assert bool(((kp.view(torch.uint8)[~ok] & 0x7F) == 0x7F).all())   # NaN stays NaN


def test_decode_fp8_tp8_and_wide_group():
    _run_case(synth.Config("fp8tp8", 4, 0, 8, 1, 128, kv_dtype="fp8_e4m3", ragged=(10, 1500),
                           k_scale=0.7, v_scale=1.3, seed=25))
    _run_case(synth.Config("fp8g16", 2, 0, 32, 2, 128, kv_dtype="fp8_e4m3", ragged=(10, 500),
                           k_scale=1.0, v_scale=1.0, seed=26))


@pytest.mark.parametrize("cap", [30.0, 5.0])
def test_decode_logit_cap(cap):
    cfg = synth.Config("cap", 4, 0, 16, 4, 128, ragged=(10, 600), seed=27)
    _run_case(cfg, logit_cap=cap, max_ctas=6, min_chunks=1)


@pytest.mark.parametrize("kv_dtype", ["bf16", "fp8_e4m3"])
@pytest.mark.parametrize("D", [64, 128, 256])
def test_decode_peaked_softmax(kv_dtype, D):
    """q scaled by 6: the softmax of a row is carried by a handful of keys, so a wrong V row, a dropped
    token or a swapped head-dim column moves the output by far more than the tolerance (with randn q the
    output is an average of hundreds of V rows, |o| ~ 0.05, and such errors could hide under atol)."""
    cfg = synth.Config("peak", 6, 0, 32, 8, D, kv_dtype=kv_dtype, ragged=(100, 1500), page_size=4,
                       k_scale=0.5 if kv_dtype != "bf16" else 1.0, v_scale=2.0 if kv_dtype != "bf16" else 1.0, seed=33)
    _run_case(cfg, q_gain=6.0, max_ctas=17, min_chunks=1, lse_tol=2e-3)


def test_uniform_c1_shape():
    """BASELINE configs[0] (C1: B16, ctx 2048, 32/8 heads, d64, page 1) at full size."""
    _run_case(synth.CONFIGS["C1"])


def test_strided_and_2d_q():
    """q arrives [B, Hq*D] as a view of the fused qkv buffer (models/llama.py:186-187)."""
    import b200attn

    cfg = synth.Config("strided", 4, 0, 32, 8, 128, ragged=(10, 300), seed=28)
    w_cpu, w = _pair(cfg)
    _check_same_inputs(w_cpu, w)
    assert not w.q.is_contiguous() and w.q.dim() == 2
    runner = make_runner(w, DEV)
    backend = b200attn.B200AttnBackend(runner)
    out = run_decode(backend, runner, w)
    assert out.shape == (cfg.batch, cfg.num_q_heads * cfg.head_dim) and out.dtype == torch.bfloat16
    ref, *_ = oracle_step(w_cpu)
    assert_close_out(out.float().cpu().numpy().reshape(ref.shape), ref)


def test_padded_graph_entries_and_cuda_graph():
    """CUDA-graph capture / replay with padded entries (seq_len 1, slot 0, req 0), int32 metadata
    buffers as CudaGraphRunner allocates them (cuda_graph_runner.py:302-306, 720-722)."""
    import b200attn

    cfg = synth.Config("graph", 6, 0, 32, 8, 128, page_size=16, ragged=(40, 900), seed=29)
    w_cpu, w = _pair(cfg)
    _check_same_inputs(w_cpu, w)
    runner = make_runner(w, DEV)
    backend = b200attn.B200AttnBackend(runner)
    max_bs = 8
    backend.init_cuda_graph_state(max_bs, max_bs)
    c = cfg
    # static buffers
    rpi = torch.zeros(max_bs, dtype=torch.int32, device=DEV)
    sl = torch.full((max_bs,), backend.get_cuda_graph_seq_len_fill_value(), dtype=torch.int32, device=DEV)
    loc = torch.zeros(max_bs, dtype=torch.int64, device=DEV)
    qkv = torch.zeros((max_bs, w.qkv.shape[1]), dtype=w.qkv.dtype, device=DEV)
    q, k, v = qkv.split([c.num_q_heads * c.head_dim, c.num_kv_heads * c.head_dim, c.num_kv_heads * c.head_dim], dim=-1)
    layer = make_layer(w)
    fb = b200attn.ForwardBatch(forward_mode=b200attn.ForwardMode.DECODE, batch_size=max_bs,
                               req_pool_indices=rpi, seq_lens=sl, out_cache_loc=loc, seq_lens_sum=max_bs,
                               req_to_token_pool=runner.req_to_token_pool,
                               token_to_kv_pool=runner.token_to_kv_pool, attn_backend=backend)
    backend.init_forward_metadata_capture_cuda_graph(max_bs, max_bs, rpi, sl, None, fb.forward_mode, None)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(2):
            layer(q, k, v, fb)
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out_static = layer(q, k, v, fb)
    # replay with the real batch in the first 6 entries, 2 padded
    B = c.batch
    rpi[:B] = w.req_pool_indices.to(torch.int32)
    sl[:B] = w.seq_lens.to(torch.int32)
    loc[:B] = w.out_cache_loc
    qkv[:B] = w.qkv
    backend.init_forward_metadata_replay_cuda_graph(max_bs, rpi, sl, int(sl.sum().item()), None,
                                                    fb.forward_mode, None, None)
    graph.replay()
    torch.cuda.synchronize()
    ref, _, kp, vp, _, _ = oracle_step(w_cpu)
    got = out_static[:B].float().cpu().numpy().reshape(ref.shape)
    assert_close_out(got, ref)
    pool = runner.token_to_kv_pool
    # real rows stored bit-exactly; slot 0 (dummy page) absorbed the padded writes
    locs = w.out_cache_loc
    assert torch.equal(pool.k_buffer[0][locs].cpu(), kp[w_cpu.out_cache_loc])
    assert torch.equal(pool.v_buffer[0][locs].cpu(), vp[w_cpu.out_cache_loc])
    assert torch.isfinite(out_static.float()).all()


def test_empty_and_len1_requests():
    import b200attn

    cfg = synth.Config("tiny", 5, 0, 8, 2, 64, ragged=(1, 3), seed=30)
    _run_case(cfg)
    # batch 0 is a no-op
    w = synth.make_workload(cfg, DEV)
    runner = make_runner(w, DEV)
    backend = b200attn.B200AttnBackend(runner)
    fb = harness.make_batch(w, runner, backend)
    fb.batch_size = 0
    fb.req_pool_indices = fb.req_pool_indices[:0]
    fb.seq_lens = fb.seq_lens[:0]
    fb.out_cache_loc = fb.out_cache_loc[:0]
    backend.init_forward_metadata(fb)
    out = make_layer(w)(w.q[:0], w.k[:0], w.v[:0], fb)
    assert out.shape == (0, cfg.num_q_heads * cfg.head_dim)


def test_merge_states_golden():
    for dname, dt in (("bf16", torch.bfloat16), ("f16", torch.float16)):
        g = np.load(os.path.join(GOLDEN_DIR, f"g8_merge_{dname}.npz"))
        oa = torch.from_numpy(g["o_a"].view(np.int16)).view(dt).to(DEV)
        ob = torch.from_numpy(g["o_b"].view(np.int16)).view(dt).to(DEV)
        la = torch.from_numpy(g["lse_a"]).to(DEV)
        lb = torch.from_numpy(g["lse_b"]).to(DEV)
        out = torch.empty_like(oa)
        lse = torch.empty_like(la)
        torch.ops.b200attn.merge_states(oa, la, ob, lb, out, lse)
        torch.cuda.synchronize()
        # reference tolerances: 1e-2 for half outputs, 1e-3 for lse (test_merge_state_v2.py:346-370)
        assert np.abs(out.float().cpu().numpy() - g["out"]).max() <= 1e-2
        fin = np.isfinite(g["lse"])
        assert np.abs(lse.cpu().numpy()[fin] - g["lse"][fin]).max() <= 1e-3


def test_merge_states_fp32():
    """float32 states (merge_attn_states.cu:111-122 dispatches float too): the bf16 golden's inputs widened to fp32,
    against the float64 merge of the same values (`merge_state_torch`'s formula, test_merge_state_v2.py:30-60)."""
    g = np.load(os.path.join(GOLDEN_DIR, "g8_merge_bf16.npz"))
    oa = torch.from_numpy(g["o_a"].view(np.int16)).view(torch.bfloat16).float().to(DEV)
    ob = torch.from_numpy(g["o_b"].view(np.int16)).view(torch.bfloat16).float().to(DEV)
    la, lb = torch.from_numpy(g["lse_a"]).to(DEV), torch.from_numpy(g["lse_b"]).to(DEV)
    out, lse = torch.empty_like(oa), torch.empty_like(la)
    torch.ops.b200attn.merge_states(oa, la, ob, lb, out, lse)
    torch.cuda.synchronize()
    a64, b64 = oa.double().cpu().numpy(), ob.double().cpu().numpy()
    la64, lb64 = la.double().cpu().numpy(), lb.double().cpu().numpy()
    la64 = np.where(np.isinf(la64), -np.inf, la64)
    lb64 = np.where(np.isinf(lb64), -np.inf, lb64)
    m = np.maximum(la64, lb64)
    with np.errstate(invalid="ignore"):
        wa, wb = np.exp(la64 - m), np.exp(lb64 - m)
    wa, wb = np.nan_to_num(wa), np.nan_to_num(wb)
    den = wa + wb
    ok = den > 0
    ref = np.where(ok[..., None], (wa[..., None] * a64 + wb[..., None] * b64) / np.where(ok, den, 1.0)[..., None], 0.0)
    got = out.double().cpu().numpy()
    assert np.abs(got[ok] - ref[ok]).max() <= 1e-5
    assert np.abs(lse.double().cpu().numpy()[ok] - (m + np.log(den))[ok]).max() <= 1e-5


def test_errors_are_loud():
    import b200attn
    from b200attn import capi

    # CPU tensors: the op has no CPU kernel
    with pytest.raises((NotImplementedError, RuntimeError)):
        torch.ops.b200attn.merge_states(torch.zeros(1, 1, 8, dtype=torch.bfloat16), torch.zeros(1, 1),
                                        torch.zeros(1, 1, 8, dtype=torch.bfloat16), torch.zeros(1, 1),
                                        torch.zeros(1, 1, 8, dtype=torch.bfloat16), None)
    # head dims without an instantiated kernel are refused, not silently mis-served
    for d in (13, 72, 512):
        with pytest.raises(capi.B200AttnError):
            capi.decode_geometry(32, 8, d, capi.BF16)
    for d in (80, 96):   # fp8 fragments come in pairs of 16-byte units: head dims 64 / 128 / 256
        with pytest.raises(capi.B200AttnError):
            capi.decode_geometry(32, 8, d, capi.FP8_E4M3)


# ----------------------------------------------------------------------------- #
# full-size properties (BASELINE configs[1], C2)
# ----------------------------------------------------------------------------- #
def test_c2_full_size_properties():
    """At B128 / ctx4096 the oracle is too slow; check size-independent properties instead:
    (1) constant V => output equals that constant for every head; (3) the result does not depend on
    the split plan: 148-CTA plan == 37-CTA plan within fp reassociation; (4) a sample of requests
    against the oracle."""
    import b200attn

    cfg = synth.CONFIGS["C2"]
    w = synth.make_workload(cfg, DEV)
    runner = make_runner(w, DEV)
    backend = b200attn.B200AttnBackend(runner, return_lse=True)
    out = run_decode(backend, runner, w).clone()
    lse = backend.last_lse(cfg.batch).clone()
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all()

    # (4) sample 3 requests on the oracle
    from oracle import decode_oracle as oracle

    sel = [0, 57, 127]
    q = w.q.reshape(cfg.batch, cfg.num_q_heads, cfg.head_dim)[sel].cpu()
    rpi = w.req_pool_indices[sel].cpu()
    r2t = w.req_to_token.cpu()
    slots = torch.cat([r2t[int(r), : cfg.ctx] for r in rpi]).to(torch.int64)
    uniq, inv = torch.unique(slots, return_inverse=True)
    kp = w.k_pool[uniq.to(DEV)].cpu()
    vp = w.v_pool[uniq.to(DEV)].cpu()
    r2t_small = inv.to(torch.int32).reshape(len(sel), cfg.ctx)
    ref, ref_lse = oracle.decode_attention(q, kp, vp, r2t_small, torch.arange(len(sel)),
                                           torch.full((len(sel),), cfg.ctx), w.sm_scale)
    got = out.float().cpu().numpy().reshape(cfg.batch, cfg.num_q_heads, cfg.head_dim)[sel]
    assert_close_out(got, ref)
    assert np.abs(lse[sel].cpu().numpy() - ref_lse).max() <= 1e-3

    # (3) different plan, same math
    backend2 = b200attn.B200AttnBackend(runner, return_lse=True, max_ctas=37)
    out2 = run_decode(backend2, runner, w, save_kv_cache=False)
    torch.cuda.synchronize()
    assert (out2.float() - out.float()).abs().max().item() <= 4e-3
    assert (backend2.last_lse(cfg.batch) - lse).abs().max().item() <= 1e-4

    # (1) constant V
    w.v_pool.fill_(0.75)
    out3 = run_decode(backend, runner, w, save_kv_cache=False)
    torch.cuda.synchronize()
    assert (out3.float() - 0.75).abs().max().item() <= 1e-2


# ----------------------------------------------------------------------------- #
# fused RoPE + KV store (K4)
# ----------------------------------------------------------------------------- #
from tests.golden.make_golden import ROPE_CASES  # noqa: E402


def _ulp_close(got: torch.Tensor, ref64: np.ndarray, dtype: torch.dtype):
    """|got - ref| <= 1 ulp of `dtype` at |ref| (one rounding from fp32 arithmetic)."""
    eps = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
    g = got.float().cpu().numpy().astype(np.float64)
    bound = eps * np.abs(ref64) + 1e-30 + 2.0 ** -24   # f16 subnormal floor
    bad = np.abs(g - ref64) > bound
    assert not bad.any(), f"{bad.sum()} elements off by more than 1 ulp; worst {np.abs(g - ref64).max()}"


@pytest.mark.parametrize("name", sorted(ROPE_CASES))
def test_rope_golden(name):
    """RoPE only (save_kv = 0) on the reference's own test shapes, against the fp32 run of the
    reference's forward_native."""
    import b200attn  # noqa: F401

    g = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    hs, rot, maxpos, base, neox, n, hq, hkv = (int(x) for x in g["meta"])
    rows = torch.from_numpy(g["cache_rows"]).to(DEV)
    bf = lambda a: torch.from_numpy(a.view(np.int16).copy()).view(torch.bfloat16)  # noqa: E731
    q = bf(g["q_in"]).view(n, hq, hs).to(DEV)
    k = bf(g["k_in"]).view(n, hkv, hs).to(DEV)
    for pos_dtype in (torch.int64, torch.int32):
        q1, k1 = q.clone(), k.clone()
        torch.ops.b200attn.rope_kv_store(q1, k1, None, torch.arange(n, device=DEV, dtype=pos_dtype), rows,
                                         bool(neox), None, None, None, None, None, None, None, 0)
        torch.cuda.synchronize()
        _ulp_close(q1.view(n, -1), g["q_out_f32"].astype(np.float64), torch.bfloat16)
        _ulp_close(k1.view(n, -1), g["k_out_f32"].astype(np.float64), torch.bfloat16)


@pytest.mark.parametrize("kv_dtype,q_dtype", [("bf16", "bf16"), ("f16", "f16"), ("fp8_e4m3", "bf16")])
@pytest.mark.parametrize("neox", [True, False])
def test_rope_kv_store_fused(kv_dtype, q_dtype, neox):
    """Fused launch == oracle RoPE (1 ulp) followed by the bit-exact store of the rotated k and v;
    q / k / v views of one fused qkv buffer (strided), scales as 0-dim CUDA tensors for fp8."""
    import b200attn
    from oracle import decode_oracle as oracle

    cfg = synth.Config("rope", 7, 0, 16, 4, 128, kv_dtype=kv_dtype, q_dtype=q_dtype, page_size=4,
                       ragged=(3, 90), k_scale=0.5, v_scale=2.0, seed=31)
    w_cpu, w = _pair(cfg)
    _check_same_inputs(w_cpu, w)
    c = cfg
    B, Hq, Hkv, D = c.batch, c.num_q_heads, c.num_kv_heads, c.head_dim
    cache = oracle.cos_sin_cache(D, 4096, 10000.0)
    pos = (w_cpu.seq_lens - 1).to(torch.int64)
    q, k, v = (t.view(B, -1, D) for t in (w.q, w.k, w.v))
    fp8 = kv_dtype == "fp8_e4m3"
    kt = torch.tensor(c.k_scale, dtype=torch.float32, device=DEV) if fp8 else None
    vt = torch.tensor(c.v_scale, dtype=torch.float32, device=DEV) if fp8 else None
    b200attn.rope_store_kv(q, k, v, pos.to(DEV), cache.to(DEV), neox, w.out_cache_loc, w.k_pool, w.v_pool, kt, vt)
    torch.cuda.synchronize()

    dt = w_cpu.q.dtype
    q_ref = oracle.rope(pos, w_cpu.q.reshape(B, Hq, D), cache, neox)
    k_ref = oracle.rope(pos, w_cpu.k.reshape(B, Hkv, D), cache, neox)
    _ulp_close(q, q_ref, dt)
    if not fp8:
        _ulp_close(k, k_ref, dt)
    # store parity: bit equality end to end against an fp32 restatement of the kernel arithmetic
    # (products and sums each rounded to fp32, one rounding to the source dtype)
    def rot32(x):
        xf = x.float().numpy().astype(np.float32)
        cs = cache.numpy()[pos.numpy()]
        co, si = cs[:, None, : D // 2], cs[:, None, D // 2:]
        o = xf.copy()
        if neox:
            a, b = xf[..., : D // 2], xf[..., D // 2:]
            o[..., : D // 2] = (a * co).astype(np.float32) - (b * si).astype(np.float32)
            o[..., D // 2:] = (b * co).astype(np.float32) + (a * si).astype(np.float32)
        else:
            a, b = xf[..., 0::2], xf[..., 1::2]
            o[..., 0::2] = (a * co).astype(np.float32) - (b * si).astype(np.float32)
            o[..., 1::2] = (b * co).astype(np.float32) + (a * si).astype(np.float32)
        return torch.from_numpy(o).to(dt)

    k_rot = rot32(w_cpu.k.reshape(B, Hkv, D))
    q_rot = rot32(w_cpu.q.reshape(B, Hq, D))
    assert np.array_equal(_bits(q), _bits(q_rot)), "rotated q differs from the fp32 restatement"
    kp, vp = w_cpu.k_pool.clone(), w_cpu.v_pool.clone()
    k_after, v_after = k_rot.clone(), w_cpu.v.reshape(B, Hkv, D).clone()
    oracle.kv_store(kp, vp, w_cpu.out_cache_loc, k_after, v_after, c.k_scale if fp8 else None,
                    c.v_scale if fp8 else None, oracle.SCALE_TRUE_DIV, True)
    assert np.array_equal(_bits(w.k_pool), _bits(kp)), "k pool bytes differ"
    assert np.array_equal(_bits(w.v_pool), _bits(vp)), "v pool bytes differ"
    # in-place semantics: k holds rot(k) (divided by the scale for fp8, like div_), v likewise
    assert np.array_equal(_bits(k), _bits(k_after))
    assert np.array_equal(_bits(v), _bits(v_after))


# ----------------------------------------------------------------------------- #
# fused merge == separate stages; dynamic ranges; plan reuse across launches
# ----------------------------------------------------------------------------- #
@pytest.mark.parametrize("dyn", [-1, 0, 50])
@pytest.mark.parametrize("tp,batch", [(1, 70), (1, 37), (8, 37)])
def test_fused_merge_equals_two_stage(dyn, tp, batch, monkeypatch):
    """phases 0 (one launch, in-kernel merge) must equal phases 1 + 2 (K2 then K3) bit for bit, for
    static-only plans, the default dynamic tail and a 50 % dynamic tail; the same plan is then reused
    for several launches (layers) to exercise the self-resetting ticket / arrival counters."""
    import b200attn
    from b200attn import capi, ops
    import ctypes as C

    cfg = synth.shard(synth.Config("fm", batch, 0, 32, 8, 128, page_size=1, ragged=(1, 3000), seed=41), tp)
    w = synth.make_workload(cfg, DEV)
    runner = make_runner(w, DEV)
    backend = b200attn.B200AttnBackend(runner, return_lse=True)
    B, Hq, D = cfg.batch, cfg.num_q_heads, cfg.head_dim
    # plan with an explicit dynamic share through the C-ABI
    a = capi.PlanArgs()
    a.req_to_token = w.req_to_token.data_ptr()
    a.req_to_token_stride = w.req_to_token.stride(0)
    a.req_pool_indices = w.req_pool_indices.data_ptr()
    a.seq_lens = w.seq_lens.data_ptr()
    a.index_dtype = capi.I64
    a.batch = B
    a.num_q_heads, a.num_kv_heads, a.head_dim, a.kv_dtype = Hq, cfg.num_kv_heads, D, capi.BF16
    a.plan = backend.plan_buf.data_ptr()
    a.plan_bytes = backend.plan_buf.numel()
    a.dynamic_percent = dyn
    a.dynamic_range_chunks = 2 if dyn > 0 else 0
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    capi.check("plan", capi.lib().b200attn_plan(C.byref(a), stream))
    q3 = w.q.view(B, Hq, D)
    # batch 70: 8 heads per CTA, one token slice; batch 37: a small batch takes 4 heads per CTA and two token
    # slices (ABI 6); tp 8: one head per CTA, eight slices
    ts = capi.decode_geometry(Hq, cfg.num_kv_heads, D, capi.BF16, batch=B)["token_slices"]
    if os.environ.get("B200ATTN_NO_NARROW") != "1":   # (the switch keeps the widest rows for every batch size)
        assert ts == {(1, 70): 1, (1, 37): 2, (8, 37): 8}[(tp, batch)]

    def run(phases, o, lse):
        torch.ops.b200attn.decode(q3, w.k_pool, w.v_pool, o, lse, backend.plan_buf, w.req_to_token,
                                  backend.workspace, w.sm_scale, 0.0, 1.0, 1.0, phases)

    o2 = torch.zeros((B, Hq, D), dtype=w.q.dtype, device=DEV)
    l2 = torch.zeros(B * Hq, dtype=torch.float32, device=DEV)
    run(1, o2, l2)
    run(2, o2, l2)
    torch.cuda.synchronize()
    for it in range(4):   # the plan (tickets, arrival counters) is reused, like layers of one forward
        o1 = torch.full((B, Hq, D), float("nan"), dtype=w.q.dtype, device=DEV)
        l1 = torch.full((B * Hq,), float("nan"), dtype=torch.float32, device=DEV)
        run(0, o1, l1)
        torch.cuda.synchronize()
        if ts == 1:
            assert torch.equal(o1, o2), f"launch {it}: fused merge differs from K2 + K3"
            assert torch.equal(l1, l2)
        else:
            # token slices (ts > 1): in the one-launch form a request that one CTA holds whole is merged in
            # shared memory from the unnormalised slice states, not through the workspace -- the same sum
            # in another order, so equal to rounding (a bf16 ulp is 2^-8), and bit-stable across launches
            torch.testing.assert_close(o1.float(), o2.float(), rtol=1.6e-2, atol=2e-3)
            torch.testing.assert_close(l1, l2, rtol=0, atol=1e-5)
            if it == 0:
                o_first, l_first = o1, l1
            else:
                assert torch.equal(o1, o_first) and torch.equal(l1, l_first), f"launch {it} differs from launch 0"
    # and against the oracle on a few requests
    from oracle import decode_oracle as oracle

    ref, ref_lse = oracle.decode_attention(q3.cpu(), w.k_pool.cpu(), w.v_pool.cpu(), w.req_to_token.cpu(),
                                           w.req_pool_indices.cpu(), w.seq_lens.cpu(), w.sm_scale)
    assert_close_out(o1.float().cpu().numpy(), ref)
    assert np.abs(l1.view(B, Hq).cpu().numpy() - ref_lse).max() <= 1e-3


# ----------------------------------------------------------------------------- #
# sliding-window layers (interleaved with full-attention layers)
# ----------------------------------------------------------------------------- #
@pytest.mark.parametrize("window,D,heads", [(1, 128, (16, 4)), (16, 128, (16, 4)), (127, 128, (16, 4)), (4096, 128, (16, 4)),
                                            (64, 256, (8, 4)), (511, 256, (16, 8))])   # the last two: Gemma-shaped
def test_sliding_window_layers(window, D, heads):
    """A window layer attends to the last min(S, W + 1) tokens (triton_backend.py:927-956), a full layer
    of the same forward to everything; window_kv_indptr / window_kv_indices are bit-exact."""
    import b200attn
    from oracle import decode_oracle as oracle

    cfg = synth.Config("swa", 9, 0, heads[0], heads[1], D, page_size=4, ragged=(1, 700), seed=51)
    w_cpu, w = _pair(cfg)
    _check_same_inputs(w_cpu, w)
    runner = make_runner(w, DEV)
    runner.sliding_window_size = window
    backend = b200attn.B200AttnBackend(runner, return_lse=True, materialize_kv_indices=True, max_ctas=11)
    c = cfg
    qc = w_cpu.q.reshape(c.batch, c.num_q_heads, c.head_dim)
    for layer_window in (window, -1):
        layer = make_layer(w)
        layer.sliding_window_size = layer_window
        out = run_decode(backend, runner, w, layer=layer, save_kv_cache=False)
        torch.cuda.synchronize()
        ref, ref_lse = oracle.decode_attention(qc, w_cpu.k_pool, w_cpu.v_pool, w_cpu.req_to_token,
                                               w_cpu.req_pool_indices, w_cpu.seq_lens, w_cpu.sm_scale,
                                               sliding_window=max(layer_window, 0))
        assert_close_out(out.float().cpu().numpy().reshape(ref.shape), ref)
        assert np.abs(backend.last_lse(c.batch).cpu().numpy() - ref_lse).max() <= 1e-3
    md = backend.forward_metadata
    indptr, indices = oracle.kv_indices(w_cpu.req_to_token, w_cpu.req_pool_indices, w_cpu.seq_lens, window)
    assert np.array_equal(md.window_kv_indptr.cpu().numpy(), indptr)
    assert np.array_equal(md.window_kv_indices.cpu().numpy()[: indptr[-1]], indices)
    indptr, indices = oracle.kv_indices(w_cpu.req_to_token, w_cpu.req_pool_indices, w_cpu.seq_lens)
    assert np.array_equal(md.kv_indptr.cpu().numpy(), indptr)
    assert np.array_equal(md.kv_indices.cpu().numpy(), indices)


# ----------------------------------------------------------------------------- #
# hybrid SWAKVPool: window layers live in a second pool behind full_to_swa_index_mapping
# ----------------------------------------------------------------------------- #
@pytest.mark.parametrize("kv_dtype,D", [("bf16", 128), ("fp8_e4m3", 128), ("bf16", 256)])
def test_swa_kv_pool_decode(kv_dtype, D):
    """``SWAKVPool`` (memory_pool.py:420-553): layer 0 is a full-attention layer in ``full_kv_pool``, layer 1 a
    sliding-window layer in ``swa_kv_pool``, whose slot of a token is ``full_to_swa_index_mapping[full slot]``.
    The new token's K/V must land at the translated slot of the swa pool only, and the window layer must attend
    over the translated slots (the reference translates a materialised kv_indices list,
    flashinfer_backend.py:754, 1001; here the decode kernel maps the ids it reads from req_to_token)."""
    import b200attn
    from oracle import decode_oracle as oracle

    W = 200
    fp8 = kv_dtype != "bf16"
    cfg = synth.Config("swapool", 9, 0, 8, 4, D, kv_dtype=kv_dtype, ragged=(1, 700), seed=77,
                       k_scale=0.5 if fp8 else 1.0, v_scale=2.0 if fp8 else 1.0)
    w_cpu, w = _pair(cfg)
    _check_same_inputs(w_cpu, w)
    c = cfg
    runner = make_runner(w, DEV)
    n = w.pool_size
    pool = b200attn.SWAKVPool(n, n, runner.token_to_kv_pool.dtype, c.num_kv_heads, D, [1], [0], False, DEV)
    gen = torch.Generator().manual_seed(5)
    perm = torch.randperm(n + 1, generator=gen)                       # a slot of the swa pool for every full slot
    pool.full_to_swa_index_mapping = perm.to(DEV)
    pool.full_kv_pool.k_buffer[0], pool.full_kv_pool.v_buffer[0] = w.k_pool.clone(), w.v_pool.clone()
    pool.swa_kv_pool.k_buffer[0][perm.to(DEV)] = w.k_pool
    pool.swa_kv_pool.v_buffer[0][perm.to(DEV)] = w.v_pool
    runner.token_to_kv_pool = pool
    runner.sliding_window_size = W
    backend = b200attn.B200AttnBackend(runner, return_lse=True, max_ctas=9)
    qc = w_cpu.q.reshape(c.batch, c.num_q_heads, D)
    kp, vp = w_cpu.k_pool.clone(), w_cpu.v_pool.clone()
    oracle.kv_store(kp, vp, w_cpu.out_cache_loc, w_cpu.k.reshape(c.batch, c.num_kv_heads, D).clone(),
                    w_cpu.v.reshape(c.batch, c.num_kv_heads, D).clone(), c.k_scale if fp8 else None,
                    c.v_scale if fp8 else None)
    loc = w.out_cache_loc.long()
    for layer_id, window in ((1, W), (0, -1)):
        layer = make_layer(w)
        layer.layer_id = layer_id
        layer.sliding_window_size = window
        qkv0 = w.qkv.clone()
        out = run_decode(backend, runner, w, layer=layer, save_kv_cache=True)
        torch.cuda.synchronize()
        w.qkv.copy_(qkv0)    # an fp8 store divides k / v in place
        ref, ref_lse = oracle.decode_attention(qc, kp, vp, w_cpu.req_to_token, w_cpu.req_pool_indices, w_cpu.seq_lens,
                                               w_cpu.sm_scale, k_scale=c.k_scale if fp8 else 1.0,
                                               v_scale=c.v_scale if fp8 else 1.0, sliding_window=max(window, 0))
        assert_close_out(out.float().cpu().numpy().reshape(ref.shape), ref)
        assert np.abs(backend.last_lse(c.batch).cpu().numpy() - ref_lse).max() <= 1e-3
        # the new token's rows: bit-exact, in the right pool at the right slot
        sub = pool.swa_kv_pool if layer_id == 1 else pool.full_kv_pool
        at = pool.full_to_swa_index_mapping[loc] if layer_id == 1 else loc
        assert torch.equal(sub.k_buffer[0][at].cpu().view(torch.uint8), kp[w_cpu.out_cache_loc.long()].view(torch.uint8))
        assert torch.equal(sub.v_buffer[0][at].cpu().view(torch.uint8), vp[w_cpu.out_cache_loc.long()].view(torch.uint8))
