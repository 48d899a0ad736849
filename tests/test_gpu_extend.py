"""GPU parity tests of the native extend (prefill with cached prefix) path, through the C-ABI:
``B200AttnBackend.forward_extend`` -> ``torch.ops.b200attn.extend`` against the CPU oracle
(``oracle/extend_oracle.py``) and the goldens produced by the reference's own
``TorchNativeAttnBackend.forward_extend``.

Bars: KV store bit-exact on the raw pool bytes; qo_indptr / kv_indptr / mask_indptr bit-exact;
output within max-abs 2e-2 / rel 1e-2 of the float64 oracle; LSE within 1e-3.
"""
from __future__ import annotations

import os
import types

import numpy as np
import pytest
import torch

import synth
from tests.golden.make_golden import EXTEND_CASES
from tests.harness import (GOLDEN_DIR, assert_close_out, make_extend_runner, oracle_extend, run_extend)

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _bits(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().contiguous().view(torch.int16).numpy().view(np.uint16)


def _pair(cfg, pl, el):
    w_cpu = synth.make_extend_workload(cfg, pl, el, "cpu")
    w = synth.make_extend_workload(cfg, pl, el, DEV)
    w.qkv.copy_(w_cpu.qkv)   # CPU and CUDA generators differ: both sides get the CPU data
    w.k_pool.copy_(w_cpu.k_pool)
    w.v_pool.copy_(w_cpu.v_pool)
    return w_cpu, w


def _layer(w, causal=True, logit_cap=0.0, window=-1):
    import b200attn

    c = w.cfg
    at = b200attn.AttentionType.DECODER if causal else b200attn.AttentionType.ENCODER_ONLY
    layer = b200attn.AttnLayer(c.num_q_heads, c.head_dim, w.sm_scale, c.num_kv_heads, layer_id=0,
                               logit_cap=logit_cap, sliding_window_size=window, attn_type=at)
    if c.kv_dtype.startswith("fp8"):   # per-tensor scales the way RadixAttention carries them (radix_attention.py:61-82)
        dev = w.k_pool.device
        layer.k_scale = torch.tensor(c.k_scale, dtype=torch.float32, device=dev)
        layer.v_scale = torch.tensor(c.v_scale, dtype=torch.float32, device=dev)
        layer.k_scale_float, layer.v_scale_float = c.k_scale, c.v_scale
    return layer


def _run(cfg, pl, el, causal=True, logit_cap=0.0, window=-1, index_dtype=torch.int64, check_pool=True):
    import b200attn

    w_cpu, w = _pair(cfg, pl, el)
    runner = make_extend_runner(w, DEV, sliding_window_size=window if window > 0 else None)
    backend = b200attn.B200AttnBackend(runner, return_lse=True)
    out = run_extend(backend, runner, w, _layer(w, causal, logit_cap, window), index_dtype=index_dtype)
    torch.cuda.synchronize()
    ref, ref_lse, kp, vp = oracle_extend(w_cpu, causal=causal, logit_cap=logit_cap, sliding_window=window)
    c = cfg
    if check_pool:
        pool = runner.token_to_kv_pool
        if kp.dtype == torch.float8_e5m2:   # (the oracle tells the fp8 kind by the view; the pool stores bytes)
            kp, vp = kp.view(torch.uint8), vp.view(torch.uint8)
        assert torch.equal(pool.k_buffer[0].cpu(), kp) and torch.equal(pool.v_buffer[0].cpu(), vp), \
            "kv_store: pool bytes differ from the oracle"
    got = out.float().cpu().numpy().reshape(ref.shape)
    err = assert_close_out(got, ref)
    lse = backend.lse_extend(w.tokens).view(w.tokens, c.num_q_heads).cpu().numpy()
    fin = np.isfinite(ref_lse)
    assert np.array_equal(np.isfinite(lse), fin)
    assert np.abs(lse[fin] - ref_lse[fin]).max() <= 1e-3, np.abs(lse[fin] - ref_lse[fin]).max()
    md = backend.extend_metadata
    from oracle import extend_oracle
    qo, kv, _ = extend_oracle.extend_indptrs(pl, el)
    assert np.array_equal(md.qo_indptr.cpu().numpy(), qo) and np.array_equal(md.kv_indptr.cpu().numpy(), kv)
    if window > 0:
        _, wkv, _ = extend_oracle.extend_indptrs(pl, el, window)
        assert np.array_equal(md.window_kv_indptr.cpu().numpy(), wkv)
    return err, got, w_cpu


@pytest.mark.parametrize("name", sorted(EXTEND_CASES))
def test_golden_extend(name):
    """The reference's own torch_native forward_extend output (bf16/f16) on identical inputs."""
    cfg, pl, el, causal = EXTEND_CASES[name]
    g = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    _, got, w_cpu = _run(cfg, pl, el, causal=causal)
    dt = w_cpu.k_pool.dtype
    ref = torch.from_numpy(g["out_bits"].view(np.int16).copy()).view(dt).float().numpy()
    assert_close_out(got, ref, what=f"{name}: CUDA vs reference torch_native")


# (name, Hq, Hkv, D, page, prefix_lens, extend_lens)
SHAPES = [
    ("mha_d128", 4, 4, 128, 1, (0, 33, 200), (130, 64, 1)),
    ("gqa2_d64", 8, 4, 64, 1, (7, 0), (129, 65)),
    ("gqa3_d128", 6, 2, 128, 1, (64, 5), (31, 70)),
    ("gqa4_d128_long_extend", 8, 2, 128, 16, (100, 0), (700, 257)),
    ("gqa4_d128_long_prefix", 32, 8, 128, 1, (1500, 64), (3, 40)),
    ("gqa5_d128", 10, 2, 128, 1, (17,), (90,)),
    ("gqa7_d128", 28, 4, 128, 1, (40, 0), (50, 20)),
    ("gqa8_d128", 64, 8, 128, 1, (300,), (140,)),
    ("gqa16_d64", 32, 2, 64, 4, (10, 90), (33, 8)),
    ("gqa4_d64_long", 32, 8, 64, 1, (300, 0, 1500), (140, 257, 3)),
    ("mha_d64_tile_edges", 2, 2, 64, 16, (0, 0, 127, 128, 129), (128, 129, 1, 256, 257)),
    ("tile_edges", 8, 2, 128, 1, (0, 0, 0, 0, 63, 64, 65), (32, 33, 64, 65, 1, 1, 1)),
    ("many_short", 8, 2, 64, 1, tuple(range(0, 24)), tuple(1 + (i * 7) % 13 for i in range(24))),
    ("gemma_d256", 8, 4, 256, 1, (0, 33, 300), (130, 64, 1)),         # head_dim 256: the mma.sync kernel
    ("gemma_d256_gqa8", 16, 2, 256, 16, (100, 0), (257, 70)),
    # head dims 80 / 96 (the decode kernel takes them too; test/srt/test_triton_attention_kernels.py:283-291 lists
    # D = 80): 10 / 12 chunks per pool row, ragged loader passes, odd k-step count at 80
    ("phi_d96_gqa3", 12, 4, 96, 1, (0, 33, 300), (130, 64, 1)),
    ("phi_d96_mha_long", 4, 4, 96, 16, (700, 0), (257, 129)),
    ("d80_gqa4", 16, 4, 80, 1, (0, 65, 1500), (129, 64, 3)),
    ("d80_mha_tile_edges", 2, 2, 80, 1, (0, 0, 63, 64, 65), (64, 65, 1, 127, 128)),
]


@pytest.mark.parametrize("shape", SHAPES, ids=[s[0] for s in SHAPES])
def test_extend_shapes(shape):
    name, hq, hkv, d, page, pl, el = shape
    _run(synth.Config(name, len(pl), 0, hq, hkv, d, page_size=page, seed=101), pl, el)


@pytest.mark.parametrize("d", [80, 96])
def test_extend_d80_d96_options(d):
    """head dims 80 / 96 through the options of the path: logit cap, sliding window, non-causal, f16."""
    _run(synth.Config(f"d{d}_cap", 2, 0, 8, 2, d, seed=131), (100, 0), (70, 129), logit_cap=30.0)
    _run(synth.Config(f"d{d}_win", 2, 0, 8, 4, d, seed=132), (300, 20), (64, 90), window=64)
    _run(synth.Config(f"d{d}_noncausal_f16", 2, 0, 4, 4, d, kv_dtype="f16", q_dtype="f16", seed=133), (10, 0), (33, 65),
         causal=False)


def test_extend_f16_and_int32_indices():
    cfg = synth.Config("f16", 2, 0, 8, 2, 128, kv_dtype="f16", q_dtype="f16", seed=102)
    _run(cfg, (20, 0), (70, 33), index_dtype=torch.int32)


# fp8-e4m3 pools (C3's cache dtype): new tokens are quantised by the store (bit-exact bytes), the kernel
# reads k_scale * K / v_scale * V (FlashInfer's paged-prefill semantics, flashinfer_backend.py:560-567).
FP8_SHAPES = [
    ("fp8_gqa4_d128", 32, 8, 128, 1, "bf16", (300, 0, 64), (140, 257, 1), 1.0, 1.0),
    ("fp8_gqa4_d128_scaled", 8, 2, 128, 16, "bf16", (1500, 64), (3, 400), 0.5, 2.0),
    ("fp8_gqa2_d64_scaled", 8, 4, 64, 1, "bf16", (7, 0, 200), (129, 65, 2), 0.5, 2.0),
    ("fp8_f16q_d128", 8, 2, 128, 1, "f16", (20, 0), (70, 133), 2.0, 0.25),
    ("fp8_f16q_d64", 32, 2, 64, 4, "f16", (10, 90), (33, 8), 1.0, 1.0),
    ("fp8_tile_edges", 8, 2, 128, 1, "bf16", (0, 0, 127, 128, 129), (128, 129, 1, 1, 1), 0.5, 2.0),
    ("fp8_gemma_d256", 8, 4, 256, 1, "bf16", (40, 0, 200), (90, 65, 2), 0.5, 2.0),
]


@pytest.mark.parametrize("shape", FP8_SHAPES, ids=[s[0] for s in FP8_SHAPES])
def test_extend_fp8_pool(shape):
    name, hq, hkv, d, page, qdt, pl, el, ks, vs = shape
    cfg = synth.Config(name, len(pl), 0, hq, hkv, d, kv_dtype="fp8_e4m3", q_dtype=qdt, page_size=page,
                       k_scale=ks, v_scale=vs, seed=601)
    _run(cfg, pl, el)


@pytest.mark.parametrize("shape", [FP8_SHAPES[1], FP8_SHAPES[2], FP8_SHAPES[4], FP8_SHAPES[6]], ids=lambda s: s[0])
def test_extend_fp8_e5m2_pool(shape):
    """The same shapes over a float8_e5m2 pool (tcgen05 kernels d128, mma.sync kernel d64 / d256)."""
    name, hq, hkv, d, page, qdt, pl, el, ks, vs = shape
    cfg = synth.Config(name + "_e5m2", len(pl), 0, hq, hkv, d, kv_dtype="fp8_e5m2", q_dtype=qdt, page_size=page,
                       k_scale=ks, v_scale=vs, seed=611)
    _run(cfg, pl, el)


def test_extend_fp8_e5m2_verify_smallq():
    """TARGET_VERIFY-shaped batch (small-q kernels) over an e5m2 pool."""
    cfg = synth.Config("e5m2_verify", 6, 0, 32, 8, 128, kv_dtype="fp8_e5m2", k_scale=0.5, v_scale=2.0, seed=612)
    _run(cfg, (700, 300, 1200, 64, 129, 511), (8, 8, 8, 8, 8, 8))


def test_extend_fp8_window_and_cap():
    cfg = synth.Config("fp8_win", 2, 0, 8, 2, 128, kv_dtype="fp8_e4m3", k_scale=0.5, v_scale=2.0, seed=602)
    _run(cfg, (150, 10), (40, 70), window=16)
    _run(cfg, (150, 10), (40, 70), logit_cap=5.0)


def test_extend_f16_d64():
    cfg = synth.Config("f16d64", 2, 0, 16, 4, 64, kv_dtype="f16", q_dtype="f16", seed=106)
    _run(cfg, (200, 0), (70, 133))


def test_extend_noncausal():
    _run(synth.Config("enc", 2, 0, 8, 2, 128, seed=103), (0, 12), (100, 77), causal=False)


@pytest.mark.parametrize("cap", [30.0, 5.0])
def test_extend_logit_cap(cap):
    _run(synth.Config("cap", 2, 0, 8, 2, 64, seed=104), (30, 0), (70, 40), logit_cap=cap)


@pytest.mark.parametrize("window", [3, 16, 200])
def test_extend_sliding_window(window):
    """The reference kernel's window rule (extend_attention.py:179-184) over the windowed prefix list
    (triton_backend.py:927-956)."""
    _run(synth.Config("win", 3, 0, 8, 2, 128, seed=105), (150, 10, 0), (40, 70, 20), window=window)


def _verify_case(n_draft, pl, seed, mask_fn, kv_dtype="bf16"):
    """TARGET_VERIFY: uniform extend = n_draft tokens per request, visibility from
    spec_info.custom_mask (triton_backend.py:227-263)."""
    import b200attn
    from oracle import extend_oracle

    cfg = synth.Config("verify", len(pl), 0, 32, 8, 128, kv_dtype=kv_dtype, k_scale=0.5 if kv_dtype != "bf16" else 1.0,
                       v_scale=2.0 if kv_dtype != "bf16" else 1.0, seed=seed)
    el = (n_draft,) * len(pl)
    w_cpu, w = _pair(cfg, pl, el)
    _, _, mi = extend_oracle.extend_indptrs(pl, el)
    rng = np.random.default_rng(seed)
    mask = np.zeros(int(mi[-1]), dtype=np.uint8)
    for b in range(len(pl)):
        mask[mi[b]:mi[b + 1]] = mask_fn(rng, n_draft, pl[b]).reshape(-1)
    runner = make_extend_runner(w, DEV, num_draft_tokens=n_draft)
    backend = b200attn.B200AttnBackend(runner, return_lse=True)
    spec = types.SimpleNamespace(custom_mask=torch.from_numpy(mask).to(DEV).view(torch.bool), draft_token_num=n_draft)
    out = run_extend(backend, runner, w, _layer(w), mode=b200attn.ForwardMode.TARGET_VERIFY, spec_info=spec)
    torch.cuda.synchronize()
    assert np.array_equal(backend.extend_metadata.mask_indptr.cpu().numpy(), mi)
    ref, ref_lse, _, _ = oracle_extend(w_cpu, custom_mask=mask)
    got = out.float().cpu().numpy().reshape(ref.shape)
    assert_close_out(got, ref)
    lse = backend.lse_extend(w.tokens).view(w.tokens, cfg.num_q_heads).cpu().numpy()
    fin = np.isfinite(ref_lse)
    assert np.array_equal(np.isfinite(lse), fin) and np.abs(lse[fin] - ref_lse[fin]).max() <= 1e-3
    return got, w_cpu


def test_target_verify_tree_mask():
    def tree(rng, n, p):
        m = np.ones((n, p + n), dtype=np.uint8)
        t = np.tril(rng.integers(0, 2, size=(n, n), dtype=np.uint8))
        np.fill_diagonal(t, 1)     # every draft token sees itself
        m[:, p:] = t
        return m
    _verify_case(8, (100, 37, 1, 260), 201, tree)


def _tree(rng, n, p):
    m = np.ones((n, p + n), dtype=np.uint8)
    t = np.tril(rng.integers(0, 2, size=(n, n), dtype=np.uint8))
    np.fill_diagonal(t, 1)
    m[:, p:] = t
    return m


def test_target_verify_fp8_pool():
    _verify_case(8, (100, 37, 1, 260), 204, _tree, kv_dtype="fp8_e4m3")


@pytest.mark.parametrize("pl", [(1000, 129, 0, 255, 256, 257), (4096, 3000)])
def test_target_verify_long_prefix(pl):
    """Few query rows over a long prefix: the shape the small-q (KV ping-pong) mode of the tcgen05 kernel
    serves -- odd and even tile counts, and a request with no prefix at all."""
    _verify_case(8, pl, 205, _tree)


def test_target_verify_causal_mask_equals_causal():
    """test_triton_attention_kernels.py:189-224: a custom mask equal to the causal mask gives the
    causal result."""
    def causal(rng, n, p):
        m = np.ones((n, p + n), dtype=np.uint8)
        m[:, p:] = np.tril(np.ones((n, n), dtype=np.uint8))
        return m
    pl = (50, 3)
    got, w_cpu = _verify_case(16, pl, 202, causal)
    ref, _, _, _ = oracle_extend(w_cpu, causal=True)
    assert_close_out(got, ref)


def test_target_verify_fully_masked_rows():
    """A row whose mask hides every draft token still sees its prefix; with an empty prefix it sees
    nothing: o = 0, lse = -inf."""
    def none(rng, n, p):
        m = np.ones((n, p + n), dtype=np.uint8)
        m[:, p:] = 0
        m[0, p] = 1
        return m
    _verify_case(4, (0, 20), 203, none)


def test_draft_extend_csr_indices():
    """DRAFT_EXTEND: kv_indices / kv_indptr / qo_indptr come from spec_info.generate_attn_arg_prefill
    (triton_backend.py:265-279): the kernel's CSR mode."""
    import b200attn
    from oracle import decode_oracle

    cfg = synth.Config("draft", 3, 0, 8, 2, 128, seed=301)
    pl, el = (40, 7, 130), (3, 1, 4)
    w_cpu, w = _pair(cfg, pl, el)

    def gen(req_pool_indices, paged_kernel_lens, paged_kernel_lens_sum, req_to_token):
        indptr, indices = decode_oracle.kv_indices(w_cpu.req_to_token, w_cpu.req_pool_indices, w_cpu.extend_prefix_lens)
        qo = np.zeros(len(pl) + 1, dtype=np.int32)
        qo[1:] = np.cumsum(el)
        return (torch.from_numpy(indices).to(DEV), torch.from_numpy(indptr).to(DEV), torch.from_numpy(qo).to(DEV), None)

    runner = make_extend_runner(w, DEV)
    backend = b200attn.B200AttnBackend(runner, return_lse=True)
    spec = types.SimpleNamespace(generate_attn_arg_prefill=gen)
    out = run_extend(backend, runner, w, _layer(w), mode=b200attn.ForwardMode.DRAFT_EXTEND, spec_info=spec)
    torch.cuda.synchronize()
    ref, _, _, _ = oracle_extend(w_cpu)
    assert_close_out(out.float().cpu().numpy().reshape(ref.shape), ref)


def test_extend_errors_are_loud():
    import b200attn

    cfg = synth.Config("err", 1, 0, 8, 2, 128, seed=401)
    w = synth.make_extend_workload(cfg, (4,), (8,), DEV)
    runner = make_extend_runner(w, DEV)
    backend = b200attn.B200AttnBackend(runner)
    with pytest.raises(RuntimeError):   # no metadata
        backend.forward_extend(w.q, w.k.view(-1, 2, 128), w.v.view(-1, 2, 128), _layer(w),
                               types.SimpleNamespace(token_to_kv_pool=runner.token_to_kv_pool,
                                                     out_cache_loc=w.out_cache_loc))
    q3 = w.q.view(-1, 8, 128)
    o3 = torch.empty_like(q3)
    qo = torch.tensor([0, 8], dtype=torch.int32, device=DEV)
    with pytest.raises(ValueError):     # neither kv_indices nor req_to_token
        torch.ops.b200attn.extend(q3, o3, None, w.k_pool, w.v_pool, qo, qo, None, None, None, None,
                                  w.out_cache_loc, None, None, 8, True, -1, True, 0.1, 0.0)
    with pytest.raises(RuntimeError):   # CPU tensors: CUDA-only op
        torch.ops.b200attn.extend(q3.cpu(), o3.cpu(), None, w.k_pool.cpu(), w.v_pool.cpu(), qo.cpu(), qo.cpu(),
                                  qo.cpu(), None, None, None, w.out_cache_loc.cpu(), None, None, 8, True, -1,
                                  True, 0.1, 0.0)


def test_extend_full_size_split_invariance():
    """Llama-3.1-8B-shaped prefill of 4096 tokens at full size, through a size-independent property:
    prefilling [0, 4096) in one pass equals prefilling [0, 2048) and then extending [2048, 4096) over
    the cached prefix (same pool rows), row for row on the second half; and the first token's output
    is its own V row."""
    import b200attn

    cfg = synth.Config("full", 1, 0, 32, 8, 128, seed=501)
    w = synth.make_extend_workload(cfg, (0,), (4096,), DEV)
    runner = make_extend_runner(w, DEV)
    backend = b200attn.B200AttnBackend(runner)
    layer = _layer(w)
    whole = run_extend(backend, runner, w, layer).view(4096, 32, 128).float()
    v0 = w.v.view(4096, 8, 128)[0].float().repeat_interleave(4, dim=0)
    assert torch.allclose(whole[0], v0, atol=1e-2)
    # second pass: same request, first 2048 tokens are now the cached prefix
    w2 = types.SimpleNamespace(**{**w.__dict__})
    w2.prefix_lens_np = np.array([2048])
    w2.extend_lens_np = np.array([2048])
    w2.extend_prefix_lens = torch.tensor([2048], device=DEV)
    w2.extend_seq_lens = torch.tensor([2048], device=DEV)
    w2.out_cache_loc = w.out_cache_loc[2048:].contiguous()
    w2.q, w2.k, w2.v = w.q[2048:], w.k[2048:], w.v[2048:]
    w2.tokens = 2048
    half = run_extend(backend, runner, w2, layer, save_kv_cache=False).view(2048, 32, 128).float()
    torch.cuda.synchronize()
    assert (half - whole[2048:]).abs().max().item() <= 2e-2


@pytest.mark.parametrize("mode", ["persistent", "oneshot"])
def test_small_q_kernels_forced(mode):
    """Small-q batches (verify / decode-like) have two kernels, chosen by pool type; B200ATTN_SMALLQ forces one
    (read once per process, hence the subprocess): the small-q cases of this file must pass on either."""
    import os, subprocess, sys
    env = dict(os.environ, B200ATTN_SMALLQ=mode)
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-q", "-x", "-m", "gpu", "-p", "no:cacheprovider",
                        "-k", "(verify or mask or fp8 or decode_like or draft or golden or shapes) and not forced"],
                       env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout


@pytest.mark.parametrize("window", [40, 300])
def test_swa_kv_pool_extend(window):
    """``SWAKVPool`` prefill: a sliding-window layer reads its cached prefix and its new tokens from the swa pool
    through ``full_to_swa_index_mapping`` (translate_loc_from_full_to_swa, memory_pool.py:515-517), a full layer
    of the same forward from the full pool; both against the extend oracle on the untranslated layout."""
    import b200attn

    cfg = synth.Config("swaext", 3, 0, 8, 4, 128, page_size=1, seed=88)
    pl, el = (0, 150, 700), (90, 33, 5)
    w_cpu, w = _pair(cfg, pl, el)
    runner = make_extend_runner(w, DEV, sliding_window_size=window)
    n = w.pool_size
    pool = b200attn.SWAKVPool(n, n, runner.token_to_kv_pool.dtype, cfg.num_kv_heads, cfg.head_dim, [1], [0], False, DEV)
    perm = torch.randperm(n + 1, generator=torch.Generator().manual_seed(9)).to(DEV)
    pool.full_to_swa_index_mapping = perm
    pool.full_kv_pool.k_buffer[0], pool.full_kv_pool.v_buffer[0] = w.k_pool.clone(), w.v_pool.clone()
    pool.swa_kv_pool.k_buffer[0][perm] = w.k_pool
    pool.swa_kv_pool.v_buffer[0][perm] = w.v_pool
    runner.token_to_kv_pool = pool
    backend = b200attn.B200AttnBackend(runner, return_lse=True)
    for layer_id, win in ((1, window), (0, -1)):
        layer = _layer(w, True, 0.0, win)
        layer.layer_id = layer_id
        out = run_extend(backend, runner, w, layer)
        torch.cuda.synchronize()
        ref, ref_lse, kp, vp = oracle_extend(w_cpu, causal=True, sliding_window=win)
        assert_close_out(out.float().cpu().numpy().reshape(ref.shape), ref)
        sub = pool.swa_kv_pool if layer_id == 1 else pool.full_kv_pool
        loc = w.out_cache_loc.long()
        at = perm[loc] if layer_id == 1 else loc
        assert torch.equal(sub.k_buffer[0][at].cpu(), kp[w_cpu.out_cache_loc.long()])
        assert torch.equal(sub.v_buffer[0][at].cpu(), vp[w_cpu.out_cache_loc.long()])
