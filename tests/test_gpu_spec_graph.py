"""GPU tests of the speculative-decoding metadata paths, eager and under CUDA graphs:

* decode with ``spec_info.kv_indptr / kv_indices`` (the CSR plan, triton_backend.py:206-208);
* capture-then-replay of DECODE (CSR), TARGET_VERIFY and DRAFT_EXTEND graphs where every replay hands the
  hooks a *fresh* ``spec_info`` whose tensors live at other addresses and hold other values, as
  ``CudaGraphRunner.replay_prepare`` does (cuda_graph_runner.py:760-768) -- the replayed result must equal the
  eager result of the same inputs bit for bit (triton_backend.py:445-501, 575-623 keeps static copies);
* ``draft_kv_indices`` against the restated ``generate_draft_decode_kv_indices`` (eagle_utils.py:844-920);
* ``B200MultiStepDraftBackend`` (``TritonMultiStepDraftBackend``, triton_backend.py:735-872).
"""
from __future__ import annotations

import types

import numpy as np
import pytest
import torch

import synth
from tests.harness import assert_close_out, make_extend_runner, make_layer, make_runner

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _csr_from(w_cpu, order):
    """CSR lists of the requests in ``order`` (a permutation / selection of batch rows)."""
    from oracle import decode_oracle as oracle

    rpi = w_cpu.req_pool_indices[order]
    sl = w_cpu.seq_lens[order]
    indptr, indices = oracle.kv_indices(w_cpu.req_to_token, rpi, sl)
    return indptr, indices


def _oracle_csr(q, k_pool, v_pool, indptr, indices, sm_scale):
    """Oracle attention over explicit CSR lists: a pseudo req_to_token with one row per list."""
    from oracle import decode_oracle as oracle

    n = len(indptr) - 1
    lens = np.diff(indptr).astype(np.int64)
    r2t = np.zeros((n, max(int(lens.max()), 1)), dtype=np.int32)
    for i in range(n):
        r2t[i, : lens[i]] = indices[indptr[i]: indptr[i + 1]]
    return oracle.decode_attention(q, k_pool, v_pool, torch.from_numpy(r2t), torch.arange(n),
                                   torch.from_numpy(lens), sm_scale)


def _decode_fb(b200attn, w, runner, backend, rows, spec, dtype=torch.int64):
    B = w.cfg.batch
    return b200attn.ForwardBatch(
        forward_mode=b200attn.ForwardMode.DECODE, batch_size=B,
        req_pool_indices=w.req_pool_indices.to(dtype), seq_lens=w.seq_lens.to(dtype),
        out_cache_loc=torch.zeros(rows, dtype=torch.int64, device=DEV), seq_lens_sum=w.seq_lens_sum,
        req_to_token_pool=runner.req_to_token_pool, token_to_kv_pool=runner.token_to_kv_pool,
        attn_backend=backend, spec_info=spec)


# ----------------------------------------------------------------------------- #
# decode over spec_info.kv_indptr / kv_indices
# ----------------------------------------------------------------------------- #
@pytest.mark.parametrize("kv_dtype", ["bf16", "fp8_e4m3"])
def test_decode_csr_spec_info(kv_dtype):
    """forward_decode fed with spec_info lists (plan_csr -> b200attn_plan with csr_indptr) == oracle over the
    same lists; the lists name the requests in another order than req_pool_indices and repeat one (topk)."""
    import b200attn

    cfg = synth.Config("csr", 6, 0, 32, 8, 128, kv_dtype=kv_dtype, page_size=4, ragged=(1, 1300),
                       k_scale=0.5 if kv_dtype != "bf16" else 1.0, v_scale=2.0 if kv_dtype != "bf16" else 1.0, seed=61)
    w_cpu = synth.make_workload(cfg, "cpu")
    w = synth.make_workload(cfg, DEV)
    w.k_pool.copy_(w_cpu.k_pool)
    w.v_pool.copy_(w_cpu.v_pool)
    order = np.array([3, 0, 0, 5, 1, 4, 2, 2])          # 8 rows over 6 requests
    indptr, indices = _csr_from(w_cpu, order)
    rows = len(order)
    g = torch.Generator().manual_seed(7)
    q_cpu = torch.randn((rows, cfg.num_q_heads, cfg.head_dim), generator=g).to(torch.bfloat16)
    runner = make_runner(w, DEV)
    runner.server_args = types.SimpleNamespace(speculative_eagle_topk=2, speculative_num_draft_tokens=None,
                                               speculative_num_steps=None, speculative_algorithm="EAGLE")
    backend = b200attn.B200AttnBackend(runner, return_lse=True, max_ctas=13, min_chunks=1)
    spec = types.SimpleNamespace(kv_indptr=torch.from_numpy(indptr).to(DEV), kv_indices=torch.from_numpy(indices).to(DEV))
    fb = _decode_fb(b200attn, w, runner, backend, rows, spec)
    backend.init_forward_metadata(fb)
    layer = make_layer(w)
    out = layer(q_cpu.to(DEV).view(rows, -1), None, None, fb, save_kv_cache=False)
    torch.cuda.synchronize()
    fp8 = kv_dtype != "bf16"
    from oracle import decode_oracle as oracle

    n = rows
    lens = np.diff(indptr).astype(np.int64)
    r2t = np.zeros((n, int(lens.max())), dtype=np.int32)
    for i in range(n):
        r2t[i, : lens[i]] = indices[indptr[i]: indptr[i + 1]]
    ref, ref_lse = oracle.decode_attention(q_cpu, w_cpu.k_pool, w_cpu.v_pool, torch.from_numpy(r2t), torch.arange(n),
                                           torch.from_numpy(lens), w_cpu.sm_scale, 0.0,
                                           cfg.k_scale if fp8 else 1.0, cfg.v_scale if fp8 else 1.0)
    assert_close_out(out.float().cpu().numpy().reshape(ref.shape), ref)
    assert np.abs(backend.last_lse(rows).cpu().numpy() - ref_lse).max() <= 1e-3


def _capture(fn):
    """Warm up on a side stream, then capture ``fn`` into a CUDA graph; returns (graph, static output)."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(2):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out = fn()
    return graph, out


def test_graph_replay_decode_csr_fresh_spec_info():
    """Capture DECODE with one spec_info, replay with others (new tensors, new addresses, new lists):
    each replay equals the eager forward of the same lists bit for bit."""
    import b200attn

    cfg = synth.Config("csrg", 5, 0, 32, 8, 128, page_size=1, ragged=(30, 900), seed=62)
    w = synth.make_workload(cfg, DEV)
    w_cpu_tables = types.SimpleNamespace(req_to_token=w.req_to_token.cpu(), req_pool_indices=w.req_pool_indices.cpu(),
                                         seq_lens=w.seq_lens.cpu())
    runner = make_runner(w, DEV)
    runner.server_args = types.SimpleNamespace(speculative_eagle_topk=2, speculative_num_draft_tokens=None,
                                               speculative_num_steps=None, speculative_algorithm="EAGLE")
    rows = 8
    orders = [np.array([0, 1, 2, 3, 4, 0, 1, 2]), np.array([4, 4, 3, 3, 2, 1, 0, 0]), np.array([2, 0, 4, 1, 3, 3, 1, 4])]
    q = torch.randn((rows, cfg.num_q_heads * cfg.head_dim), device=DEV).to(torch.bfloat16)
    layer = make_layer(w)

    def spec_for(order, pad_front):
        indptr, indices = _csr_from(w_cpu_tables, order)
        # fresh allocations at shifted addresses: slices of larger buffers
        ip = torch.zeros(len(indptr) + pad_front, dtype=torch.int32, device=DEV)
        ix = torch.zeros(len(indices) + pad_front, dtype=torch.int32, device=DEV)
        ip[pad_front:] = torch.from_numpy(indptr).to(DEV)
        ix[pad_front:] = torch.from_numpy(indices).to(DEV)
        return types.SimpleNamespace(kv_indptr=ip[pad_front:], kv_indices=ix[pad_front:])

    # eager results, one backend per call so nothing is shared with the graph path
    eager = []
    for i, order in enumerate(orders):
        be = b200attn.B200AttnBackend(runner)
        fb = _decode_fb(b200attn, w, runner, be, rows, spec_for(order, 0))
        be.init_forward_metadata(fb)
        eager.append(layer(q, None, None, fb, save_kv_cache=False).clone())
    torch.cuda.synchronize()

    backend = b200attn.B200AttnBackend(runner)
    backend.init_cuda_graph_state(cfg.batch, rows)
    rpi32, sl32 = w.req_pool_indices.to(torch.int32), w.seq_lens.to(torch.int32)
    spec0 = spec_for(orders[0], 4)
    fb = _decode_fb(b200attn, w, runner, backend, rows, spec0, torch.int32)
    backend.init_forward_metadata_capture_cuda_graph(cfg.batch, rows, rpi32, sl32, None, fb.forward_mode, spec0)
    graph, out = _capture(lambda: layer(q, None, None, fb, save_kv_cache=False))
    for i in (1, 2, 0, 1):
        spec = spec_for(orders[i], 4 * (i + 2))
        spec0.kv_indices.fill_(0)        # the capture-time tensors are dead: a stale pointer would read zeros
        spec0.kv_indptr.fill_(0)
        backend.init_forward_metadata_replay_cuda_graph(cfg.batch, rpi32, sl32, -1, None, fb.forward_mode, spec, None)
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, eager[i]), f"replay with spec_info {i} differs from the eager forward"


def _tree_mask(rng, n, p):
    m = np.ones((n, p + n), dtype=np.uint8)
    t = np.tril(rng.integers(0, 2, size=(n, n), dtype=np.uint8))
    np.fill_diagonal(t, 1)
    m[:, p:] = t
    return m


@pytest.mark.parametrize("kv_dtype", ["bf16", "fp8_e4m3"])
def test_graph_replay_target_verify_fresh_mask(kv_dtype):
    """TARGET_VERIFY graph: the custom mask of every replay comes in a new tensor; results equal eager."""
    import b200attn
    from oracle import extend_oracle

    n_draft = 8
    pl = (100, 37, 1, 260, 700)
    el = (n_draft,) * len(pl)
    cfg = synth.Config("vg", len(pl), 0, 32, 8, 128, kv_dtype=kv_dtype, k_scale=0.5 if kv_dtype != "bf16" else 1.0,
                       v_scale=2.0 if kv_dtype != "bf16" else 1.0, seed=63)
    w = synth.make_extend_workload(cfg, pl, el, DEV)
    _, _, mi = extend_oracle.extend_indptrs(pl, el)
    runner = make_extend_runner(w, DEV, num_draft_tokens=n_draft)
    layer = b200attn.AttnLayer(cfg.num_q_heads, cfg.head_dim, w.sm_scale, cfg.num_kv_heads, layer_id=0)
    if kv_dtype != "bf16":
        layer.k_scale_float, layer.v_scale_float = cfg.k_scale, cfg.v_scale

    def mask_for(seed):
        rng = np.random.default_rng(seed)
        mask = np.zeros(int(mi[-1]), dtype=np.uint8)
        for b in range(len(pl)):
            mask[mi[b]:mi[b + 1]] = _tree_mask(rng, n_draft, pl[b]).reshape(-1)
        return torch.from_numpy(mask).to(DEV).view(torch.bool)

    bs = len(pl)
    mode = b200attn.ForwardMode.TARGET_VERIFY
    prefix = w.extend_prefix_lens            # TARGET_VERIFY: seq_lens = the committed lengths
    rpi32, sl32 = w.req_pool_indices.to(torch.int32), prefix.to(torch.int32)

    def fb_for(backend, spec):
        return b200attn.ForwardBatch(
            forward_mode=mode, batch_size=bs, req_pool_indices=rpi32, seq_lens=sl32, out_cache_loc=w.out_cache_loc,
            seq_lens_sum=0,   # no host-side length hint: a captured graph has none either (same small-q kernel)
            req_to_token_pool=runner.req_to_token_pool,
            token_to_kv_pool=runner.token_to_kv_pool, attn_backend=backend, spec_info=spec)

    # K/V of the draft tokens are stored once; every forward below reads them back (save_kv_cache=False)
    b200attn.store_kv(w.k_pool, w.v_pool, w.out_cache_loc, w.k.view(-1, cfg.num_kv_heads, cfg.head_dim).clone(),
                      w.v.view(-1, cfg.num_kv_heads, cfg.head_dim).clone(),
                      cfg.k_scale if kv_dtype != "bf16" else None, cfg.v_scale if kv_dtype != "bf16" else None)
    seeds = [11, 12, 13]
    eager = []
    for sd in seeds:
        be = b200attn.B200AttnBackend(runner)
        spec = types.SimpleNamespace(custom_mask=mask_for(sd), draft_token_num=n_draft)
        fb = fb_for(be, spec)
        be.init_forward_metadata(fb)
        eager.append(layer(w.q, w.k, w.v, fb, save_kv_cache=False).clone())
    torch.cuda.synchronize()
    assert not torch.equal(eager[0], eager[1])   # the masks matter

    backend = b200attn.B200AttnBackend(runner)
    backend.init_cuda_graph_state(bs, bs * n_draft)
    spec0 = types.SimpleNamespace(custom_mask=mask_for(seeds[0]), draft_token_num=n_draft)
    fb = fb_for(backend, spec0)
    backend.init_forward_metadata_capture_cuda_graph(bs, bs * n_draft, rpi32, sl32, None, mode, spec0)
    graph, out = _capture(lambda: layer(w.q, w.k, w.v, fb, save_kv_cache=False))
    for i in (1, 2, 0):
        spec0.custom_mask.fill_(False)       # dead capture-time tensor
        pad = torch.zeros(64 * (i + 1), dtype=torch.bool, device=DEV)   # shifts the next allocation
        spec = types.SimpleNamespace(custom_mask=mask_for(seeds[i]), draft_token_num=n_draft)
        backend.init_forward_metadata_replay_cuda_graph(bs, rpi32, sl32, int(prefix.sum()), None, mode, spec, None)
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, eager[i]), f"replay with mask {i} differs from the eager forward"
        del pad


def test_graph_replay_draft_extend_accept_lengths():
    """DRAFT_EXTEND graph: captured with speculative_num_steps + 1 tokens per request, replayed with
    spec_info.accept_length (triton_backend.py:476-501, 607-623); nothing allocates on replay.  Equals the eager
    forward whose lists come from generate_attn_arg_prefill, bit for bit, and the oracle."""
    import b200attn
    from oracle import decode_oracle, extend_oracle

    steps = 3
    per = steps + 1
    cfg = synth.Config("dg", 4, 0, 32, 8, 128, seed=64)
    pl = (90, 7, 300, 41)                      # seq_lens handed to the hooks: the prefix list of each request
    el_cap = (per,) * len(pl)
    w = synth.make_extend_workload(cfg, pl, el_cap, DEV)
    bs = len(pl)
    runner = make_extend_runner(w, DEV)
    runner.server_args = types.SimpleNamespace(speculative_num_draft_tokens=None, speculative_num_steps=steps,
                                               speculative_eagle_topk=1, speculative_algorithm="EAGLE")
    layer = b200attn.AttnLayer(cfg.num_q_heads, cfg.head_dim, w.sm_scale, cfg.num_kv_heads, layer_id=0)
    mode = b200attn.ForwardMode.DRAFT_EXTEND
    prefix = w.extend_prefix_lens
    rpi32, sl32 = w.req_pool_indices.to(torch.int32), prefix.to(torch.int32)
    r2t_cpu, rpi_cpu = w.req_to_token.cpu(), w.req_pool_indices.cpu()
    H, D = cfg.num_kv_heads, cfg.head_dim

    def inputs_for(accept):
        """q / k / v rows and out_cache_loc of a step that accepted `accept[b]` tokens of request b: request b's
        rows are the first accept[b] of its `per` captured rows, packed back to back (qo_indptr = cumsum)."""
        rows = np.concatenate([np.arange(b * per, b * per + a) for b, a in enumerate(accept)])
        n = len(rows)
        qkv = torch.zeros_like(w.qkv)
        qkv[:n] = w.qkv[torch.from_numpy(rows).to(DEV)]
        loc = torch.zeros_like(w.out_cache_loc)
        loc[:n] = w.out_cache_loc[torch.from_numpy(rows).to(DEV)]
        return qkv, loc, n

    def fb_for(backend, spec, loc):
        return b200attn.ForwardBatch(
            forward_mode=mode, batch_size=bs, req_pool_indices=rpi32, seq_lens=sl32, out_cache_loc=loc,
            seq_lens_sum=0, req_to_token_pool=runner.req_to_token_pool, token_to_kv_pool=runner.token_to_kv_pool,
            attn_backend=backend, spec_info=spec)

    # all captured-size K/V rows are in the pool once
    b200attn.store_kv(w.k_pool, w.v_pool, w.out_cache_loc, w.k.view(-1, H, D).clone(), w.v.view(-1, H, D).clone())

    def eager_spec(accept):
        def gen(req_pool_indices, paged_kernel_lens, paged_kernel_lens_sum, req_to_token):
            indptr, indices = decode_oracle.kv_indices(r2t_cpu, rpi_cpu, prefix.cpu())
            qo = np.zeros(bs + 1, dtype=np.int32)
            qo[1:] = np.cumsum(accept)
            return (torch.from_numpy(indices).to(DEV), torch.from_numpy(indptr).to(DEV), torch.from_numpy(qo).to(DEV), None)
        return types.SimpleNamespace(generate_attn_arg_prefill=gen,
                                     accept_length=torch.tensor(accept, dtype=torch.int32, device=DEV))

    accepts = [(4, 4, 4, 4), (1, 3, 2, 4), (2, 1, 1, 1)]
    eager = []
    for acc in accepts:
        qkv, loc, n = inputs_for(acc)
        q, k, v = qkv.split([cfg.num_q_heads * D, H * D, H * D], dim=-1)
        be = b200attn.B200AttnBackend(runner)
        fb = fb_for(be, eager_spec(acc), loc)
        be.init_forward_metadata(fb)
        eager.append((layer(q[:n], k[:n], v[:n], fb, save_kv_cache=False).clone(), n))
    torch.cuda.synchronize()

    backend = b200attn.B200AttnBackend(runner)
    backend.init_cuda_graph_state(bs, bs * per)
    s_qkv = torch.zeros_like(w.qkv)
    s_loc = torch.zeros_like(w.out_cache_loc)
    sq, sk, sv = s_qkv.split([cfg.num_q_heads * D, H * D, H * D], dim=-1)
    spec0 = types.SimpleNamespace(accept_length=torch.full((bs,), per, dtype=torch.int32, device=DEV))
    fb = fb_for(backend, spec0, s_loc)
    qkv, loc, _ = inputs_for(accepts[0])
    s_qkv.copy_(qkv)
    s_loc.copy_(loc)
    backend.init_forward_metadata_capture_cuda_graph(bs, bs * per, rpi32, sl32, None, mode, spec0)
    graph, out = _capture(lambda: layer(sq, sk, sv, fb, save_kv_cache=False))
    for i in (1, 2, 0, 1):
        qkv, loc, n = inputs_for(accepts[i])
        s_qkv.copy_(qkv)
        s_loc.copy_(loc)
        spec = types.SimpleNamespace(accept_length=torch.tensor(accepts[i], dtype=torch.int32, device=DEV))
        before = torch.cuda.memory_allocated()
        backend.init_forward_metadata_replay_cuda_graph(bs, rpi32, sl32, int(prefix.sum()), None, mode, spec, None)
        assert torch.cuda.memory_allocated() == before, "the replay hook allocated device memory"
        graph.replay()
        torch.cuda.synchronize()
        ref_out, n_ref = eager[i]
        assert n == n_ref and torch.equal(out[:n], ref_out), f"replay with accept lengths {accepts[i]} differs from eager"
    # and the last replay against the oracle: prefix list + causal extend tokens
    acc = accepts[1]
    qkv, loc, n = inputs_for(acc)
    ps, _ = extend_oracle.slots_from_req_to_token(r2t_cpu, rpi_cpu, np.asarray(pl), np.asarray(acc), -1)
    loc_np = loc[:n].cpu().numpy()
    es, o = [], 0
    for a in acc:
        es.append(loc_np[o:o + a])
        o += a
    q = qkv[:n, : cfg.num_q_heads * D].cpu().reshape(n, cfg.num_q_heads, D)
    ref, _ = extend_oracle.extend_attention(q, w.k_pool.cpu(), w.v_pool.cpu(), ps, es, w.sm_scale, causal=True)
    assert_close_out(out[:n].float().cpu().numpy().reshape(ref.shape), ref)


# ----------------------------------------------------------------------------- #
# multi-step draft decode
# ----------------------------------------------------------------------------- #
def _draft_workload(steps, topk, page, seed, batch=5):
    """Requests whose req_to_token rows hold, after the committed S tokens, the slots the draft branches
    will fill (topk * steps at page 1; whole pages per branch otherwise)."""
    room = topk * (steps + 2 * page) + page
    cfg = synth.Config("draft", batch, 0, 32, 8, 128, page_size=page, ragged=(20 + room, 700 + room), seed=seed)
    w_cpu = synth.make_workload(cfg, "cpu")
    w = synth.make_workload(cfg, DEV)
    w.k_pool.copy_(w_cpu.k_pool)
    w.v_pool.copy_(w_cpu.v_pool)
    committed = w_cpu.seq_lens - room
    return cfg, w_cpu, w, committed


@pytest.mark.parametrize("steps,topk,page,idx", [(3, 1, 1, torch.int64), (4, 2, 1, torch.int32), (2, 4, 16, torch.int64),
                                                 (5, 3, 4, torch.int32), (1, 8, 1, torch.int64)])
def test_draft_kv_indices(steps, topk, page, idx):
    """One launch == the restated generate_draft_decode_kv_indices (eagle_utils.py:844-920), every entry."""
    import b200attn  # noqa: F401
    from oracle import decode_oracle as oracle

    cfg, w_cpu, w, committed = _draft_workload(steps, topk, page, 70 + steps)
    n = cfg.batch
    positions = committed.repeat_interleave(topk)
    width = int(committed.sum()) * topk + n * topk * steps + 64
    ref_ptr, ref_idx = oracle.draft_decode_kv_indices(w_cpu.req_to_token, w_cpu.req_pool_indices, committed, positions,
                                                      steps, topk, page, width)
    kv_indices = torch.zeros((steps, width), dtype=torch.int32, device=DEV)
    kv_indptr = torch.zeros((steps, n * topk + 1), dtype=torch.int32, device=DEV)
    torch.ops.b200attn.draft_kv_indices(w.req_pool_indices.to(idx), w.req_to_token, committed.to(DEV).to(idx),
                                        positions.to(DEV), kv_indices, kv_indptr, topk, page)
    torch.cuda.synchronize()
    assert np.array_equal(kv_indptr.cpu().numpy(), ref_ptr)
    assert np.array_equal(kv_indices.cpu().numpy(), ref_idx)
    # the lists tile the buffer back to back: row z of step it spans [indptr[z], indptr[z + 1])
    for it in range(steps):
        lens = np.diff(ref_ptr[it])
        assert np.array_equal(lens, (committed.numpy() + it + 1).repeat(topk))


@pytest.mark.parametrize("graph", [False, True])
def test_multi_step_draft_backend(graph):
    """B200MultiStepDraftBackend: steps x decode over the shared kv_indptr / kv_indices rows, eager and
    captured (one graph per step, replayed after the batch changed) -- every step against the oracle."""
    import b200attn
    from oracle import decode_oracle as oracle

    steps, topk, page = 3, 2, 1
    cfg, w_cpu, w, committed = _draft_workload(steps, topk, page, 81)
    n = cfg.batch
    rows = n * topk
    runner = make_runner(w, DEV)
    runner.server_args = types.SimpleNamespace(speculative_eagle_topk=topk, speculative_num_draft_tokens=None,
                                               speculative_num_steps=steps, speculative_algorithm="EAGLE", page_size=page)
    draft = b200attn.B200MultiStepDraftBackend(runner, topk, steps, return_lse=True)
    assert len(draft.attn_backends) == steps and draft.kv_indptr.shape == (steps, runner.req_to_token_pool.size * topk + 1)
    layer = make_layer(w)
    g = torch.Generator().manual_seed(9)
    q_cpu = torch.randn((rows, cfg.num_q_heads, cfg.head_dim), generator=g).to(torch.bfloat16)
    q = q_cpu.to(DEV).view(rows, -1)

    def batch_for(sl_cpu, dtype):
        spec = types.SimpleNamespace(kv_indptr=None, kv_indices=None)
        return b200attn.ForwardBatch(
            forward_mode=b200attn.ForwardMode.DECODE, batch_size=n, req_pool_indices=w.req_pool_indices.to(dtype),
            seq_lens=sl_cpu.to(DEV).to(dtype), out_cache_loc=torch.zeros(rows, dtype=torch.int64, device=DEV),
            seq_lens_sum=int(sl_cpu.sum()), req_to_token_pool=runner.req_to_token_pool,
            token_to_kv_pool=runner.token_to_kv_pool, attn_backend=None, spec_info=spec,
            positions=sl_cpu.repeat_interleave(topk).to(DEV))

    def check(outs, sl_cpu):
        width = int(sl_cpu.sum()) * topk + rows * steps
        ref_ptr, ref_idx = oracle.draft_decode_kv_indices(w_cpu.req_to_token, w_cpu.req_pool_indices, sl_cpu,
                                                          sl_cpu.repeat_interleave(topk), steps, topk, page, width)
        for i in range(steps):
            ref, ref_lse = _oracle_csr(q_cpu, w_cpu.k_pool, w_cpu.v_pool, ref_ptr[i], ref_idx[i], w_cpu.sm_scale)
            assert_close_out(outs[i].float().cpu().numpy().reshape(ref.shape), ref, what=f"draft step {i}")

    if not graph:
        fb = batch_for(committed, torch.int64)
        draft.init_forward_metadata(fb)
        outs = []
        for i in range(steps):
            fb.attn_backend = draft.attn_backends[i]
            outs.append(layer(q, None, None, fb, save_kv_cache=False).clone())
        torch.cuda.synchronize()
        check(outs, committed)
        return

    draft.init_cuda_graph_state(n, rows)
    fb = batch_for(committed, torch.int32)
    draft.init_forward_metadata_capture_cuda_graph(fb)
    graphs = []
    for i in range(steps):
        fb_i = batch_for(committed, torch.int32)
        fb_i.attn_backend = draft.attn_backends[i]
        graphs.append(_capture(lambda fb_i=fb_i: layer(q, None, None, fb_i, save_kv_cache=False)))
    for shrink in (0, 7, 3):
        sl = committed - shrink                      # the batch moved on: other lengths, same requests
        fb = batch_for(sl, torch.int32)
        draft.init_forward_metadata_replay_cuda_graph(fb, n)
        outs = []
        for gr, out in graphs:
            gr.replay()
            outs.append(out.clone())
        torch.cuda.synchronize()
        check(outs, sl)
