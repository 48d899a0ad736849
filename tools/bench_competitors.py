"""Same-box library competitors of the decode path (informational): flashinfer's
BatchDecodeWithPagedKVCacheWrapper (the image has 0.6.11, not the reference's pinned 0.2.9rc2: a competitor,
not an oracle) and torch SDPA over a gathered, contiguous copy of the KV (what torch_native does, on the GPU)
on bench.py's configurations, page_size 1, next to this repo's decode step.  One JSON line per config.
usage: python tools/bench_competitors.py [C2 C4 ...]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ltp-sglang_b200")]
import torch
import b200attn, synth
from bench_harness import make_batch, make_layer, make_runner

dev = torch.device("cuda:0")


def timed(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


for name in (sys.argv[1:] or ["C2"]):
    base, _, tp = name.partition("/")   # "C2/8": the per-rank shape of an 8-way KV-head shard (batch x 8, heads / 8)
    c = synth.CONFIGS[base]
    if tp:
        from dataclasses import replace
        c = synth.shard(replace(c, batch=c.batch * int(tp)), int(tp))
    w = synth.make_workload(c, dev)
    runner = make_runner(w, dev)
    backend = b200attn.B200AttnBackend(runner, materialize_kv_indices=True)
    layer = make_layer(w)
    fb = make_batch(w, runner, backend)
    backend.init_forward_metadata(fb)
    B, Hq, Hkv, D = c.batch, c.num_q_heads, c.num_kv_heads, c.head_dim
    # outputs are compared on the SAME pool state: no KV store here (a store divides k / v in place on an fp8 pool,
    # so every later call writes another new-token row: round 1's 0.10 "gap" on C3 was exactly that, see
    # tools/fp8_vs_flashinfer.py and profiles/r02_fp8_vs_flashinfer.jsonl)
    ours = layer(w.q, w.k, w.v, fb, save_kv_cache=False).view(B, Hq, D).clone()
    pool_k0, pool_v0 = w.k_pool[w.out_cache_loc.long()].clone(), w.v_pool[w.out_cache_loc.long()].clone()
    e = 1 if c.kv_dtype == "fp8_e4m3" else 2
    kv_bytes = w.seq_lens_sum * Hkv * 2 * D * e
    rec = {"config": name, "kv_bytes": kv_bytes, "page_size": c.page_size}
    def graphed(fn):  # both sides are timed as CUDA-graph replays: no Python or launch overhead on either
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            fn()
        torch.cuda.current_stream().wait_stream(s)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        return g

    g_ours = graphed(lambda: layer(w.q, w.k, w.v, fb))
    t = timed(g_ours.replay)
    rec["b200attn"] = {"ms": t, "gb_per_s": kv_bytes / t / 1e6, "note": "kv_store + decode + merge, graph replay"}
    md = backend.forward_metadata
    kv_indptr, kv_indices = md.kv_indptr[: B + 1], md.kv_indices
    fp8 = c.kv_dtype == "fp8_e4m3"
    w.k_pool[w.out_cache_loc.long()] = pool_k0   # the timed replays stored new-token rows: back to the compared state
    w.v_pool[w.out_cache_loc.long()] = pool_v0
    if c.page_size == 1:
        try:
            import flashinfer
            t0 = time.time()
            ws = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
            fw = flashinfer.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD", use_tensor_cores=True)
            fw.plan(kv_indptr, kv_indices, torch.ones(B, dtype=torch.int32, device=dev), Hq, Hkv, D, 1,
                    q_data_type=w.q.dtype, kv_data_type=torch.float8_e4m3fn if fp8 else w.k_pool.dtype,
                    sm_scale=w.sm_scale)
            q3 = w.q.view(B, Hq, D).contiguous()
            kc, vc = w.k_pool.unsqueeze(1), w.v_pool.unsqueeze(1)
            kw = {}
            if fp8:
                kc, vc = kc.view(torch.float8_e4m3fn), vc.view(torch.float8_e4m3fn)
                kw = {"k_scale": float(c.k_scale), "v_scale": float(c.v_scale)}
            o = fw.run(q3, (kc, vc), **kw)
            torch.cuda.synchronize()
            first = time.time() - t0
            o_buf = torch.empty_like(q3)
            g_fi = graphed(lambda: fw.run(q3, (kc, vc), out=o_buf, **kw))
            t = timed(g_fi.replay)
            rec["flashinfer"] = {"version": flashinfer.__version__, "ms": t, "gb_per_s": kv_bytes / t / 1e6,
                                 "max_abs_diff_vs_b200attn": float((o.float() - ours.float()).abs().max()),
                                 "first_call_s": first, "note": "run() only (no KV store), plan() outside, graph replay; tensor-core decode, page_size 1"}
        except Exception as ex:  # noqa: BLE001 -- a competitor that does not run here is recorded, not fatal
            rec["flashinfer"] = {"unavailable": repr(ex)[:300]}
    print(json.dumps(rec), flush=True)
