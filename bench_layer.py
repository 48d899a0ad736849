"""bench_layer.py -- one Llama attention block in decode, tensor-parallel over N B200s.

    python bench_layer.py --gpus 1
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 ... bench_layer.py --gpus N

The layer-level bench SURVEY.md section 8(d)/(e) names for C4 (Llama-3.1-70B-shaped: hidden 8192,
64 q / 8 kv heads, head_dim 128, batch 256, ctx 8192, bf16).  Per rank (sglang TP: q/kv heads and the
o_proj input dim sharded, reference python/sglang/srt/models/llama.py:100-190):

    qkv  = x @ Wqkv_r^T                      cuBLAS (plain library GEMM)          llama.py:186
    rope + KV store                          b200attn rope_kv_store (K4, 1 launch) llama.py:187-188 + memory_pool.py:369-407
    attn = paged decode attention            b200attn decode (K2 + K3)            radix_attention.py:102-110
    out  = attn @ Wo_r^T                     cuBLAS                               llama.py:189
    all_reduce(out)                          NCCL over NVLink (N > 1)             layers/linear.py:1302-1303

This is the ONE place a collective appears: attention itself shards by KV head with no exchange
(the all-reduce sums o_proj partials).  The step is captured in one CUDA graph; time = CUDA events,
max over ranks.  Strong scaling: the batch is fixed, every rank holds 1/N of the heads and KV bytes.
Prints one JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from dataclasses import replace

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "ltp-sglang_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import bench  # noqa: E402  (Rank, ClockSampler, L2_BYTES)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--config", default="C4")
    ap.add_argument("--hidden", type=int, default=0, help="model hidden size (default: q heads x head_dim)")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--ctx", type=int, default=0)
    ap.add_argument("--unfused", action="store_true", help="RoPE-only launch + separate kv_store (reference call shape)")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    import b200attn
    import synth
    from b200attn import capi
    from bench_harness import make_batch, make_layer, make_runner

    R = bench.Rank(args)
    world, dev = R.world, R.dev
    gcfg = synth.CONFIGS[args.config]
    if args.batch:
        gcfg = replace(gcfg, batch=args.batch)
    if args.ctx:
        gcfg = replace(gcfg, ctx=args.ctx)
    hidden = args.hidden or gcfg.num_q_heads * gcfg.head_dim
    c = synth.shard(gcfg, world)
    w = synth.make_workload(c, dev)
    runner = make_runner(w, dev, tp_size=world)
    backend = b200attn.B200AttnBackend(runner)
    layer = make_layer(w)
    fb = make_batch(w, runner, backend)
    B, Hq, Hkv, D = c.batch, c.num_q_heads, c.num_kv_heads, c.head_dim
    dt = torch.bfloat16
    g = torch.Generator(device=dev)
    g.manual_seed(4321 + R.rank)
    x = torch.randn((B, hidden), generator=g, device=dev, dtype=torch.float32).to(dt)
    w_qkv = (torch.randn(((Hq + 2 * Hkv) * D, hidden), generator=g, device=dev, dtype=torch.float32) * hidden ** -0.5).to(dt)
    w_o = (torch.randn((hidden, Hq * D), generator=g, device=dev, dtype=torch.float32) * (Hq * D * world) ** -0.5).to(dt)
    qkv = torch.empty((B, (Hq + 2 * Hkv) * D), device=dev, dtype=dt)
    out = torch.empty((B, hidden), device=dev, dtype=dt)
    cache = synth.cos_sin_cache(D, max(gcfg.ctx, 1) + 8, 500000.0).to(dev)
    positions = (w.seq_lens - 1).to(torch.int64)
    pool = runner.token_to_kv_pool
    k_pool, v_pool = pool.k_buffer[0], pool.v_buffer[0]
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    q3, k3, v3 = q.view(B, Hq, D), k.view(B, Hkv, D), v.view(B, Hkv, D)
    backend.init_forward_metadata(fb)

    def step(allreduce=True):
        torch.matmul(x, w_qkv.t(), out=qkv)
        if args.unfused:
            b200attn.rope_store_kv(q3, k3, None, positions, cache, True)
            attn = layer(q, k, v, fb, save_kv_cache=True)
        else:
            b200attn.rope_store_kv(q3, k3, v3, positions, cache, True, w.out_cache_loc, k_pool, v_pool)
            attn = layer(q, k, v, fb, save_kv_cache=False)
        torch.matmul(attn, w_o.t(), out=out)
        if world > 1 and allreduce:
            dist.all_reduce(out)
        return out

    def graphed(fn):
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(2):
                fn()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            fn()
        return gr

    kv_bytes = w.seq_lens_sum * Hkv * 2 * D * 2
    flush = kv_bytes < 4 * bench.L2_BYTES

    def timed(gr, steps):
        for _ in range(max(args.warmup, 3)):
            gr.replay()
        R.barrier()
        ms = R.timed(gr.replay, steps, flush)
        R.barrier()
        return R.max_over_ranks(ms)

    n0 = capi.launch_count()
    step()
    launches = capi.launch_count() - n0
    g_full = graphed(step)
    sampler = bench.ClockSampler(R.local_rank)
    sampler.start()
    ms_full = timed(g_full, args.steps)
    clocks = sampler.stop()
    # breakdown (each part alone, same graphs machinery)
    g_attn = graphed(lambda: layer(q, k, v, fb, save_kv_cache=False))
    ms_attn = timed(g_attn, args.steps)
    g_rope = graphed(lambda: b200attn.rope_store_kv(q3, k3, v3, positions, cache, True, w.out_cache_loc, k_pool, v_pool))
    ms_rope = timed(g_rope, args.steps)
    g_gemm = graphed(lambda: (torch.matmul(x, w_qkv.t(), out=qkv), torch.matmul(q, w_o.t(), out=out)))
    ms_gemm = timed(g_gemm, args.steps)
    ms_ar = None
    if world > 1:
        g_ar = graphed(lambda: dist.all_reduce(out))
        ms_ar = timed(g_ar, args.steps)
    res = step()
    torch.cuda.synchronize()
    assert torch.isfinite(res.float()).all()

    if R.rank == 0:
        attn_bytes = kv_bytes + 2 * B * Hq * D * 2 + 4 * w.seq_lens_sum
        line = {
            "metric": "decode attention-block tokens/s (qkv_proj -> RoPE + KV store -> paged decode attention -> o_proj -> all-reduce)",
            "value": B / (ms_full * 1e-3), "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_full, "higher_is_better": True, "scaling": "strong",
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"{gcfg.name} layer: hidden {hidden}, {gcfg.num_q_heads}q/{gcfg.num_kv_heads}kv heads, "
                                   f"head_dim {D}, batch {B}, ctx {gcfg.ctx}, bf16 KV, page_size {gcfg.page_size}",
                       "global_batch": B, "seq_len": gcfg.ctx,
                       "parallelism": f"tp{world} (heads sharded; NCCL all-reduce of [{B}, {hidden}] bf16 after o_proj)",
                       "rope_store": "separate launches" if args.unfused else "fused (K4)"},
            "breakdown_ms": {"attention(K2+K3)": ms_attn, "rope+kv_store(K4)": ms_rope, "qkv_proj+o_proj(cuBLAS)": ms_gemm,
                             "all_reduce(NCCL)": ms_ar},
            "attention_gb_per_s_per_gpu": attn_bytes / (ms_attn * 1e-3) / 1e9,
            "tokens_per_s_model_equiv": B / (ms_full * 1e-3) / gcfg.layers_equiv, "layers_equiv": gcfg.layers_equiv,
            "clocks": clocks, "gpu_launches_own_per_step": launches,
        }
        print(json.dumps(line), flush=True)
    del g_full, g_attn, g_rope, g_gemm
    if world > 1:
        del g_ar
    R.shutdown()


if __name__ == "__main__":
    main()
