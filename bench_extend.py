"""bench_extend.py -- extend / prefill attention throughput of the native path.

    python bench_extend.py [--config P1|P2|P3|all] [--steps K] [--warmup W] [--sdpa]

Secondary bench (the driver's contract is bench.py): one JSON line per configuration.  A "step" is
one ``forward_extend`` of one attention layer (multi-token KV store + the extend kernel) on a
Llama-3.1-8B-shaped (32 q / 8 kv heads, d128, bf16) EXTEND batch:

    P1  one request, 8192 new tokens, no cached prefix              (pure prefill; tensor bound)
    P2  16 requests x 2048 new tokens over a 2048-token prefix      (chunked prefill / cache hit)
    P3  128 requests x 8 draft tokens over a 4096-token prefix      (TARGET_VERIFY shape; HBM bound)

``roofline`` follows the bound of each shape: tensor (algorithmic flops = 4 D Hq x visible q-k pairs,
against the measured dense bf16 rate in MEASURED_PEAKS.json) for P1 / P2, HBM (unique KV bytes) for
P3.  ``--sdpa`` adds torch SDPA (cuDNN / flash library kernels) on the same requests laid out
contiguously -- a library competitor on the same box, not the reference.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "ltp-sglang_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

CONFIGS = {
    "P1": dict(prefix=[0], extend=[8192], note="1 x 8192 new tokens, no prefix"),
    "P2": dict(prefix=[2048] * 16, extend=[2048] * 16, note="16 x 2048 new tokens over 2048 cached"),
    "P3": dict(prefix=[4096] * 128, extend=[8] * 128, note="128 x 8 draft tokens over 4096 cached (verify shape)"),
    # Llama-3.2-1B shape (head_dim 64, BASELINE configs[0]'s model): the mma.sync kernel
    "P1d64": dict(prefix=[0], extend=[8192], d=64, note="1 x 8192 new tokens, no prefix, head_dim 64"),
    # the same shapes over an fp8-e4m3 pool with per-tensor scales (C3's cache dtype)
    "P2f": dict(prefix=[2048] * 16, extend=[2048] * 16, kv="fp8_e4m3",
                note="16 x 2048 new tokens over 2048 cached, fp8-e4m3 pool"),
    "P3f": dict(prefix=[4096] * 128, extend=[8] * 128, kv="fp8_e4m3",
                note="128 x 8 draft tokens over 4096 cached (verify shape), fp8-e4m3 pool"),
}


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return float(d["bf16_tflops"]), float(d["hbm_gbs"]), "MEASURED_PEAKS.json (burst)"
    except Exception:
        return 1600.0, 6650.0, "B200_PROFILING.md fallback"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="all")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--sdpa", action="store_true")
    ap.add_argument("--flashinfer", action="store_true",
                    help="also time flashinfer's BatchPrefillWithPagedKVCacheWrapper on the same paged pool (the reference's "
                         "prefill route on sm_100, flashinfer_backend.py:155-160, 450-532); a competitor, not an oracle")
    args = ap.parse_args()

    import numpy as np
    import torch

    import b200attn
    import synth
    from bench import ClockSampler
    from bench_harness import make_extend_runner

    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    tf_peak, hbm_peak, src = _peaks()
    names = list(CONFIGS) if args.config == "all" else args.config.split(",")
    for name in names:
        spec = CONFIGS[name]
        fp8 = spec.get("kv") == "fp8_e4m3"
        D = spec.get("d", 128)
        cfg = synth.Config(name, len(spec["prefix"]), 0, 32, 8, D, seed=77, kv_dtype=spec.get("kv", "bf16"),
                           k_scale=0.5 if fp8 else 1.0, v_scale=2.0 if fp8 else 1.0)
        w = synth.make_extend_workload(cfg, spec["prefix"], spec["extend"], dev)
        runner = make_extend_runner(w, dev)
        backend = b200attn.B200AttnBackend(runner)
        layer = b200attn.AttnLayer(32, D, w.sm_scale, 8, layer_id=0)
        if fp8:
            layer.k_scale = torch.tensor(cfg.k_scale, dtype=torch.float32, device=dev)
            layer.v_scale = torch.tensor(cfg.v_scale, dtype=torch.float32, device=dev)
            layer.k_scale_float, layer.v_scale_float = cfg.k_scale, cfg.v_scale
        B = len(spec["prefix"])
        fb = b200attn.ForwardBatch(
            forward_mode=b200attn.ForwardMode.EXTEND, batch_size=B, req_pool_indices=w.req_pool_indices,
            seq_lens=w.seq_lens, out_cache_loc=w.out_cache_loc, seq_lens_sum=int(w.seq_lens.sum()),
            req_to_token_pool=runner.req_to_token_pool, token_to_kv_pool=runner.token_to_kv_pool,
            attn_backend=backend, extend_prefix_lens=w.extend_prefix_lens, extend_seq_lens=w.extend_seq_lens,
            extend_seq_lens_cpu=[int(x) for x in w.extend_lens_np])
        backend.init_forward_metadata(fb)

        def step():
            return layer(w.q, w.k, w.v, fb, save_kv_cache=True)

        def kernel_only():
            return layer(w.q, w.k, w.v, fb, save_kv_cache=False)

        def timed(fn, steps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / steps

        for _ in range(max(args.warmup, 3)):
            step()
        torch.cuda.synchronize()
        n0 = b200attn.capi.launch_count()
        clocks = ClockSampler(0)
        clocks.start()
        ms = timed(step, args.steps)
        launches = b200attn.capi.launch_count() - n0
        kms = timed(kernel_only, args.steps)
        clk = clocks.stop()

        flops = w.flops(causal=True)
        kv_bytes = int((w.prefix_lens_np + w.extend_lens_np).sum()) * 8 * 2 * D * (1 if fp8 else 2)
        io_bytes = kv_bytes + 2 * w.tokens * 32 * D * 2
        tflops = flops / (kms * 1e-3) / 1e12
        gbs = io_bytes / (kms * 1e-3) / 1e9
        tensor_bound = flops / (tf_peak * 1e12) >= io_bytes / (hbm_peak * 1e9)
        line = {
            "metric": "extend-attn tokens/s (one attention layer step: KV store + extend kernel)",
            "value": w.tokens / (ms * 1e-3), "unit": "tokens/s", "n_gpus": 1, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": f"{name}: {spec['note']}; 32q/8kv heads, head_dim {D}, {'fp8-e4m3' if fp8 else 'bf16'} pool, page_size 1, "
                                   "permuted slots", "l2": "working set > L2 for P2/P3; P1 (32 MB KV) is L2-resident "
                                                           "by nature: a prefill re-reads its own KV"},
            "roofline": {"bound": "tensor" if tensor_bound else "hbm",
                         "achieved": tflops if tensor_bound else gbs,
                         "peak": tf_peak if tensor_bound else hbm_peak,
                         "unit": "TFLOP/s" if tensor_bound else "GB/s",
                         "frac": (tflops / tf_peak) if tensor_bound else (gbs / hbm_peak),
                         "traffic": None, "kernel": "extend_kernel", "kernel_ms": kms,
                         "algorithmic_flops": flops, "algorithmic_bytes": io_bytes, "peak_source": src,
                         "tflops": tflops, "gb_per_s": gbs},
            "gpu_launches": int(launches), "clocks": clk,
        }
        if args.sdpa:
            # library competitor: contiguous (unpaged) causal SDPA per request shape, bottom-right aligned
            import torch.nn.functional as F

            Le, Lp = spec["extend"][0], spec["prefix"][0]
            q = torch.randn(B, 32, Le, D, device=dev, dtype=torch.bfloat16)
            k = torch.randn(B, 8, Lp + Le, D, device=dev, dtype=torch.bfloat16)
            v = torch.randn_like(k)
            mask = None
            if Lp > 0:
                i = torch.arange(Le, device=dev)[:, None]
                j = torch.arange(Lp + Le, device=dev)[None, :]
                mask = (j - Lp <= i)

            def sdpa():
                return F.scaled_dot_product_attention(q, k, v, attn_mask=mask, is_causal=mask is None, enable_gqa=True)

            try:
                for _ in range(3):
                    sdpa()
                sms = timed(sdpa, args.steps)
                line["sdpa_contiguous"] = {"ms": sms, "tflops": flops / (sms * 1e-3) / 1e12,
                                           "note": "torch SDPA on contiguous K/V (no paging, no KV store)"}
            except Exception as e:  # noqa: BLE001
                line["sdpa_contiguous"] = {"error": repr(e)[:200]}
        if args.flashinfer and not fp8:
            # the same requests through flashinfer's paged prefill: kv list = the request's prefix + new slots
            # (req_to_token[req, :prefix + extend], the new tokens stored by the step above), causal, page_size 1
            try:
                import flashinfer

                sl = (w.extend_prefix_lens + w.extend_seq_lens).to(torch.int64)
                kv_indptr = torch.zeros(B + 1, dtype=torch.int32, device=dev)
                kv_indptr[1:] = torch.cumsum(sl, 0)
                rows = torch.repeat_interleave(w.req_pool_indices.to(torch.int64), sl)
                pos = torch.arange(int(sl.sum()), device=dev) - torch.repeat_interleave(kv_indptr[:-1].to(torch.int64), sl)
                kv_indices = runner.req_to_token_pool.req_to_token[rows, pos].to(torch.int32)
                qo_indptr = torch.zeros(B + 1, dtype=torch.int32, device=dev)
                qo_indptr[1:] = torch.cumsum(w.extend_seq_lens.to(torch.int64), 0)
                q3 = w.q.view(w.tokens, 32, D).contiguous()
                kc, vc = w.k_pool.unsqueeze(1), w.v_pool.unsqueeze(1)
                ours = kernel_only().view(w.tokens, 32, D).float()
                res = {}
                for be in ("cutlass", "auto"):
                    try:
                        ws = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
                        fw = flashinfer.BatchPrefillWithPagedKVCacheWrapper(ws, "NHD", backend=be)
                        fw.plan(qo_indptr, kv_indptr, kv_indices, torch.ones(B, dtype=torch.int32, device=dev), 32, 8, D, 1,
                                causal=True, sm_scale=w.sm_scale, q_data_type=w.q.dtype, kv_data_type=w.k_pool.dtype)
                        o = fw.run(q3, (kc, vc))
                        torch.cuda.synchronize()
                        for _ in range(3):
                            fw.run(q3, (kc, vc))
                        fms = timed(lambda: fw.run(q3, (kc, vc)), args.steps)
                        res[be] = {"ms": fms, "tflops": flops / (fms * 1e-3) / 1e12,
                                   "max_abs_diff_vs_b200attn": float((o.float() - ours).abs().max())}
                    except Exception as e:  # noqa: BLE001 -- a backend that does not run here is recorded
                        res[be] = {"error": repr(e)[:200]}
                if int(w.extend_prefix_lens.sum()) == 0:
                    # no cached prefix: the reference's sm_100 route for the extend part is the RAGGED wrapper with
                    # backend="cutlass" (flashinfer_backend.py:155-160) on contiguous k / v of the new tokens
                    k3 = w.k.view(w.tokens, 8, D).contiguous()
                    v3 = w.v.view(w.tokens, 8, D).contiguous()
                    for be in ("cutlass", "auto"):
                        try:
                            ws = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
                            rw = flashinfer.BatchPrefillWithRaggedKVCacheWrapper(ws, "NHD", backend=be)
                            rw.plan(qo_indptr, qo_indptr, 32, 8, D, causal=True, sm_scale=w.sm_scale,
                                    q_data_type=w.q.dtype, kv_data_type=w.q.dtype)
                            o = rw.run(q3, k3, v3)
                            torch.cuda.synchronize()
                            for _ in range(3):
                                rw.run(q3, k3, v3)
                            fms = timed(lambda: rw.run(q3, k3, v3), args.steps)
                            res["ragged_" + be] = {"ms": fms, "tflops": flops / (fms * 1e-3) / 1e12,
                                                   "max_abs_diff_vs_b200attn": float((o.float() - ours).abs().max())}
                        except Exception as e:  # noqa: BLE001
                            res["ragged_" + be] = {"error": repr(e)[:200]}
                line["flashinfer_paged_prefill"] = {"version": flashinfer.__version__, "backends": res,
                                                    "note": "run() only (no KV store), plan() outside the timed region, page_size 1"}
            except Exception as e:  # noqa: BLE001
                line["flashinfer_paged_prefill"] = {"error": repr(e)[:200]}
        print(json.dumps(line), flush=True)
        del w, runner, backend
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
