"""Per-CTA timeline of one decode_kernel launch (the decode counterpart of tools/trace_tc.py).

Needs the instrumented library:  make -C ltp-sglang_b200/csrc trace
    B200ATTN_LIB=ltp-sglang_b200/b200attn/libb200attn_trace.so python tools/trace_decode.py --config C2 [--tp-sim 8]
                 [--batch N] [--flush] [--out gpurun_out/trace_decode_C2.json]

Slots (csrc/decode_kernel.cuh, TR_*): %globaltimer at CTA entry / exit (ns, comparable across SMs), and the
CTA's own clock64 at: entry, barriers ready, producer's first piece known (plan reads done), first chunk's
copies issued (slot ids arrived), last chunk issued, consumer q loaded, first stage landed, first chunk
consumed, last chunk consumed, last epilogue written, merge started, exit.  Prints where the time of a launch
goes outside the steady-state stream: launch skew, prologue chain, pipeline fill, tail.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
from dataclasses import replace

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "ltp-sglang_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

SLOTS = ["gt_entry", "entry", "bars_ready", "p_walk", "p_ids", "p_first_issue", "p_last_issue", "c_q", "c_first_full",
         "c_first_done", "c_last_done", "c_epilogue", "c_merged", "c_exit", "counts", "gt_exit"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C2")
    ap.add_argument("--tp-sim", type=int, default=1)
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--page-size", type=int, default=0)
    ap.add_argument("--flush", action="store_true", help="flush L2 before the traced launch")
    ap.add_argument("--out", default="")
    ap.add_argument("--per-cta", action="store_true", help="also dump per-CTA arrays (SM id, steady us / chunk, lifetime)")
    ap.add_argument("--repeat", type=int, default=1, help="trace this many launches (per-CTA arrays of each are kept)")
    args = ap.parse_args()

    import torch

    import b200attn
    import synth
    from b200attn import capi
    from synth import CONFIGS
    from bench_harness import make_batch, make_layer, make_runner

    lib = capi.lib()
    if not hasattr(lib, "b200attn_debug_set_decode_trace"):
        raise SystemExit("not a trace build: make -C ltp-sglang_b200/csrc trace, then set B200ATTN_LIB")
    lib.b200attn_debug_set_decode_trace.argtypes = [C.c_void_p]
    lib.b200attn_debug_set_decode_trace.restype = None

    dev = torch.device("cuda:0")
    cfg = CONFIGS[args.config]
    if args.batch:
        cfg = replace(cfg, batch=args.batch)
    if args.page_size:
        cfg = replace(cfg, page_size=args.page_size)
    c = synth.shard(cfg, args.tp_sim)
    w = synth.make_workload(c, dev)
    runner = make_runner(w, dev)
    backend = b200attn.B200AttnBackend(runner)
    layer = make_layer(w)
    fb = make_batch(w, runner, backend)
    backend.init_forward_metadata(fb)
    B = c.batch
    fp8 = c.kv_dtype == "fp8_e4m3"
    q3 = w.q.view(B, c.num_q_heads, c.head_dim)
    o3 = torch.empty_like(q3)
    dargs = (q3, w.k_pool, w.v_pool, o3, None, backend.plan_buf, backend.forward_metadata.indices, backend.workspace,
             float(layer.scaling), 0.0, c.k_scale if fp8 else 1.0, c.v_scale if fp8 else 1.0, 0)
    for _ in range(3):
        torch.ops.b200attn.decode(*dargs)
    torch.cuda.synchronize()
    n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
    trace = torch.zeros((n_sm, 24), dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if args.flush:
        flush.zero_()
        flush.sum()
    per_cta = []
    for rep_i in range(args.repeat):
        backend.init_forward_metadata(fb)   # a server plans every step: ranges follow the recorded per-SM rates
        trace.zero_()
        if args.flush:
            flush.zero_()
            flush.sum()
        lib.b200attn_debug_set_decode_trace(C.c_void_p(trace.data_ptr()))
        e0.record()
        torch.ops.b200attn.decode(*dargs)
        e1.record()
        torch.cuda.synchronize()
        lib.b200attn_debug_set_decode_trace(None)
        tt = trace.cpu().numpy().astype(np.int64)
        ch = np.maximum((tt[:, 14] & 0xFFFFFFFF) - 1, 1)
        per_cta.append({"smid": tt[:, 4].tolist(), "clk_per_chunk": ((tt[:, 10] - tt[:, 9]) / ch).round(1).tolist(),
                        "lifetime_ns": (tt[:, 15] - tt[:, 0]).tolist(), "start_ns": (tt[:, 0] - tt[:, 0].min()).tolist()})
    t = trace.cpu().numpy().astype(np.int64)
    t = t[t[:, 1] != 0]          # CTAs that ran
    t = t[t[:, 13] != 0]         # ... and had work
    col = {n: t[:, i] for i, n in enumerate(SLOTS)}
    span_ns = (col["gt_exit"] - col["gt_entry"]).astype(np.float64)
    span_clk = (col["c_exit"] - col["entry"]).astype(np.float64)
    ghz = float(np.median(span_clk / np.maximum(span_ns, 1)))
    us = lambda clk: clk / ghz / 1e3   # noqa: E731

    def st(x):
        x = np.asarray(x, dtype=np.float64)
        return {"min": round(float(x.min()), 2), "med": round(float(np.median(x)), 2), "max": round(float(x.max()), 2)}

    chunks = col["counts"] & 0xFFFFFFFF
    pieces = col["counts"] >> 32
    g = capi.decode_geometry(c.num_q_heads, c.num_kv_heads, c.head_dim, backend.kv_dtype, batch=B)
    e_kv = 1 if fp8 else 2
    chunk_bytes = 2 * g["chunk_tokens"] * g["heads_per_cta"] * c.head_dim * e_kv
    steady = us(col["c_last_done"] - col["c_first_done"]) / np.maximum(chunks - 1, 1)
    kv_bytes = w.seq_lens_sum * c.num_kv_heads * 2 * c.head_dim * e_kv
    kernel_us = (col["gt_exit"].max() - col["gt_entry"].min()) / 1e3
    rep = {
        "config": f"{c.name} batch {B} tp_sim {args.tp_sim} page {c.page_size} {c.kv_dtype}", "ctas": int(len(t)),
        "geometry": g, "event_us": round(e0.elapsed_time(e1) * 1e3, 2), "kernel_us_globaltimer": round(float(kernel_us), 2),
        "kv_bytes": int(kv_bytes), "ideal_us_at_6545GBs": round(kv_bytes / 6545e3, 2), "sm_ghz": round(ghz, 3),
        "launch_skew_us": st((col["gt_entry"] - col["gt_entry"].min()) / 1e3),
        "entry_to_bars_ready_us": st(us(col["bars_ready"] - col["entry"])),
        "entry_to_first_piece_known_us": st(us(col["p_walk"] - col["entry"])),
        "entry_to_first_copies_issued_us": st(us(col["p_first_issue"] - col["entry"])),
        "entry_to_first_stage_landed_us": st(us(col["c_first_full"] - col["entry"])),
        "entry_to_first_chunk_consumed_us": st(us(col["c_first_done"] - col["entry"])),
        "chunks_per_cta": st(chunks), "pieces_per_cta": st(pieces),
        "steady_us_per_chunk": st(steady), "steady_gb_per_s_per_sm": st(chunk_bytes / np.maximum(steady, 1e-9) / 1e3),
        "stream_us_first_to_last_chunk": st(us(col["c_last_done"] - col["c_first_done"])),
        "producer_last_issue_to_consumer_last_done_us": st(us(col["c_last_done"] - col["p_last_issue"])),
        "last_chunk_to_exit_us": st(us(col["c_exit"] - col["c_last_done"])),
        "last_chunk_to_last_epilogue_us": st(us(col["c_epilogue"] - col["c_last_done"])),
        "last_epilogue_to_exit_us": st(us(col["c_exit"] - col["c_epilogue"])),
        "cta_lifetime_us": st(span_ns / 1e3),
        "idle_after_exit_us": st((col["gt_exit"].max() - col["gt_exit"]) / 1e3),
        "merging_ctas": int((col["c_merged"] != 0).sum()),
    }
    # one steady-state chunk of consumer warp 0 (slots 16-20), CTAs that consumed more than 64 chunks
    tk = t[t[:, 20] != 0]
    if len(tk):
        rep["chunk65_clk"] = {
            "wait_full": st(tk[:, 17] - tk[:, 16]), "qk": st(tk[:, 18] - tk[:, 17]), "softmax": st(tk[:, 19] - tk[:, 18]),
            "pv": st(tk[:, 20] - tk[:, 19]), "arrive": st(tk[:, 21] - tk[:, 20]), "to_next_wait": st(tk[:, 22] - tk[:, 21]),
            "total": st(tk[:, 22] - tk[:, 16])}
    if args.per_cta:
        rep["per_cta"] = per_cta
        if len(per_cta) > 1:   # is a slow SM always slow?  rank correlation of clk / chunk by SM id between launches
            a = dict(zip(per_cta[0]["smid"], per_cta[0]["clk_per_chunk"]))
            cors = []
            for other in per_cta[1:]:
                b = dict(zip(other["smid"], other["clk_per_chunk"]))
                ks = sorted(set(a) & set(b))
                cors.append(round(float(np.corrcoef([a[k] for k in ks], [b[k] for k in ks])[0, 1]), 3))
            rep["steady_rate_correlation_by_sm_vs_first_launch"] = cors
    print(json.dumps({k: v for k, v in rep.items() if k != "per_cta"}, indent=1))
    if args.out:
        with open(args.out, "w") as f:
            json.dump(rep, f, indent=1)
            f.write("\n")


if __name__ == "__main__":
    main()
