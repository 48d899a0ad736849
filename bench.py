"""bench.py -- paged decode-attention throughput on N B200s (one process per GPU).

    python bench.py --gpus 1 --steps K --warmup W
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the reference's own CPU kernel on the host cores

A "step" is one pass of the hot path over one decode batch: ``forward_decode`` of one attention
layer (KV store of the new token + split-KV decode + LSE merge) on the workload named in
``config.workload`` (BASELINE.json configs[1], "C2", unless --config says otherwise).  With N > 1
the KV heads are sharded over the ranks exactly as sglang's tensor parallelism does
(models/llama.py:119-132): every rank serves the same requests for its head slice, no collective
on the data path.  Default "weak" scaling: the decode batch is N x the config's, so each GPU keeps
the single-GPU configuration's KV bytes (what TP buys a server is room for N x the requests);
value = global batch / max-over-ranks time.  The fixed-batch ("strong") figure of the same
configuration is measured in the same run and reported under ``strong_scaling``.

JSON keys beyond the base contract: ``roofline`` (decode kernel alone, algorithmic bytes / CUDA-event
time vs the measured HBM peak), ``cpu_baseline`` (reference CPU kernel on a bounded sample),
``e2e`` (same metric through the backend's public calls with HOST buffers, every step: H2D of qkv +
request metadata from pinned memory, init_forward_metadata_replay_cuda_graph, CUDA-graph replay of the
layer call, D2H of o, stream sync -- the sequence CudaGraphRunner.replay drives; the eager variant is
reported next to it), ``clocks``, ``gpu_launches``.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time
from dataclasses import replace

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "ltp-sglang_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "paged decode-attn tokens/s (one attention layer step; HBM GB/s in roofline)"
UNIT = "tokens/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="C2", help="C1..C5 (synth.CONFIGS); the metric is quoted on C2")
    ap.add_argument("--page-size", type=int, default=0, help="override the config's page size")
    ap.add_argument("--batch", type=int, default=0, help="override the config's decode batch (a tuning aid; config.note says so)")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of a CUDA graph")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sustain", type=float, default=2.0,
                    help="seconds of back-to-back replays of the step graph after the timed steps (clocks / power "
                         "under sustained load, reported under 'sustained'); 0 = skip")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="N > 1: weak = batch x N with KV heads sharded (per-GPU bytes fixed); strong = fixed batch")
    ap.add_argument("--no-strong", action="store_true", help="N > 1: skip the extra strong-scaling measurement")
    ap.add_argument("--cpu-sample-requests", type=int, default=32)
    ap.add_argument("--tp-sim", type=int, default=1,
                    help="single-GPU dry run of ONE rank of an N-way KV-head sharding (per-rank shape of --gpus N); "
                         "a tuning aid, not a bench line: the JSON says so in config.note")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------- #
# clocks (sampled DURING the timed region, in-process through NVML)
# ----------------------------------------------------------------------------------------------- #
class ClockSampler:
    REASONS = {
        0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x10: "sync_boost",
        0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown",
        0x100: "display_clock_setting",
    }

    def __init__(self, index: int):
        self.samples, self.reason_bits, self.power = [], 0, []
        self.max_mhz = None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:  # NVML missing: report that, do not fake numbers
            self.nv, self.err = None, repr(e)

    def _loop(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                self.reason_bits |= nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
            except Exception:
                pass
            time.sleep(0.0005)

    def start(self):
        if self.nv:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self) -> dict:
        if not self.nv:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "error": self.err}
        self._stop.set()
        self._thread.join()
        reasons = [n for b, n in self.REASONS.items() if self.reason_bits & b]
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(self.samples),
                "power_w_max": max(self.power) if self.power else None}


# ----------------------------------------------------------------------------------------------- #
# reference arm / cpu baseline: the reference's own CPU decode kernel on the host cores
# ----------------------------------------------------------------------------------------------- #
def cpu_reference_run(cfg, sample_requests: int, steps: int, warmup: int):
    """Times the reference CPU path on a bounded sample (the first `sample_requests` requests of the
    same-shaped workload, full context length).  Returns (tokens/s, ms/step, info dict)."""
    import torch

    import synth
    from oracle import decode_oracle, ref_cpu_kernel

    n = min(sample_requests, cfg.batch)
    scfg = replace(cfg, batch=n)
    kind = "reference"
    note = ""
    if scfg.kv_dtype == "fp8_e4m3":
        # decode.cpp dispatches reduced floating types only (decode.cpp:1465): time it on an exactly
        # dequantised bf16 copy, as BASELINE.md section 3 prescribes.
        scfg = replace(scfg, kv_dtype="bf16")
        note = "; fp8 pool dequantised to bf16 (reference CPU kernel has no fp8 KV path)"
    torch.set_num_threads(os.cpu_count() or 1)
    w = synth.make_workload(scfg, "cpu")
    c = scfg
    q = w.q.reshape(c.batch, c.num_q_heads, c.head_dim)
    k = w.k.reshape(c.batch, c.num_kv_heads, c.head_dim)
    v = w.v.reshape(c.batch, c.num_kv_heads, c.head_dim)
    try:
        path = ref_cpu_kernel.load()
        attn_logits = torch.empty((c.batch, c.num_q_heads, ref_cpu_kernel.NUM_KV_SPLITS, c.head_dim + 1),
                                  dtype=torch.float32)

        def step():
            return ref_cpu_kernel.decode_attention_cpu(q, w.k_pool, w.v_pool, k, v, w.out_cache_loc,
                                                       w.req_to_token, w.req_pool_indices, w.seq_lens,
                                                       w.sm_scale, 0.0, attn_logits)
        impl = f"sgl-kernel decode_attention_cpu ({os.path.basename(path)})"
    except Exception as e:  # oracle/_ref missing or not runnable on this CPU: time the port
        kind = "port"
        impl = f"torch_native port (oracle/_ref unavailable: {str(e)[:80]})"

        def step():
            decode_oracle.kv_store(w.k_pool, w.v_pool, w.out_cache_loc, k, v)
            return decode_oracle.torch_native_decode(q, w.k_pool, w.v_pool, w.req_to_token,
                                                     w.req_pool_indices, w.seq_lens, w.sm_scale)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    tn = torch_native_cpu_run(scfg, min(8, n))
    info = {"kind": kind, "cores": torch.get_num_threads(), "torch_native": tn,
            "sample": f"{n} of {cfg.batch} requests of {cfg.name} at full context ({w.seq_lens_sum} tokens, "
                      f"{w.algorithmic_bytes() / 1e6:.0f} MB algorithmic), {steps} steps; {impl}{note}",
            "gb_per_s": w.algorithmic_bytes() / dt / 1e9}
    return n / dt, dt * 1e3, info


def torch_native_cpu_run(scfg, n_req: int, steps: int = 2):
    """The reference's OTHER CPU path: ``TorchNativeAttnBackend.forward_decode``
    (layers/attention/torch_native_backend.py:226-267: set_kv_buffer + one SDPA per request on gathered K/V),
    timed beside the sgl-kernel kernel on a smaller sample.  The reference's own class is used where its tree
    is mounted (this container); on the GPU box, where it is not, the line is the oracle's restatement of the
    same loop (``decode_oracle.torch_native_decode``) and says so."""
    import torch

    import synth
    from oracle import decode_oracle, ref_loader

    c = replace(scfg, batch=n_req)
    w = synth.make_workload(c, "cpu")
    q = w.q.reshape(c.batch, c.num_q_heads, c.head_dim)
    k = w.k.reshape(c.batch, c.num_kv_heads, c.head_dim)
    v = w.v.reshape(c.batch, c.num_kv_heads, c.head_dim)
    kind = "port"
    step = None
    if ref_loader.available():
        try:
            import types

            ref = ref_loader.load()
            r2t = ref.ReqToTokenPool(size=w.req_to_token.shape[0], max_context_len=w.req_to_token.shape[1], device="cpu",
                                     enable_memory_saver=False)
            r2t.req_to_token.copy_(w.req_to_token)
            pool = ref.MHATokenToKVPool(size=w.pool_size, page_size=c.page_size, dtype=w.k_pool.dtype, head_num=c.num_kv_heads,
                                        head_dim=c.head_dim, layer_num=1, device="cpu", enable_memory_saver=False)
            pool.k_buffer[0].copy_(w.k_pool)
            pool.v_buffer[0].copy_(w.v_pool)
            backend = ref.TorchNativeAttnBackend(types.SimpleNamespace(device="cpu"))
            fb = ref.ForwardBatch(forward_mode=ref.ForwardMode.DECODE, batch_size=c.batch, input_ids=torch.zeros(c.batch, dtype=torch.int64),
                                  req_pool_indices=w.req_pool_indices, seq_lens=w.seq_lens, out_cache_loc=w.out_cache_loc,
                                  seq_lens_sum=w.seq_lens_sum, req_to_token_pool=r2t, token_to_kv_pool=pool, attn_backend=backend)
            layer = ref.RadixAttention(c.num_q_heads, c.head_dim, w.sm_scale, c.num_kv_heads, layer_id=0)

            def step():
                return layer(w.q.clone(), w.k.clone(), w.v.clone(), fb)
            step()
            kind = "reference"
        except Exception as e:   # constructor signatures drift between reference versions: fall back, say why
            step = None
            kind = f"port (reference class not usable here: {str(e)[:60]})"
    if step is None:
        def step():
            decode_oracle.kv_store(w.k_pool, w.v_pool, w.out_cache_loc, k, v)
            return decode_oracle.torch_native_decode(q, w.k_pool, w.v_pool, w.req_to_token, w.req_pool_indices,
                                                     w.seq_lens, w.sm_scale)
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return {"value": n_req / dt, "unit": UNIT, "ms_per_step": dt * 1e3, "kind": kind, "cores": torch.get_num_threads(),
            "sample": f"{n_req} requests of {scfg.name} at full context, {steps} steps; TorchNativeAttnBackend.forward_decode"}


def run_reference_arm(args, cfg):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(args.steps, 20))
    tps, ms, info = cpu_reference_run(cfg, args.cpu_sample_requests, steps, min(args.warmup, 3))
    line = {
        "impl": "reference", "metric": METRIC, "value": tps, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": min(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True,
        "scaling": args.scaling, "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": workload_config(replace(cfg, batch=cfg.batch * (args.gpus if args.scaling == "weak" else 1)),
                                  args.gpus, note="reference CPU path on host cores; bounded sample"),
        "cpu_baseline": {"value": tps, "unit": UNIT, "cores": info["cores"], "kind": info["kind"],
                         "sample": info["sample"], "torch_native": info["torch_native"]},
        "e2e": {"value": tps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(cfg, n_gpus, note="", l2=""):
    return {"workload": f"{cfg.name}: batch {cfg.batch}, ctx "
                        f"{'%d-%d ragged' % cfg.ragged if cfg.ragged else cfg.ctx}, {cfg.num_q_heads}q/"
                        f"{cfg.num_kv_heads}kv heads, head_dim {cfg.head_dim}, KV {cfg.kv_dtype}, page_size "
                        f"{cfg.page_size}, random-init pools, permuted slots",
            "global_batch": cfg.batch, "seq_len": cfg.ctx if not cfg.ragged else list(cfg.ragged),
            "parallelism": f"tp{n_gpus} (KV-head sharding, no data-path collective)",
            "l2": l2, "note": note}


# ----------------------------------------------------------------------------------------------- #
# own arm
# ----------------------------------------------------------------------------------------------- #
L2_BYTES = 126 * 1024 * 1024


class Rank:
    """Process-wide state of one rank (one process per GPU)."""

    def __init__(self, args, device_type: str = "cuda"):
        """``device_type="cpu"`` (gloo) exists for the world_size-2 CPU tests of this plumbing."""
        import torch
        import torch.distributed as dist

        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        if self.world != args.gpus:
            raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={self.world}: launch with torchrun for N > 1")
        self.cuda = device_type == "cuda"
        if self.cuda:
            self.dev = torch.device("cuda", self.local_rank)
            torch.cuda.set_device(self.dev)
        else:
            self.dev = torch.device("cpu")
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            if self.cuda:
                dist.init_process_group("nccl", device_id=self.dev)
            else:
                dist.init_process_group("gloo")
        self.flush_buf = None

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        if self.cuda:
            self.torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        t = self.torch.tensor([x], device=self.dev, dtype=self.torch.float64)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def shutdown(self):
        """Leaves the process group without ever hanging the launcher: everything this rank had to
        report is printed by now, so if NCCL teardown stalls (seen with CUDA graphs that captured a
        collective still alive) a watchdog ends the process with exit code 0 after 20 s."""
        sys.stdout.flush()
        if self.world <= 1:
            return
        import gc

        def _bail():
            time.sleep(20)
            os._exit(0)

        threading.Thread(target=_bail, daemon=True).start()
        gc.collect()
        if self.cuda:
            self.torch.cuda.synchronize()
        self.dist.barrier()
        self.dist.destroy_process_group()

    def flush_l2(self):
        """Writes a buffer twice the size of L2, then reads a second one of the same size, so the
        next step starts from HBM with an L2 full of CLEAN lines.  (A write-only flush leaves
        126 MB of dirty lines whose write-back lands inside the next timed step: ~20 us of HBM
        time, as much as a small config's whole step.)"""
        torch = self.torch
        if self.flush_buf is None:
            self.flush_buf = torch.empty(2 * L2_BYTES, dtype=torch.uint8, device=self.dev)
            self.flush_src = torch.zeros(2 * L2_BYTES // 8, dtype=torch.int64, device=self.dev)
            self.flush_sink = torch.zeros((), dtype=torch.int64, device=self.dev)
        self.flush_buf.zero_()
        torch.sum(self.flush_src, dim=0, out=self.flush_sink)

    def timed(self, fn, steps: int, flush: bool) -> float:
        """ms per step by CUDA events on the current stream.  With `flush` every step is timed on
        its own (events around the step only) with an L2 flush in front of it."""
        torch = self.torch
        if not flush:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / steps
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for a, b in evs:
            self.flush_l2()
            a.record()
            fn()
            b.record()
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs) / steps


def measure(R: Rank, args, gcfg, scaling: str, with_clocks: bool = True) -> dict:
    """One workload on all ranks: device-resident value, decode-kernel roofline, e2e."""
    import synth
    import b200attn
    from b200attn import capi
    from bench_harness import make_batch, make_layer, make_runner

    torch, world, dev = R.torch, R.world, R.dev
    c = synth.shard(gcfg, world * args.tp_sim)           # this rank's head slice
    w = synth.make_workload(c, dev)                      # request tables identical on every rank
    runner = make_runner(w, dev, tp_size=world)
    backend = b200attn.B200AttnBackend(runner)
    layer = make_layer(w)
    fb = make_batch(w, runner, backend)
    B = c.batch
    fp8 = c.kv_dtype == "fp8_e4m3"
    e_kv = 1 if fp8 else 2
    kv_bytes = w.seq_lens_sum * c.num_kv_heads * 2 * c.head_dim * e_kv
    flush = kv_bytes < 4 * L2_BYTES                      # working set could (partly) live in L2
    # set_kv_buffer divides k / v by the scale IN PLACE when the pool is fp8 (memory_pool.py:385-389);
    # in a model the qkv projection rewrites that buffer every step, so the step does the same here
    # (one 0.5-3 MB device copy inside the timed region) instead of letting k drift to inf.
    qkv0 = w.qkv.clone()
    refresh = fp8
    l2_note = ("L2 flushed (252 MB memset, then a 252 MB read so no dirty lines remain) before every timed step; step time = "
               "(graph of 4 x [flush, step] - graph of 4 x [flush]) / 4 by CUDA events (ms_per_step_single_replay: event pair "
               "around one replay of a one-step graph, graph launch included)"
               if flush else "per-GPU KV working set (%d MB) far larger than the 126 MB L2; no flush" % (kv_bytes >> 20))

    # ---------------- device-resident timing: `value` ----------------
    # Eager warm-up the way a server runs: every decode step plans (init_forward_metadata), then runs the layers.
    # The plan sizes the per-SM ranges by the streaming rates the previous launches left in the plan buffer
    # (csrc/plan.cu), so a few steps are needed before the ranges are the steady-state ones.
    for _ in range(10):
        backend.init_forward_metadata(fb)
        if refresh:
            w.qkv.copy_(qkv0)
        out = layer(w.q, w.k, w.v, fb)
    torch.cuda.synchronize()
    graph = None
    launches_per_step = None
    if not args.no_graph:
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            layer(w.q, w.k, w.v, fb)
        torch.cuda.current_stream().wait_stream(s)
        n0 = capi.launch_count()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            if refresh:
                w.qkv.copy_(qkv0)
            out = layer(w.q, w.k, w.v, fb)
        launches_per_step = capi.launch_count() - n0

    def step():
        if graph is not None:
            graph.replay()
        else:
            if refresh:
                w.qkv.copy_(qkv0)
            layer(w.q, w.k, w.v, fb)

    if launches_per_step is None:
        n0 = capi.launch_count()
        step()
        launches_per_step = capi.launch_count() - n0
    warm = max(args.warmup, 3)
    for _ in range(warm):
        step()
    sampler = ClockSampler(R.local_rank) if with_clocks else None
    R.barrier()
    if sampler:
        sampler.start()
    ms = R.timed(step, args.steps, flush)
    R.barrier()
    n_timed = len(sampler.samples) if sampler else 0
    ms_single = None
    if flush and graph is not None:
        # Flushed configs (KV working set < 4 x L2): a step is 25-60 us, and an event pair around ONE replay of a
        # one-step graph also holds the graph launch (5-9 us) -- which a decode step inside a model's graph of
        # hundreds of kernels does not pay.  So the step is timed inside a graph: A = 4 x [L2 flush, step],
        # B = 4 x [L2 flush]; step = (A - B) / 4.  The single-replay figure is reported beside it.
        ms_single = R.max_over_ranks(ms)

        def _flush_step_graph(with_step: bool):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(4):
                    R.flush_l2()
                    if with_step:
                        if refresh:
                            w.qkv.copy_(qkv0)
                        layer(w.q, w.k, w.v, fb)
            for _ in range(2):
                g.replay()
            torch.cuda.synchronize()
            return g
        R.flush_l2()
        torch.cuda.synchronize()
        ga, gb = _flush_step_graph(True), _flush_step_graph(False)
        reps = max(10, min(args.steps, 30))
        ta = min(R.timed(ga.replay, reps, False) for _ in range(3))
        tb = min(R.timed(gb.replay, reps, False) for _ in range(3))
        ms = (ta - tb) / 4
        del ga, gb
        R.barrier()
    ms_max = R.max_over_ranks(ms)

    # ---------------- decode kernel alone: roofline ----------------
    md = backend.forward_metadata
    q3 = w.q.view(B, c.num_q_heads, c.head_dim)
    o3 = torch.empty((B, c.num_q_heads, c.head_dim), dtype=w.q.dtype, device=dev)
    pool = runner.token_to_kv_pool
    dargs = (q3, pool.k_buffer[0], pool.v_buffer[0], o3, None, backend.plan_buf, md.indices, backend.workspace,
             float(layer.scaling), 0.0, c.k_scale if fp8 else 1.0, c.v_scale if fp8 else 1.0, 0)
    for _ in range(3):
        torch.ops.b200attn.decode(*dargs)
    # K_PER launches per graph replay (a model's decode step is one graph of many kernels, not one graph per
    # kernel): the average launch duration excludes most of the graph-launch gap.  With an L2 flush in front of
    # every step only one launch per replay can be timed cold.
    K_PER = 1 if flush else 4
    kgraph = torch.cuda.CUDAGraph()
    torch.cuda.synchronize()
    with torch.cuda.graph(kgraph):
        for _ in range(K_PER):
            torch.ops.b200attn.decode(*dargs)
    for _ in range(3):
        kgraph.replay()
    torch.cuda.synchronize()
    k_ms = R.max_over_ranks(R.timed(kgraph.replay, max(10, min(args.steps, 50)), flush)) / K_PER
    k_ms_single = None
    if flush:
        # Flushed configs: one launch per replay can be timed cold, and the event pair around a one-kernel graph
        # also holds the graph launch (5-9 us, as much as a third of a 25-50 us kernel; ncu's launch list and the
        # globaltimer trace both show the kernel itself that much shorter).  A decode step runs the kernel inside a
        # graph of many, so it is timed that way: graph A = 4 x [L2 flush, decode], graph B = 4 x [L2 flush];
        # kernel = (A - B) / 4.  The single-launch figure stays beside it.
        k_ms_single = k_ms

        def _flush_graph(with_decode: bool):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(4):
                    R.flush_l2()
                    if with_decode:
                        torch.ops.b200attn.decode(*dargs)
            for _ in range(2):
                g.replay()
            torch.cuda.synchronize()
            return g
        R.flush_l2()
        torch.cuda.synchronize()
        ga, gb = _flush_graph(True), _flush_graph(False)
        reps = max(10, min(args.steps, 30))
        ta = min(R.timed(ga.replay, reps, False) for _ in range(3))
        tb = min(R.timed(gb.replay, reps, False) for _ in range(3))
        k_ms = R.max_over_ranks((ta - tb) / 4)
        del ga, gb
    clocks = None
    if sampler:  # sampled from the start of the timed steps to the end of the K2-alone timing
        clocks = sampler.stop()
        clocks["samples_in_timed_steps"] = n_timed
    # SURVEY.md 8(d) per-unit figure for the decode kernel: KV rows + q + o + index words
    decode_bytes = kv_bytes + 2 * B * c.num_q_heads * c.head_dim * 2 + 4 * w.seq_lens_sum
    peak, peak_src = _measured_peak()
    achieved = decode_bytes / (k_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": _ncu_traffic(gcfg.name, world), "kernel": "decode_kernel (split-KV decode with the LSE merge fused in; one launch), per GPU",
                "kernel_ms": k_ms, "kernel_ms_single_launch_graph": k_ms_single,
                "kernel_ms_method": ("(graph of 4 x [L2 flush, decode] - graph of 4 x [L2 flush]) / 4, CUDA events" if flush else
                                     "graph of %d back-to-back launches / %d, CUDA events" % (K_PER, K_PER)),
                "launches_per_timed_graph": K_PER, "algorithmic_bytes": decode_bytes, "peak_source": peak_src,
                "frac_of_nominal_8TBs": achieved / 8000.0}

    # ---------------- sustained replay: >= args.sustain seconds of the same step graph ----------------
    sustained = None
    if args.sustain > 0 and graph is not None and not flush and world == 1:
        ssamp = ClockSampler(R.local_rank)
        reps = max(100, int(args.sustain * 1e3 / ms) + 1)
        ssamp.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            graph.replay()
        e1.record()
        torch.cuda.synchronize()
        sms = e0.elapsed_time(e1) / reps
        sustained = {"seconds": e0.elapsed_time(e1) / 1e3, "steps": reps, "ms_per_step": sms, "value": B / (sms * 1e-3),
                     "hbm_gb_per_s": w.algorithmic_bytes() / (sms * 1e-3) / 1e9, "clocks": ssamp.stop()}

    # ---------------- the small kernels of the path: K1 kv_store, K4 rope + store, merge_states ----------------
    other = small_kernel_rooflines(R, w, c, pool, peak) if world == 1 and args.tp_sim == 1 else None

    # ---------------- end to end with host buffers: `e2e` ----------------
    # The call sequence CudaGraphRunner.replay makes every decode step (cuda_graph_runner.py:698-790):
    # copy the step's inputs into the static buffers (here: from pinned HOST memory), re-plan through
    # init_forward_metadata_replay_cuda_graph, replay the captured layer, read the result on the host.
    # One packed pinned staging buffer [qkv | req_pool_indices | seq_lens | out_cache_loc] and one device
    # mirror: a step's inputs cross PCIe / C2C as ONE copy (four separate copies cost ~4 x 8 us of launch +
    # DMA set-up for 2.6 MB).
    def _al(n):
        return (n + 255) // 256 * 256
    nb_qkv = w.qkv.numel() * w.qkv.element_size()
    parts = [("qkv", w.qkv.dtype, tuple(w.qkv.shape), nb_qkv)] + \
            [(n, torch.int64, (B,), 8 * B) for n in ("rpi", "sl", "loc")]
    offs, total_b = {}, 0
    for name, dt, shape, nb in parts:
        offs[name] = (total_b, nb, dt, shape)
        total_b = _al(total_b + nb)
    host_buf = torch.empty(total_b, dtype=torch.uint8).pin_memory()
    dev_buf = torch.empty(total_b, dtype=torch.uint8, device=dev)

    def _view(buf, name):
        o, nb, dt, shape = offs[name]
        return buf[o:o + nb].view(dt).view(shape)
    _view(host_buf, "qkv").copy_(qkv0.cpu())
    _view(host_buf, "rpi").copy_(w.req_pool_indices.cpu())
    _view(host_buf, "sl").copy_(w.seq_lens.cpu())
    _view(host_buf, "loc").copy_(w.out_cache_loc.cpu())
    host_out = torch.empty((B, c.num_q_heads * c.head_dim), dtype=w.q.dtype).pin_memory()
    d_qkv, d_rpi, d_sl, d_loc = (_view(dev_buf, n) for n in ("qkv", "rpi", "sl", "loc"))
    qs, ks, vs = d_qkv.split([c.num_q_heads * c.head_dim, c.num_kv_heads * c.head_dim,
                              c.num_kv_heads * c.head_dim], dim=-1)
    fb2 = make_batch(w, runner, backend)
    fb2.req_pool_indices, fb2.seq_lens, fb2.out_cache_loc = d_rpi, d_sl, d_loc
    h2d = total_b
    d2h = host_out.numel() * host_out.element_size()
    mode = b200attn.ForwardMode.DECODE

    def h2d_copies():
        dev_buf.copy_(host_buf, non_blocking=True)

    def e2e_eager():
        h2d_copies()
        backend.init_forward_metadata(fb2)
        o = layer(qs, ks, vs, fb2)
        host_out.copy_(o, non_blocking=True)
        torch.cuda.current_stream().synchronize()   # the caller reads the result on the host

    e2e_eager()
    egraph, e_out = None, None
    if not args.no_graph:
        backend.init_cuda_graph_state(B, B)
        backend.init_forward_metadata_capture_cuda_graph(B, B, d_rpi, d_sl, None, mode, None)
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            layer(qs, ks, vs, fb2)
        torch.cuda.current_stream().wait_stream(s)
        egraph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(egraph):
            e_out = layer(qs, ks, vs, fb2)

    def e2e_graph():
        h2d_copies()
        backend.init_forward_metadata_replay_cuda_graph(B, d_rpi, d_sl, w.seq_lens_sum, None, mode, None, None)
        egraph.replay()
        host_out.copy_(e_out, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    # One graph for the whole step: the H2D copies out of the pinned host buffers, the plan (the
    # replay hook is capture-safe: no sync, no allocation), the layer call and the D2H of o are all
    # nodes of one CUDA graph, so a decode step costs one graph launch + one stream sync on the host.
    wgraph = None
    if not args.no_graph:
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            h2d_copies()
            backend.init_forward_metadata_replay_cuda_graph(B, d_rpi, d_sl, w.seq_lens_sum, None, mode, None, None)
            layer(qs, ks, vs, fb2)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        wgraph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(wgraph):
            h2d_copies()
            backend.init_forward_metadata_replay_cuda_graph(B, d_rpi, d_sl, w.seq_lens_sum, None, mode, None, None)
            w_out = layer(qs, ks, vs, fb2)
            host_out.copy_(w_out, non_blocking=True)

    def e2e_one_graph():
        wgraph.replay()
        torch.cuda.current_stream().synchronize()

    def time_e2e(fn):
        for _ in range(3):
            fn()
        R.barrier()
        n0 = capi.launch_count()
        if flush:
            t = 0.0
            for _ in range(args.steps):
                R.flush_l2()
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                fn()                       # ends with a stream synchronize: wall clock == device + host time
                t += time.perf_counter() - t0
            ms_ = t * 1e3 / args.steps
        else:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.steps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            ms_ = e0.elapsed_time(e1) / args.steps
        n = capi.launch_count() - n0
        R.barrier()
        return R.max_over_ranks(ms_), n

    eager_ms, eager_launches = time_e2e(e2e_eager)
    variants = {"eager": eager_ms}
    if egraph is not None:
        hook_ms, hook_launches = time_e2e(e2e_graph)
        variants["graph_replay_after_eager_hook"] = hook_ms
        host_out.zero_()
        e2e_ms, e2e_launches = time_e2e(e2e_one_graph)
        variants["one_graph"] = e2e_ms
        # Headline: the sequence the unmodified CudaGraphRunner drives (replay_prepare runs the backend's replay
        # hook eagerly, then graph.replay(): cuda_graph_runner.py:760-768, 783-790).  The one-graph form (copies,
        # plan, layer and D2H as nodes of a single graph) is reported beside it under variants_ms_per_step.
        e2e_ms, e2e_launches = hook_ms, hook_launches
        e2e_path = ("ONE packed H2D(qkv | req_pool_indices | seq_lens | out_cache_loc) from pinned host memory -> "
                    "init_forward_metadata_replay_cuda_graph (eager plan launch) -> CUDA-graph replay of "
                    "RadixAttention.forward (kv_store, decode + merge) -> D2H(o) -> stream sync")
    else:
        e2e_ms, e2e_launches = eager_ms, eager_launches
        e2e_path = "H2D -> init_forward_metadata -> RadixAttention.forward (eager) -> D2H(o) -> stream sync"
    torch.cuda.synchronize()
    assert torch.isfinite(host_out.float()).all(), "e2e output is not finite"
    # the e2e result must be the same attention output the device-resident path produced -- to rounding: every
    # plan sizes the CTA ranges by the streaming rates recorded so far, so two plans of the same batch may cut the
    # KV sequence at different places (fp32 reassociation, 1 bf16 ulp at most on |o| <= 0.25)
    if not fp8:
        diff = (host_out.to(dev).view(-1).float() - out.view(-1).float()).abs().max().item()
        assert diff <= 4e-3, f"e2e output differs from the device-resident step by {diff}"

    total_bytes = w.algorithmic_bytes() * world
    res = {
        "value": B / (ms_max * 1e-3), "ms_per_step": ms_max, "scaling": scaling,
        "dtype": "bf16" if not fp8 else "fp8_e4m3 KV / bf16 q",
        "config": workload_config(gcfg, world, l2=l2_note,
                                  note="" if args.tp_sim == 1 else f"DRY RUN: one rank of a tp{args.tp_sim} sharding on 1 GPU"),
        "hbm_gb_per_s": total_bytes / (ms_max * 1e-3) / 1e9,
        "tokens_per_s_model_equiv": B / (ms_max * 1e-3) / gcfg.layers_equiv,
        "layers_equiv": gcfg.layers_equiv,
        "launch_mode": "cuda_graph" if graph is not None else "eager",
        "ms_per_step_single_replay": ms_single,
        "roofline": roofline,
        "e2e": {"value": B / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_ms, "launches": e2e_launches, "path": e2e_path,
                "eager_value": B / (eager_ms * 1e-3), "eager_ms_per_step": eager_ms,
                "variants_ms_per_step": variants},
        "gpu_launches": launches_per_step * args.steps,
        "clocks": clocks,
    }
    if sustained is not None:
        res["sustained"] = sustained
    if other is not None:
        res["other_kernels"] = other
    if w.unique_tokens != w.seq_lens_sum:
        # shared-prefix pages are read once per request that names them (logical bytes) but live in
        # HBM once: the unique-byte figure is what DRAM must deliver at least, the rest can come from L2
        ub = w.unique_tokens * c.num_kv_heads * 2 * c.head_dim * e_kv + decode_bytes - kv_bytes
        res["roofline"]["unique_bytes"] = ub
        res["roofline"]["achieved_unique"] = ub / (k_ms * 1e-3) / 1e9
        res["roofline"]["frac_unique"] = ub / (k_ms * 1e-3) / 1e9 / peak
    del graph, kgraph, egraph, wgraph, backend, runner, w, fb, fb2
    torch.cuda.empty_cache()
    return res


def main():
    args = parse_args()
    import synth

    cfg = synth.CONFIGS[args.config]
    if args.page_size:
        cfg = replace(cfg, page_size=args.page_size)
    if args.batch:
        cfg = replace(cfg, name=f"{cfg.name}(batch {args.batch})", batch=args.batch)
    if args.impl == "reference":
        return run_reference_arm(args, cfg)

    R = Rank(args)
    world = R.world
    # N > 1: KV heads are sharded over the ranks (sglang TP, models/llama.py:119-132), no data-path
    # collective.  Weak scaling (default): the decode batch grows with N (global batch = N x the
    # config's), so every GPU keeps the single-GPU configuration's KV bytes -- TP is what lets a
    # server hold N x the requests.  Strong scaling (fixed batch, 1/N of the bytes per GPU) is
    # measured in the same run and reported under "strong_scaling".
    if args.scaling == "weak":
        gcfg = replace(cfg, batch=cfg.batch * world * args.tp_sim)
    else:
        gcfg = cfg
    res = measure(R, args, gcfg, args.scaling)
    strong = None
    if world > 1 and args.scaling == "weak" and not args.no_strong:
        s = measure(R, args, cfg, "strong", with_clocks=False)
        strong = {k: s[k] for k in ("value", "ms_per_step", "hbm_gb_per_s", "config")}
        strong["unit"] = UNIT
        strong["roofline"] = {k: s["roofline"][k] for k in ("achieved", "frac", "kernel_ms", "algorithmic_bytes")}
        strong["e2e"] = {k: s["e2e"][k] for k in ("value", "ms_per_step")}

    if R.rank == 0:
        line = {"metric": METRIC, "value": res.pop("value"), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": res.pop("ms_per_step"), "higher_is_better": True,
                "scaling": res.pop("scaling"), "vs_baseline": None, "dtype": res.pop("dtype"),
                "data": "synthetic", "config": res.pop("config")}
        line.update(res)
        if strong is not None:
            line["strong_scaling"] = strong
        if world == 1 and not args.no_cpu_baseline:
            tps, cms, info = cpu_reference_run(cfg, args.cpu_sample_requests, 5, 2)
            line["cpu_baseline"] = {"value": tps, "unit": UNIT, "cores": info["cores"], "kind": info["kind"],
                                    "sample": info["sample"], "ms_per_step": cms, "gb_per_s": info["gb_per_s"],
                                    "torch_native": info["torch_native"]}
        print(json.dumps(line), flush=True)
    R.shutdown()


def small_kernel_rooflines(R, w, c, pool, peak):
    """K1 kv_store, K4 rope_kv_store and merge_states at the workload's shape: bytes moved / CUDA-event time of
    8 launches per graph replay.  MB-sized launches: their time is launch latency, not bandwidth -- the lines
    exist so that nothing on the path goes unmeasured."""
    import b200attn

    torch, dev = R.torch, R.dev
    B, Hq, Hkv, D = c.batch, c.num_q_heads, c.num_kv_heads, c.head_dim
    fp8 = c.kv_dtype == "fp8_e4m3"
    e_kv = 1 if fp8 else 2
    q3, k3, v3 = w.q.view(B, Hq, D), w.k.view(B, Hkv, D), w.v.view(B, Hkv, D)
    kp, vp = pool.k_buffer[0], pool.v_buffer[0]
    pos = (w.seq_lens - 1).to(torch.int64)
    import synth
    cache = synth.cos_sin_cache(D, int(w.seq_lens_np.max()) + 8).to(dev)
    ks = c.k_scale if fp8 else None
    vs = c.v_scale if fp8 else None
    oa = torch.randn((B, Hq, D), device=dev).to(w.q.dtype)
    ob = torch.randn((B, Hq, D), device=dev).to(w.q.dtype)
    la, lb = torch.randn((B, Hq), device=dev), torch.randn((B, Hq), device=dev)
    oo, lo = torch.empty_like(oa), torch.empty_like(la)
    cases = [
        ("kv_store_kernel (K1)", lambda: b200attn.store_kv(kp, vp, w.out_cache_loc, k3, v3, ks, vs),
         2 * B * Hkv * D * (2 + e_kv) + 8 * B),
        ("rope_kv_store_kernel (K4)", lambda: b200attn.rope_store_kv(q3, k3, v3, pos, cache, True, w.out_cache_loc, kp, vp, ks, vs),
         2 * B * Hq * D * 2 + 2 * B * Hkv * D * (2 + e_kv) + B * (8 + 8 + 4 * D)),
        ("merge_states_kernel (merge_state_v2)", lambda: torch.ops.b200attn.merge_states(oa, la, ob, lb, oo, lo),
         3 * B * Hq * D * 2 + 3 * B * Hq * 4),
    ]
    out = []
    qkv0 = w.qkv.clone()
    for name, fn, nbytes in cases:
        fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(8):
                fn()
        for _ in range(3):
            g.replay()
        ms = R.timed(g.replay, 20, False) / 8
        w.qkv.copy_(qkv0)          # RoPE / fp8 div_ rotate and scale q, k, v in place
        gbs = nbytes / (ms * 1e-3) / 1e9
        out.append({"kernel": name, "kernel_ms": ms, "algorithmic_bytes": nbytes, "achieved": gbs, "unit": "GB/s",
                    "frac": gbs / peak, "bound": "hbm (launch-latency dominated at this size)"})
    return out


def _measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy, of measured)"
    except Exception:
        return 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"


def _ncu_traffic(config: str, world: int):
    """dram read+write bytes per launch of the decode kernel from the committed ncu capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(f"{config}_tp{world}")
    except Exception:
        return None


if __name__ == "__main__":
    main()
