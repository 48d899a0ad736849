"""bench.py -- paged decode-attention throughput on N B200s (one process per GPU).

    python bench.py --gpus 1 --steps K --warmup W
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the reference's own CPU kernel on the host cores

A "step" is one pass of the hot path over one decode batch: ``forward_decode`` of one attention
layer (KV store of the new token + split-KV decode + LSE merge) on the workload named in
``config.workload`` (BASELINE.json configs[1], "C2", unless --config says otherwise).  With N > 1
the KV heads are sharded over the ranks exactly as sglang's tensor parallelism does
(models/llama.py:119-132): every rank serves the same requests for its head slice, no collective
on the data path, value = requests / max-over-ranks time ("strong" scaling of a fixed batch).

JSON keys beyond the base contract: ``roofline`` (decode kernel alone, algorithmic bytes / CUDA-event
time vs the measured HBM peak), ``cpu_baseline`` (reference CPU kernel on a bounded sample),
``e2e`` (same metric through the backend's public calls with HOST buffers: H2D of q/k/v + metadata,
plan, forward_decode, D2H of o, every step), ``clocks``, ``gpu_launches``.
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
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of a CUDA graph")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample-requests", type=int, default=32)
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
            time.sleep(0.002)

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
    info = {"kind": kind, "cores": torch.get_num_threads(),
            "sample": f"{n} of {cfg.batch} requests of {cfg.name} at full context ({w.seq_lens_sum} tokens, "
                      f"{w.algorithmic_bytes() / 1e6:.0f} MB algorithmic), {steps} steps; {impl}{note}",
            "gb_per_s": w.algorithmic_bytes() / dt / 1e9}
    return n / dt, dt * 1e3, info


def run_reference_arm(args, cfg):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(args.steps, 20))
    tps, ms, info = cpu_reference_run(cfg, args.cpu_sample_requests, steps, min(args.warmup, 3))
    line = {
        "impl": "reference", "metric": METRIC, "value": tps, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": min(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": workload_config(cfg, args.gpus, note="reference CPU path on host cores; bounded sample"),
        "cpu_baseline": {"value": tps, "unit": UNIT, "cores": info["cores"], "kind": info["kind"],
                         "sample": info["sample"]},
        "e2e": {"value": tps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(cfg, n_gpus, note=""):
    return {"workload": f"{cfg.name}: batch {cfg.batch}, ctx "
                        f"{'%d-%d ragged' % cfg.ragged if cfg.ragged else cfg.ctx}, {cfg.num_q_heads}q/"
                        f"{cfg.num_kv_heads}kv heads, head_dim {cfg.head_dim}, KV {cfg.kv_dtype}, page_size "
                        f"{cfg.page_size}, random-init pools, permuted slots",
            "global_batch": cfg.batch, "seq_len": cfg.ctx if not cfg.ragged else list(cfg.ragged),
            "parallelism": f"tp{n_gpus} (KV-head sharding, no data-path collective)",
            "l2": "per-GPU KV working set larger than the 126 MB L2 (no flush needed)", "note": note}


# ----------------------------------------------------------------------------------------------- #
# own arm
# ----------------------------------------------------------------------------------------------- #
def main():
    args = parse_args()
    import synth

    cfg = synth.CONFIGS[args.config]
    if args.page_size:
        cfg = replace(cfg, page_size=args.page_size)
    if args.impl == "reference":
        return run_reference_arm(args, cfg)

    import torch
    import torch.distributed as dist

    import b200attn
    from b200attn import capi
    from tests.harness import make_batch, make_layer, make_runner

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun for N > 1")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    rcfg = synth.shard(cfg, world)                       # this rank's head slice
    rcfg = replace(rcfg, seed=cfg.seed + 1000 * rank)    # own pool contents, identical request tables
    w = synth.make_workload(replace(rcfg, seed=cfg.seed), dev) if world == 1 else _rank_workload(synth, rcfg, cfg, dev)
    runner = make_runner(w, dev, tp_size=world)
    backend = b200attn.B200AttnBackend(runner)
    layer = make_layer(w)
    fb = make_batch(w, runner, backend)
    c = rcfg
    B = c.batch

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident timing: `value` ----------------
    backend.init_forward_metadata(fb)
    out = layer(w.q, w.k, w.v, fb)            # eager warm-up, also sizes everything
    torch.cuda.synchronize()
    launches0 = capi.launch_count()
    graph = None
    if not args.no_graph:
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            layer(w.q, w.k, w.v, fb)
        torch.cuda.current_stream().wait_stream(s)
        launches0 = capi.launch_count()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = layer(w.q, w.k, w.v, fb)
    launches_per_step = capi.launch_count() - launches0 if graph is not None else 3

    def step():
        if graph is not None:
            graph.replay()
        else:
            layer(w.q, w.k, w.v, fb)

    for _ in range(max(args.warmup, 3)):
        step()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1) / args.steps
    ms_t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
    ms_max = float(ms_t.item())

    # ---------------- decode kernel alone: roofline ----------------
    md = backend.forward_metadata
    q3 = w.q.view(B, c.num_q_heads, c.head_dim)
    o3 = torch.empty((B, c.num_q_heads, c.head_dim), dtype=w.q.dtype, device=dev)
    pool = runner.token_to_kv_pool
    fp8 = c.kv_dtype == "fp8_e4m3"
    dargs = (q3, pool.k_buffer[0], pool.v_buffer[0], o3, None, backend.plan_buf, md.indices, backend.workspace,
             float(layer.scaling), 0.0, c.k_scale if fp8 else 1.0, c.v_scale if fp8 else 1.0, 1)
    for _ in range(3):
        torch.ops.b200attn.decode(*dargs)
    kgraph = torch.cuda.CUDAGraph()
    torch.cuda.synchronize()
    with torch.cuda.graph(kgraph):
        torch.ops.b200attn.decode(*dargs)
    for _ in range(3):
        kgraph.replay()
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = max(10, min(args.steps, 50))
    torch.cuda.synchronize()
    k0.record()
    for _ in range(reps):
        kgraph.replay()
    k1.record()
    torch.cuda.synchronize()
    k_ms = k0.elapsed_time(k1) / reps
    e_kv = 1 if fp8 else 2
    # SURVEY.md 8(d) per-unit figure for the decode kernel: KV rows + q + o + index words
    decode_bytes = (w.seq_lens_sum * c.num_kv_heads * 2 * c.head_dim * e_kv
                    + 2 * B * c.num_q_heads * c.head_dim * 2 + 4 * w.seq_lens_sum)
    peaks, peak_src = _measured_peak()
    achieved = decode_bytes / (k_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peaks, "unit": "GB/s", "frac": achieved / peaks,
                "traffic": _ncu_traffic(args.config, world), "kernel": "decode_kernel (K2, phases=1)",
                "kernel_ms": k_ms, "algorithmic_bytes": decode_bytes, "peak_source": peak_src,
                "frac_of_nominal_8TBs": achieved / 8000.0}

    # ---------------- end to end with host buffers: `e2e` ----------------
    host_qkv = torch.empty(w.qkv.shape, dtype=w.qkv.dtype).pin_memory()
    host_qkv.copy_(w.qkv.cpu())
    host_rpi = w.req_pool_indices.cpu().pin_memory()
    host_sl = w.seq_lens.cpu().pin_memory()
    host_loc = w.out_cache_loc.cpu().pin_memory()
    host_out = torch.empty((B, c.num_q_heads * c.head_dim), dtype=w.q.dtype).pin_memory()
    d_qkv = torch.empty_like(w.qkv)
    d_rpi, d_sl, d_loc = (torch.empty_like(t) for t in (w.req_pool_indices, w.seq_lens, w.out_cache_loc))
    qs, ks, vs = d_qkv.split([c.num_q_heads * c.head_dim, c.num_kv_heads * c.head_dim,
                              c.num_kv_heads * c.head_dim], dim=-1)
    fb2 = make_batch(w, runner, backend)
    fb2.req_pool_indices, fb2.seq_lens, fb2.out_cache_loc = d_rpi, d_sl, d_loc
    h2d = sum(t.numel() * t.element_size() for t in (host_qkv, host_rpi, host_sl, host_loc))
    d2h = host_out.numel() * host_out.element_size()

    def e2e_step():
        d_qkv.copy_(host_qkv, non_blocking=True)
        d_rpi.copy_(host_rpi, non_blocking=True)
        d_sl.copy_(host_sl, non_blocking=True)
        d_loc.copy_(host_loc, non_blocking=True)
        backend.init_forward_metadata(fb2)
        o = layer(qs, ks, vs, fb2)
        host_out.copy_(o, non_blocking=True)
        torch.cuda.current_stream().synchronize()   # the caller reads the result on the host

    for _ in range(3):
        e2e_step()
    barrier()
    launches_a = capi.launch_count()
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    barrier()
    e2e_launches = capi.launch_count() - launches_a
    e2e_ms = torch.tensor([e0.elapsed_time(e1) / args.steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_ms = float(e2e_ms.item())

    if rank == 0:
        total_bytes = w.algorithmic_bytes() * world
        line = {
            "metric": METRIC, "value": B / (ms_max * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_max, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16" if not fp8 else "fp8_e4m3 KV / bf16 q",
            "data": "synthetic", "config": workload_config(cfg, world),
            "hbm_gb_per_s": total_bytes / (ms_max * 1e-3) / 1e9,
            "tokens_per_s_model_equiv": B / (ms_max * 1e-3) / cfg.layers_equiv,
            "layers_equiv": cfg.layers_equiv,
            "launch_mode": "cuda_graph" if graph is not None else "eager",
            "roofline": roofline,
            "e2e": {"value": B / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms, "launches": e2e_launches},
            "gpu_launches": launches_per_step * args.steps,
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            tps, cms, info = cpu_reference_run(cfg, args.cpu_sample_requests, 5, 2)
            line["cpu_baseline"] = {"value": tps, "unit": UNIT, "cores": info["cores"], "kind": info["kind"],
                                    "sample": info["sample"], "ms_per_step": cms, "gb_per_s": info["gb_per_s"]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _rank_workload(synth, rcfg, cfg, dev):
    """Head-sliced workload: request tables from the shared seed, pool contents per rank."""
    w = synth.make_workload(replace(rcfg, seed=cfg.seed), dev)
    return w


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
