"""Single-request repro for the fp8 decode gap against flashinfer that round 1 recorded (0.10 max-abs on C3,
profiles/r01_competitors_flashinfer.jsonl): this repo's decode, flashinfer's BatchDecodeWithPagedKVCacheWrapper and
the float64 oracle on the SAME fp8-e4m3 pool, no KV store in between (the round-1 tool compared outputs taken
before / after two dozen in-place `k.div_(k_scale)` stores of the new token: the pools were not the same).
usage: python tools/fp8_vs_flashinfer.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ltp-sglang_b200")]
import numpy as np
import torch
import b200attn, synth
from bench_harness import make_batch, make_layer, make_runner
from oracle import decode_oracle as oracle

dev = torch.device("cuda:0")
for ks, vs in ((1.0, 1.0), (0.5, 2.0)):
    cfg = synth.Config("fp8gap", 4, 0, 32, 8, 128, kv_dtype="fp8_e4m3", ragged=(1500, 2500), k_scale=ks, v_scale=vs, seed=3)
    w_cpu = synth.make_workload(cfg, "cpu")
    w = synth.make_workload(cfg, dev)
    w.qkv.copy_(w_cpu.qkv); w.k_pool.copy_(w_cpu.k_pool); w.v_pool.copy_(w_cpu.v_pool)
    runner = make_runner(w, dev)
    backend = b200attn.B200AttnBackend(runner, materialize_kv_indices=True)
    layer = make_layer(w)
    fb = make_batch(w, runner, backend)
    backend.init_forward_metadata(fb)
    B, Hq, Hkv, D = cfg.batch, cfg.num_q_heads, cfg.num_kv_heads, cfg.head_dim
    ours = layer(w.q, w.k, w.v, fb, save_kv_cache=False).view(B, Hq, D).float().cpu().numpy()
    ref, _ = oracle.decode_attention(w_cpu.q.reshape(B, Hq, D), w_cpu.k_pool, w_cpu.v_pool, w_cpu.req_to_token,
                                     w_cpu.req_pool_indices, w_cpu.seq_lens, w_cpu.sm_scale, k_scale=ks, v_scale=vs)
    rec = {"k_scale": ks, "v_scale": vs, "b200attn_vs_oracle_max_abs": float(np.abs(ours - ref).max()),
           "oracle_abs_max": float(np.abs(ref).max())}
    try:
        import flashinfer
        md = backend.forward_metadata
        wsb = torch.empty(128 << 20, dtype=torch.uint8, device=dev)
        for tc in (True, False):
            fw = flashinfer.BatchDecodeWithPagedKVCacheWrapper(wsb, "NHD", use_tensor_cores=tc)
            fw.plan(md.kv_indptr[: B + 1], md.kv_indices, torch.ones(B, dtype=torch.int32, device=dev), Hq, Hkv, D, 1,
                    q_data_type=w.q.dtype, kv_data_type=torch.float8_e4m3fn, sm_scale=w.sm_scale)
            kc = w.k_pool.unsqueeze(1).view(torch.float8_e4m3fn)
            vc = w.v_pool.unsqueeze(1).view(torch.float8_e4m3fn)
            o = fw.run(w.q.view(B, Hq, D).contiguous(), (kc, vc), k_scale=float(ks), v_scale=float(vs)).float().cpu().numpy()
            rec[f"flashinfer_tc{int(tc)}_vs_oracle_max_abs"] = float(np.abs(o - ref).max())
            rec[f"flashinfer_tc{int(tc)}_vs_b200attn_max_abs"] = float(np.abs(o - ours).max())
        rec["flashinfer_version"] = flashinfer.__version__
    except Exception as ex:  # noqa: BLE001
        rec["flashinfer"] = repr(ex)[:200]
    print(json.dumps(rec), flush=True)
