"""Kernel time of the small-q (verify / decode-like) extend path for a few prefix lengths (tuning aid).
usage: [B200ATTN_SMALLQ=oneshot] python tools/time_smallq.py [kv=bf16|fp8_e4m3]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ltp-sglang_b200")]
import torch
import b200attn, synth

dev = torch.device("cuda:0")
kvd = sys.argv[1] if len(sys.argv) > 1 else "bf16"
for Lp, Le, B in ((128, 8, 1024), (512, 8, 512), (1024, 8, 256), (4096, 8, 128), (16384, 8, 32), (1024, 1, 512), (4096, 4, 64)):
    w = synth.make_extend_workload(synth.Config("t", B, 0, 32, 8, 128, seed=7, kv_dtype=kvd), [Lp] * B, [Le] * B, dev)
    T = B * Le
    q = w.q.view(T, 32, 128); o = torch.empty_like(q)
    qo = torch.arange(0, T + 1, Le, dtype=torch.int32, device=dev)
    kv = torch.arange(0, B * Lp + 1, Lp, dtype=torch.int32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ts = []
    for i in range(8):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.ops.b200attn.extend(q, o, None, w.k_pool, w.v_pool, qo, kv, None, w.req_to_token, w.req_pool_indices,
                                  w.extend_prefix_lens, w.out_cache_loc, None, None, Le, True, -1, True, w.sm_scale, 0.0)
        e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts = sorted(ts[2:])
    eb = 1 if kvd.startswith("fp8") else 2
    byt = B * (Lp + Le) * 8 * 128 * 2 * eb
    print(f"{kvd} prefix {Lp:6d} x {Le} tokens x batch {B:5d}: {ts[len(ts)//2]*1e3:8.1f} us  {byt/ts[len(ts)//2]/1e6:7.0f} GB/s", flush=True)
