import os, sys
sys.path[:0] = [os.getcwd(), os.path.join(os.getcwd(), "ltp-sglang_b200")]
import torch, b200attn, synth
dev = torch.device("cuda:0")
for kvd in ("bf16",):
  for Lp, Le, B in ((1536, 8, 192), (2048, 8, 192), (3072, 8, 128)):
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
    print(os.environ.get("B200ATTN_SMALLQ"), kvd, Lp, B, round(ts[len(ts)//2]*1e3,1), "us", flush=True)
