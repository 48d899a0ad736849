"""Event log of CTA 0 of extend_tc_smallq_kernel (persistent small-q extend kernel) on the TARGET_VERIFY-like
shape of bench_extend.py P3 (tuning aid; needs the -DB200ATTN_TC_TRACE build, see tools/trace_tc.py).
usage: trace_tc_smallq.py [prefix=4096] [extend=8] [batch=128] [kv=bf16|fp8_e4m3]
Tags: softmax 1 item start, 2 saw s_full, 3 P published, 4 epilogue start, 5 both sub-tiles done, 6 saw o_full,
7 item done | loader 10 lagged publish done, 11 got the empty slot, 12 stage issued, 13 got the Q buffer |
MMA 20 item start, 21 Q and K0 landed, 22 V landed, 23 P ready (P.V issued next), 25 o_full committed."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ltp-sglang_b200")]
import numpy as np, torch
import b200attn, synth

dev = torch.device("cuda:0")
Lp = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
Le = int(sys.argv[2]) if len(sys.argv) > 2 else 8
B = int(sys.argv[3]) if len(sys.argv) > 3 else 128
kvd = sys.argv[4] if len(sys.argv) > 4 else "bf16"
w = synth.make_extend_workload(synth.Config("t", B, 0, 32, 8, 128, seed=7, kv_dtype=kvd), [Lp] * B, [Le] * B, dev)
T = B * Le
q = w.q.view(T, 32, 128); o = torch.empty_like(q)
b200attn.store_kv(w.k_pool, w.v_pool, w.out_cache_loc, w.k.view(T, 8, 128), w.v.view(T, 8, 128), 1.0, 1.0)
qo = torch.arange(0, T + 1, Le, dtype=torch.int32, device=dev)
kv = torch.arange(0, B * Lp + 1, Lp, dtype=torch.int32, device=dev)
lse = torch.zeros(max(T * 32, 4 * 2048 * 2), dtype=torch.float32, device=dev)
for _ in range(2):
    lse.zero_()
    torch.ops.b200attn.extend(q, o, lse, w.k_pool, w.v_pool, qo, kv, None, w.req_to_token, w.req_pool_indices,
                              w.extend_prefix_lens, w.out_cache_loc, None, None, Le, True, -1, True, w.sm_scale, 0.0)
torch.cuda.synchronize()
t = lse.view(torch.int64)[: 4 * 2048].cpu().numpy().reshape(4, 2048)
ev = []
for role in range(4):
    for x in t[role]:
        if x == 0: break
        ev.append((int(x) & ((1 << 48) - 1), role, int(x) >> 48))
t0 = min(e[0] for e in ev)
names = ["sm0", "sm1", "ld ", "mma"]
per = {r: [(c - t0, tag) for c, rr, tag in ev if rr == r] for r in range(4)}
for r in range(4):
    print(names[r], "events", len(per[r]), "last", per[r][-1][0] if per[r] else None)
# item boundaries seen by sub-tile 0: tags 1 .. 7
items = [c for c, tag in per[0] if tag == 1]
ends = [c for c, tag in per[0] if tag == 7]
print("item starts (sm0):", items[:8])
print("item ends   (sm0):", ends[:8])
for r in range(4):
    print("----", names[r])
    last = None
    for c, tag in per[r][: (150 if r != 2 else 260)]:
        print(f"{c:9d} +{(c - last) if last is not None else 0:6d} tag {tag}")
        last = c
