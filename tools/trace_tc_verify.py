"""Pipeline timeline of CTA (0,0,0) of extend_tc_kernel in its small-q (KV ping-pong) mode, on the
TARGET_VERIFY-like shape of bench_extend.py P3 (tuning aid; needs the -DB200ATTN_TC_TRACE build, see
tools/trace_tc.py).  usage: trace_tc_verify.py [prefix=4096] [extend=8] [batch=128] [kv=bf16|fp8_e4m3]"""
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
lse = torch.zeros(max(T * 32, 8192 + 8 * (B * 8 + 8)), dtype=torch.float32, device=dev)
for _ in range(2):
    lse.zero_()
    torch.ops.b200attn.extend(q, o, lse, w.k_pool, w.v_pool, qo, kv, None, w.req_to_token, w.req_pool_indices,
                              w.extend_prefix_lens, w.out_cache_loc, None, None, Le, True, -1, True, w.sm_scale, 0.0)
torch.cuda.synchronize()
t = lse.view(torch.int64)[: 14 * 256].cpu().numpy().reshape(14, 256)
N = (Lp + Le + 127) // 128
t0 = t[0, 0]
print("tiles", N)
print("CTA entry %d exit %d (clk rel. to first s_full)" % (t[12, 255] - t0, t[13, 255] - t0))
print("prologue: roles start %d, Q issued %d, first slot ids %d | tail: loader done %d, last MMA committed %d, O complete %d, "
      "outputs stored %d" % tuple(int(x - t0) for x in (t[12, 250], t[12, 251], t[12, 252], t[13, 252], t[13, 253], t[13, 250], t[13, 251])))
for n in range(0, 64):
    if n >= N: break
    v = n & 1
    r = [int(t[k, n] - t0) if t[k, n] else -1 for k in (v * 3, v * 3 + 1, v * 3 + 2, 6, 7, 8)]
    print(f"n={n:3d} sub{v}: s {r[0]:7d} ld {r[1]:7d} p {r[2]:7d} | mma: Vfull {r[3]:7d} pfull {r[4]:7d} K(n+2)full {r[5]:7d}")
for st in list(range(0, 10)) + list(range(40, 48)):
    if st >= 2 * N: break
    print(f"st={st:3d} loader: iter start {int(t[13, st]-t0):7d} landed(st-2) +{int(t[11, st-2]-t[13, st]) if st >= 2 else 0:5d} "
          f"empty +{int(t[10, st]-(t[11, st-2] if st >= 2 else t[13, st])):5d} issue +{int(t[12, st]-t[10, st]):5d}")
pub = t[11, 8:2 * N - 4].astype(float)
print("steady clk per stage (loader publish): %.0f" % np.diff(pub).mean())

# every CTA's entry / exit (globaltimer, ns) and SM
c = lse.view(torch.int64)[4096: 4096 + 4 * B * 8].cpu().numpy().reshape(B * 8, 4)
st, en, sm = c[:, 0] - c[:, 0].min(), c[:, 1] - c[:, 0].min(), c[:, 2]
dur = en - st
print("CTAs %d  kernel span %.1f us  CTA duration us: mean %.1f min %.1f max %.1f" % (len(c), en.max() / 1e3, dur.mean() / 1e3, dur.min() / 1e3, dur.max() / 1e3))
order = np.argsort(st)
for k in (0, 1, 147, 148, 149, 300, 600, 900, len(c) - 1):
    i = order[min(k, len(c) - 1)]
    print("  launch order %4d: cta %4d sm %3d start %.1f us dur %.1f us" % (k, i, sm[i], st[i] / 1e3, dur[i] / 1e3))
gaps = []
for m in np.unique(sm):
    idx = np.where(sm == m)[0]
    idx = idx[np.argsort(st[idx])]
    gaps += list(st[idx][1:] - en[idx][:-1])
print("CTAs per SM: min %d max %d; gap between consecutive CTAs on an SM us: mean %.2f max %.2f" % (
    min(np.sum(sm == m) for m in np.unique(sm)), max(np.sum(sm == m) for m in np.unique(sm)), np.mean(gaps) / 1e3, np.max(gaps) / 1e3))
