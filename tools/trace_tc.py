"""Pipeline timeline of one CTA of extend_tc_kernel (tuning aid; needs the -DB200ATTN_TC_TRACE build):
    make -C ltp-sglang_b200/csrc EXTRA=-DB200ATTN_TC_TRACE BUILD=/tmp/trace_build OUT=$PWD/tools/bin/libb200attn_trace.so
    B200ATTN_LIB=tools/bin/libb200attn_trace.so python tools/trace_tc.py
Kinds: 0/3 softmax0/1 saw s_full, 1/4 S row loaded, 2/5 P published, 6 MMA saw V full, 7 saw p_full0,
8 saw K(n+1) full, 9 saw p_full1, 10 loader got the empty slot (per stage), 11 loader published stage."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ltp-sglang_b200")]
import numpy as np, torch
import b200attn, synth

dev = torch.device("cuda:0")
L = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
w = synth.make_extend_workload(synth.Config("t", 1, 0, 32, 8, 128, seed=7), [0], [L], dev)
q = w.q.view(L, 32, 128); o = torch.empty_like(q)
b200attn.store_kv(w.k_pool, w.v_pool, w.out_cache_loc, w.k.view(L, 8, 128), w.v.view(L, 8, 128))
qo = torch.tensor([0, L], dtype=torch.int32, device=dev); kv = torch.zeros(2, dtype=torch.int32, device=dev)
lse = torch.zeros(L * 32, dtype=torch.float32, device=dev)
for _ in range(2):
    lse.zero_()
    torch.ops.b200attn.extend(q, o, lse, w.k_pool, w.v_pool, qo, kv, None, w.req_to_token, w.req_pool_indices,
                              w.extend_prefix_lens, w.out_cache_loc, None, None, L, True, -1, True, w.sm_scale, 0.0)
torch.cuda.synchronize()
t = lse.view(torch.int64)[: 16 * 256].cpu().numpy().reshape(16, 256)
N = (L + 127) // 128
t0 = t[0, 0]
np.set_printoptions(linewidth=200)
print("tiles", N)
for n in list(range(0, 3)) + list(range(30, 34)):
    if n >= N: break
    r = [int(t[k, n] - t0) if t[k, n] else -1 for k in range(10)]
    print(f"n={n:3d} sm0: s {r[0]:7d} ld {r[1]:7d} p {r[2]:7d} | sm1: s {r[3]:7d} ld {r[4]:7d} p {r[5]:7d} | mma: V {r[6]:7d} p0 {r[7]:7d} K {r[8]:7d} p1 {r[9]:7d}")
for st in list(range(0, 8)) + list(range(60, 66)):
    print(f"st={st:3d} loader: iter start {int(t[13, st]-t0):7d} landed(st-2) +{int(t[11, st-2]-t[13, st]) if st >= 2 else 0:5d} "
          f"empty +{int(t[10, st]-(t[11, st-2] if st >= 2 else t[13, st])):5d} issue +{int(t[12, st]-t[10, st]):5d}")
d = np.diff(t[0, 8:N-2]).astype(float)
print("steady period per tile (clk): mean %.0f" % d.mean(), " softmax0 busy (s->p): %.0f" % (t[2, 8:N-2] - t[0, 8:N-2]).mean(),
      " ld: %.0f" % (t[1, 8:N-2] - t[0, 8:N-2]).mean(), " p0->next s0: %.0f" % (t[0, 9:N-1] - t[2, 8:N-2]).mean())

if t[14, 8] and t[15, 8]:
    # single-pass body: kind 14 = row max exchanged (after the named barrier), 15 = exps + P stores issued (before wait::st)
    a = t[:, 8:N-2].astype(float)
    print("softmax0 per tile: s_full -> max %.0f | exchange + barrier %.0f | rescale vote + exps + st issue %.0f | wait::st + fence + arrive %.0f" % (
        (a[1]-a[0]).mean(), (a[14]-a[1]).mean(), (a[15]-a[14]).mean(), (a[2]-a[15]).mean()))
    print("mma: p0 published -> seen %.0f | PV0 issue (+ K wait) %.0f | p1 published -> seen %.0f | s1 commit -> softmax1 sees s_full(n+1): %.0f after loop top" % (
        (a[7]-a[2]).mean(), (a[8]-a[7]).mean(), (a[9]-a[5]).mean(), (t[3, 9:N-1]-t[6, 9:N-1]).mean()))

tw = lse.view(torch.int64)[32 * 256: 64 * 256].cpu().numpy().reshape(16, 2, 256)
if tw[0, 0, 8]:
    print("per-warp (relative to warp 0's s_full of the tile): s_full seen / P published")
    for n in (30, 31):
        base = tw[0, 0, n]
        print(f" n={n}: " + "  ".join(f"w{w}:{int(tw[w,0,n]-base)}/{int(tw[w,1,n]-base)}" for w in range(16)))
    busy = (tw[:, 1, 8:N-2] - tw[:, 0, 8:N-2]).mean(axis=1)
    print(" mean busy per warp:", " ".join(f"{b:.0f}" for b in busy))
    first = tw[:8, 1, 8:N-2].min(axis=0); last = tw[:8, 1, 8:N-2].max(axis=0)
    print(" sub-tile 0: first warp's P -> last warp's P: mean %.0f clk; s_full skew %.0f" % ((last - first).mean(), (tw[:8, 0, 8:N-2].max(axis=0) - tw[:8, 0, 8:N-2].min(axis=0)).mean()))
