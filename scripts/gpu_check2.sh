# 1-GPU: smoke + fp8 bench + tp-sim dry runs (per-rank shapes of 2/4/8-way sharding)
set -x
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/b2_C2.json 2> gpurun_out/b2_C2.err; tail -3 gpurun_out/b2_C2.err
python bench.py --steps 20 --warmup 3 --config C3 --no-cpu-baseline > gpurun_out/b2_C3.json 2> gpurun_out/b2_C3.err; tail -3 gpurun_out/b2_C3.err
python bench.py --steps 20 --warmup 3 --config C1 --no-cpu-baseline > gpurun_out/b2_C1.json 2> gpurun_out/b2_C1.err; tail -3 gpurun_out/b2_C1.err
for t in 2 4 8; do
python bench.py --steps 20 --warmup 3 --tp-sim $t --no-cpu-baseline > gpurun_out/b2_C2_tpsim$t.json 2> gpurun_out/b2_C2_tpsim$t.err; tail -3 gpurun_out/b2_C2_tpsim$t.err
python bench.py --steps 20 --warmup 3 --tp-sim $t --config C4 --scaling strong --no-cpu-baseline > gpurun_out/b2_C4_tpsim$t.json 2> gpurun_out/b2_C4_tpsim$t.err; tail -3 gpurun_out/b2_C4_tpsim$t.err
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b2_*.json')):
    for l in open(f):
        if not l.startswith('{'): continue
        d=json.loads(l); r=d['roofline']; e=d['e2e']
        print(f, 'value %.0f ms %.4f | K2 %.4f ms %.0f GB/s frac %.3f | e2e %.0f %s | clk %s'%(d['value'],d['ms_per_step'],r['kernel_ms'],r['achieved'],r['frac'],e['value'],e['variants_ms_per_step'],d['clocks']))
P
