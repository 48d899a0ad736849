set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu10.log 2>&1; tail -8 gpurun_out/t_gpu10.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
for c in C2 C1 C3 C4 C5; do
timeout 300 python bench.py --steps 20 --warmup 3 --config $c --no-cpu-baseline > gpurun_out/b10_$c.json 2> gpurun_out/b10_$c.err || tail -3 gpurun_out/b10_$c.err
done
for t in 2 4 8; do
timeout 300 python bench.py --steps 20 --warmup 3 --tp-sim $t --scaling strong --no-cpu-baseline > gpurun_out/b10_C2_tpsim$t.json 2> gpurun_out/b10_C2_tpsim$t.err
done
timeout 300 python bench.py --steps 20 --warmup 3 --tp-sim 8 --config C3 --scaling strong --no-cpu-baseline > gpurun_out/b10_C3_tpsim8.json 2> gpurun_out/b10_C3_tpsim8.err
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b10_*.json')):
    for l in open(f):
        if not l.startswith('{'): continue
        d=json.loads(l); r=d['roofline']; e=d['e2e']
        print(f, 'value %.0f ms %.4f | K2 %.4f ms %.0f GB/s frac %.3f | e2e %.0f %s'%(d['value'],d['ms_per_step'],r['kernel_ms'],r['achieved'],r['frac'],e['value'],e['variants_ms_per_step']))
P
