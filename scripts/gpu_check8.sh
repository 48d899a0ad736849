# ranges + dynamic tickets + fused merge: parity, then bench on every config
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu8.log 2>&1; tail -15 gpurun_out/t_gpu8.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
for c in C2 C1 C3 C4 C5; do
timeout 300 python bench.py --steps 20 --warmup 3 --config $c --no-cpu-baseline > gpurun_out/b8_$c.json 2> gpurun_out/b8_$c.err || tail -3 gpurun_out/b8_$c.err
done
for d in -1 10 35; do
B200ATTN_DYN_PCT=$d timeout 300 python bench.py --steps 20 --warmup 3 --config C2 --no-cpu-baseline > gpurun_out/b8_C2_dyn$d.json 2> gpurun_out/b8_C2_dyn$d.err
done
B200ATTN_DYN_BYTES=262144 timeout 300 python bench.py --steps 20 --warmup 3 --config C2 --no-cpu-baseline > gpurun_out/b8_C2_dynb256k.json 2> gpurun_out/b8_C2_dynb256k.err
B200ATTN_DYN_BYTES=1048576 timeout 300 python bench.py --steps 20 --warmup 3 --config C2 --no-cpu-baseline > gpurun_out/b8_C2_dynb1m.json 2> gpurun_out/b8_C2_dynb1m.err
for t in 4 8; do
timeout 300 python bench.py --steps 20 --warmup 3 --tp-sim $t --scaling strong --no-cpu-baseline > gpurun_out/b8_C2_tpsim$t.json 2> gpurun_out/b8_C2_tpsim$t.err
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b8_*.json')):
    for l in open(f):
        if not l.startswith('{'): continue
        d=json.loads(l); r=d['roofline']; e=d['e2e']
        print(f, 'value %.0f ms %.4f | K2 %.4f ms %.0f GB/s frac %.3f | e2e %.0f %s'%(d['value'],d['ms_per_step'],r['kernel_ms'],r['achieved'],r['frac'],e['value'],e['variants_ms_per_step']))
P
