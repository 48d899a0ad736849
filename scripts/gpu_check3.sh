# 1-GPU: parity tests, smoke, bench on every config, reference arm, ncu launch list of the default bench.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu.log 2>&1; tail -5 gpurun_out/t_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
python bench.py --steps 30 --warmup 5 > gpurun_out/b3_C2.json 2> gpurun_out/b3_C2.err; tail -3 gpurun_out/b3_C2.err
for c in C1 C3 C4 C5; do
python bench.py --steps 20 --warmup 3 --config $c --no-cpu-baseline > gpurun_out/b3_$c.json 2> gpurun_out/b3_$c.err; tail -3 gpurun_out/b3_$c.err
done
python bench.py --steps 20 --warmup 3 --config C2 --page-size 16 --no-cpu-baseline > gpurun_out/b3_C2p16.json 2> gpurun_out/b3_C2p16.err
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/bref.json 2> gpurun_out/bref.err; tail -3 gpurun_out/bref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_C2.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b3_*.json')):
    for l in open(f):
        if not l.startswith('{'): continue
        d=json.loads(l); r=d['roofline']; e=d['e2e']
        print(f, 'value %.0f ms %.4f | K2 %.4f ms %.0f GB/s frac %.3f | e2e %.0f %s | clk %s'%(d['value'],d['ms_per_step'],r['kernel_ms'],r['achieved'],r['frac'],e['value'],e['variants_ms_per_step'],d['clocks']))
P
nproc; lscpu | grep -i 'model name'
