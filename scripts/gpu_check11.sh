set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extend.py -x -q > gpurun_out/t_ext11.log 2>&1; tail -15 gpurun_out/t_ext11.log
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/t_gpu11.log 2>&1; tail -5 gpurun_out/t_gpu11.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke11.log 2>&1; tail -2 gpurun_out/smoke11.log
timeout 600 python bench_extend.py --steps 10 --warmup 3 --sdpa > gpurun_out/bext11.json 2> gpurun_out/bext11.err; tail -3 gpurun_out/bext11.err
python - <<'P'
import json
for l in open('gpurun_out/bext11.json'):
    if not l.startswith('{'): continue
    d=json.loads(l); r=d['roofline']
    print(d['config']['workload'][:40], 'ms %.3f kernel %.3f | %.0f TFLOP/s %.0f GB/s frac %.3f (%s) | sdpa %s'%(d['ms_per_step'], r['kernel_ms'], r['tflops'], r['gb_per_s'], r['frac'], r['bound'], d.get('sdpa_contiguous')))
P
