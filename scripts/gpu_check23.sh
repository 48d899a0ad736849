mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extend.py -x -q --timeout 120 > gpurun_out/t_ext23.log 2>&1; tail -3 gpurun_out/t_ext23.log
B200ATTN_EXTEND_IMPL=mma timeout 900 python -m pytest tests/test_gpu_extend.py -x -q --timeout 120 > gpurun_out/t_ext23_mma.log 2>&1; tail -3 gpurun_out/t_ext23_mma.log
timeout 400 python bench_extend.py --steps 10 --warmup 3 > gpurun_out/bext23.json 2> gpurun_out/bext23.err; tail -3 gpurun_out/bext23.err
B200ATTN_EXTEND_IMPL=mma timeout 400 python bench_extend.py --config P1d64 --steps 10 --warmup 3 > gpurun_out/bext23_mma.json 2> gpurun_out/bext23_mma.err; tail -3 gpurun_out/bext23_mma.err
python - <<'P'
import json
for f in ('gpurun_out/bext23.json','gpurun_out/bext23_mma.json'):
  for l in open(f):
    if not l.startswith('{'): continue
    d=json.loads(l); r=d['roofline']
    print(d['config']['workload'][:40], 'ms %.3f kernel %.3f | %.0f TFLOP/s %.0f GB/s frac %.3f (%s)'%(d['ms_per_step'], r['kernel_ms'], r['tflops'], r['gb_per_s'], r['frac'], r['bound']))
P
