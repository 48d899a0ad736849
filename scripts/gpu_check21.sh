mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extend.py -x -q --timeout 121 > gpurun_out/t_ext21.log 2>&1; tail -3 gpurun_out/t_ext21.log
timeout 400 python bench_extend.py --steps 10 --warmup 3 > gpurun_out/bext21.json 2> gpurun_out/bext21.err; tail -3 gpurun_out/bext21.err
python - <<'P'
import json
for f in ('gpurun_out/bext21.json',):
  for l in open(f):
    if not l.startswith('{'): continue
    d=json.loads(l); r=d['roofline']
    print(d['config']['workload'][:40], 'ms %.3f kernel %.3f | %.0f TFLOP/s %.0f GB/s frac %.3f (%s)'%(d['ms_per_step'], r['kernel_ms'], r['tflops'], r['gb_per_s'], r['frac'], r['bound']))
P
export B210ATTN_LIB=$PWD/tools/bin/libb210attn_trace.so
timeout 121 python tools/trace_tc_verify.py > gpurun_out/trace_verify.txt 2>&1
timeout 121 python tools/trace_tc.py > gpurun_out/trace_p1.txt 2>&1
grep -E "^CTA|n=  0|n= 3[12]|^st=  0" gpurun_out/trace_verify.txt; tail -12 gpurun_out/trace_p1.txt
