mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extend.py -x -q --timeout 120 > gpurun_out/t_ext22.log 2>&1; tail -3 gpurun_out/t_ext22.log
timeout 400 python bench_extend.py --steps 10 --warmup 3 > gpurun_out/bext22.json 2> gpurun_out/bext22.err; tail -3 gpurun_out/bext22.err
python - <<'P'
import json
for l in open('gpurun_out/bext22.json'):
    if not l.startswith('{'): continue
    d=json.loads(l); r=d['roofline']
    print(d['config']['workload'][:40], 'ms %.3f kernel %.3f | %.0f TFLOP/s %.0f GB/s frac %.3f (%s)'%(d['ms_per_step'], r['kernel_ms'], r['tflops'], r['gb_per_s'], r['frac'], r['bound']))
P
export B200ATTN_LIB=$PWD/tools/bin/libb200attn_trace.so
timeout 120 python tools/trace_tc_verify.py > gpurun_out/trace_verify.txt 2>&1
timeout 120 python tools/trace_tc.py > gpurun_out/trace_p1.txt 2>&1
head -4 gpurun_out/trace_verify.txt; grep -E "n= 2[0-3]|^CTAs" gpurun_out/trace_verify.txt; tail -9 gpurun_out/trace_p1.txt
