set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_extend.py -m gpu -x -q --timeout 120 > gpurun_out/t_ext33.log 2>&1; tail -3 gpurun_out/t_ext33.log
for c in P3 P3f; do
timeout 300 python bench_extend.py --steps 20 --warmup 3 --config $c > gpurun_out/bext33_$c.json 2>gpurun_out/bext33_$c.err
done
python - <<'P'
import json
for c in ('P3','P3f'):
    d=json.loads(open(f'gpurun_out/bext33_{c}.json').read().strip().splitlines()[-1]); r=d['roofline']; print(c, r['kernel_ms'], round(r['gb_per_s']), round(r['tflops']))
P
B200ATTN_LIB=$PWD/tools/bin/libb200attn_trace.so timeout 120 python tools/trace_tc_smallq.py > gpurun_out/trace_smallq2.txt 2>&1; head -7 gpurun_out/trace_smallq2.txt
