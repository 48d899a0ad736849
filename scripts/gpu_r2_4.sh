# round 2, call 4: tests on weighted ranges + last-arriver shortcut; traces; bench A/B of the balancing
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -x > gpurun_out/r2_t4.log 2>&1; tail -n 8 gpurun_out/r2_t4.log
(
export B200ATTN_LIB=$PWD/ltp-sglang_b200/b200attn/libb200attn_trace.so
for i in 1 2; do
timeout 300 python tools/trace_decode.py --config C2 --repeat 6 --out gpurun_out/r2d_trace_C2.json > /dev/null 2> gpurun_out/r2d_trace.err
done
timeout 300 python tools/trace_decode.py --config C2 --tp-sim 8 --repeat 6 --out gpurun_out/r2d_trace_C2_tp8_b128.json > /dev/null 2>> gpurun_out/r2d_trace.err
timeout 300 python tools/trace_decode.py --config C1 --flush --repeat 3 --out gpurun_out/r2d_trace_C1.json > /dev/null 2>> gpurun_out/r2d_trace.err
)
tail -n 3 gpurun_out/r2d_trace.err
timeout 600 python bench.py > gpurun_out/r2d_bench_C2.json 2> gpurun_out/r2d_bench_C2.err; tail -n 2 gpurun_out/r2d_bench_C2.err
B200ATTN_NO_BALANCE=1 timeout 300 python bench.py --no-cpu-baseline --sustain 0 > gpurun_out/r2d_bench_C2_nobalance.json 2>/dev/null
timeout 300 python bench.py --config C1 --no-cpu-baseline --sustain 0 > gpurun_out/r2d_bench_C1.json 2>/dev/null
timeout 300 python bench.py --config C2 --tp-sim 8 --scaling strong --no-cpu-baseline --sustain 0 > gpurun_out/r2d_bench_C2_tp8_strong.json 2>/dev/null
B200ATTN_NO_BALANCE=1 timeout 300 python bench.py --config C2 --tp-sim 8 --scaling strong --no-cpu-baseline --sustain 0 > gpurun_out/r2d_bench_C2_tp8_strong_nobalance.json 2>/dev/null
timeout 300 python bench.py --config C3 --no-cpu-baseline --sustain 0 > gpurun_out/r2d_bench_C3.json 2>/dev/null
timeout 300 python bench.py --config C4 --no-cpu-baseline --sustain 0 > gpurun_out/r2d_bench_C4.json 2>/dev/null
timeout 300 python bench.py --config C5 --no-cpu-baseline --sustain 0 > gpurun_out/r2d_bench_C5.json 2>/dev/null
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2d_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r2d_bench_')[1], 'ms', round(d['ms_per_step'],4), 'K2ms', round(d['roofline']['kernel_ms'],4), 'GB/s', round(d['roofline']['achieved']), 'e2e ms', round(d['e2e']['ms_per_step'],4), d['e2e']['variants_ms_per_step'], d.get('sustained',{}).get('ms_per_step'))
    except Exception as e:
        print(f, 'ERR', e)
P
