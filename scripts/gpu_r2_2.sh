# round 2, call 2: tests on the reworked decode epilogue / range descriptors / PDL, timelines, bench + dynamic-tail A/B
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -x > gpurun_out/r2_t2.log 2>&1; tail -n 8 gpurun_out/r2_t2.log
(
export B200ATTN_LIB=$PWD/ltp-sglang_b200/b200attn/libb200attn_trace.so
timeout 300 python tools/trace_decode.py --config C2 --out gpurun_out/r2b_trace_C2.json > /dev/null 2> gpurun_out/r2b_trace.err
timeout 300 python tools/trace_decode.py --config C2 --tp-sim 8 --out gpurun_out/r2b_trace_C2_tp8_b128.json > /dev/null 2>> gpurun_out/r2b_trace.err
timeout 300 python tools/trace_decode.py --config C1 --flush --out gpurun_out/r2b_trace_C1.json > /dev/null 2>> gpurun_out/r2b_trace.err
)
tail -n 3 gpurun_out/r2b_trace.err
timeout 600 python bench.py > gpurun_out/r2b_bench_C2.json 2> gpurun_out/r2b_bench_C2.err; tail -n 2 gpurun_out/r2b_bench_C2.err
for pct in 8 15; do for kb in 256 512; do
B200ATTN_DYN_PCT=$pct B200ATTN_DYN_BYTES=$((kb*1024)) timeout 300 python bench.py --no-cpu-baseline --sustain 0 --steps 30 > gpurun_out/r2b_bench_C2_dyn${pct}_${kb}k.json 2>/dev/null
done; done
B200ATTN_NO_PDL=1 timeout 300 python bench.py --no-cpu-baseline --sustain 0 --steps 30 > gpurun_out/r2b_bench_C2_nopdl.json 2>/dev/null
timeout 300 python bench.py --config C1 --no-cpu-baseline --sustain 0 > gpurun_out/r2b_bench_C1.json 2>/dev/null
timeout 300 python bench.py --config C2 --tp-sim 8 --scaling strong --no-cpu-baseline --sustain 0 > gpurun_out/r2b_bench_C2_tp8_strong.json 2>/dev/null
timeout 300 python bench.py --config C3 --tp-sim 8 --no-cpu-baseline --sustain 0 > gpurun_out/r2b_bench_C3_tp8_p1.json 2>/dev/null
timeout 300 python bench.py --config C3 --tp-sim 8 --page-size 16 --no-cpu-baseline --sustain 0 > gpurun_out/r2b_bench_C3_tp8_p16.json 2>/dev/null
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2b_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r2b_bench_')[1], 'ms', round(d['ms_per_step'],4), 'K2ms', round(d['roofline']['kernel_ms'],4), 'GB/s', round(d['roofline']['achieved']), 'e2e ms', round(d['e2e']['ms_per_step'],4), d['e2e']['variants_ms_per_step'])
    except Exception as e:
        print(f, 'ERR', e)
P
