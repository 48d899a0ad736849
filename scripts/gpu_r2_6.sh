# round 2, call 6: hot-loop fix (rate stamp / early check out of the chunk loop), first-chunk ids in the plan, balanced ranges engaged in the bench warm-up
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -x > gpurun_out/r2_t6.log 2>&1; tail -n 4 gpurun_out/r2_t6.log
for cfgs in "C1:--config C1" "C2:--config C2" "C3:--config C3" "C2tp8s:--config C2 --tp-sim 8 --scaling strong" "C3tp8:--config C3 --tp-sim 8" "C4:--config C4"; do
  name=${cfgs%%:*}; a=${cfgs#*:}
  timeout 300 python bench.py --no-cpu-baseline --sustain 0 $a > gpurun_out/r2f_bal_$name.json 2> gpurun_out/r2f_bal_$name.err
  B200ATTN_NO_BALANCE=1 timeout 300 python bench.py --no-cpu-baseline --sustain 0 $a > gpurun_out/r2f_nobal_$name.json 2> gpurun_out/r2f_nobal_$name.err
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2f_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); r=d['roofline']
        print(f.split('/')[-1][4:-5].ljust(20), 'step ms', round(d['ms_per_step'],4), 'K2 ms', round(r.get('kernel_ms',0),4), 'GB/s', round(r['achieved']), 'frac', round(r['frac'],3), 'e2e ms', round(d['e2e'].get('ms_per_step',0),4))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-400:])
P
export B200ATTN_LIB=$PWD/ltp-sglang_b200/b200attn/libb200attn_trace.so
for c in C2 C3; do
timeout 300 python tools/trace_decode.py --config $c --repeat 8 --out gpurun_out/r2f_trace_$c.json > /dev/null 2>> gpurun_out/r2f_trace.err
done
timeout 300 python tools/trace_decode.py --config C1 --flush --repeat 8 --out gpurun_out/r2f_trace_C1.json > /dev/null 2>> gpurun_out/r2f_trace.err
timeout 300 python tools/trace_decode.py --config C2 --tp-sim 8 --repeat 8 --out gpurun_out/r2f_trace_C2_tp8_b128.json > /dev/null 2>> gpurun_out/r2f_trace.err
tail -n 3 gpurun_out/r2f_trace.err
