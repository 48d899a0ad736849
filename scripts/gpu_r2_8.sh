# round 2, call 8: batch-aware (narrow) geometry for small batches, in-graph kernel timing of flushed configs
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -x > gpurun_out/r2_t8.log 2>&1; tail -n 4 gpurun_out/r2_t8.log
B200ATTN_NO_NARROW=1 timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 900 -x > gpurun_out/r2_t8_nonarrow.log 2>&1; tail -n 2 gpurun_out/r2_t8_nonarrow.log
for cfgs in "C1:--config C1" "C2tp8s:--config C2 --tp-sim 8 --scaling strong" "C1b4:--config C1 --batch 4" ; do
  name=${cfgs%%:*}; a=${cfgs#*:}
  timeout 300 python bench.py --no-cpu-baseline --sustain 0 $a > gpurun_out/r2h_narrow_$name.json 2> gpurun_out/r2h_narrow_$name.err
  B200ATTN_NO_NARROW=1 timeout 300 python bench.py --no-cpu-baseline --sustain 0 $a > gpurun_out/r2h_wide_$name.json 2> gpurun_out/r2h_wide_$name.err
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2h_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); r=d['roofline']
        print(f.split('/')[-1][4:-5].ljust(20), 'step ms', round(d['ms_per_step'],4), 'K2 ms', round(r.get('kernel_ms',0),4), 'single', r.get('kernel_ms_single_launch_graph'), 'GB/s', round(r['achieved']), 'frac', round(r['frac'],3), 'e2e ms', round(d['e2e'].get('ms_per_step',0),4))
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
P
export B200ATTN_LIB=$PWD/ltp-sglang_b200/b200attn/libb200attn_trace.so
timeout 300 python tools/trace_decode.py --config C1 --flush --repeat 8 --out gpurun_out/r2h_trace_C1.json > /dev/null 2>> gpurun_out/r2h_trace.err
tail -n 3 gpurun_out/r2h_trace.err
