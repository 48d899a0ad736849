# usage: gpu_quick.sh <tag>: GPU tests, bench of the decode shapes that matter (no CPU baseline), trace of C3
TAG=$1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/${TAG}_tests.log 2>&1; tail -n 12 gpurun_out/${TAG}_tests.log
for cfgs in "C1:--config C1" "C2:--config C2" "C3:--config C3" "C4:--config C4" "C5:--config C5" "C2tp8s:--config C2 --tp-sim 8 --scaling strong" "C2tp8:--config C2 --tp-sim 8" "C3tp2:--config C3 --tp-sim 2" "C3tp4:--config C3 --tp-sim 4" "C3tp8:--config C3 --tp-sim 8"; do
  name=${cfgs%%:*}; a=${cfgs#*:}
  timeout 300 python bench.py --no-cpu-baseline --sustain 0 $a > gpurun_out/${TAG}_bench_$name.json 2> gpurun_out/${TAG}_bench_$name.err
done
python - <<P
import json,glob
for f in sorted(glob.glob('gpurun_out/${TAG}_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); r=d['roofline']
        print(f.split('_bench_')[1][:-5].ljust(10), 'step ms', round(d['ms_per_step'],4), 'K2 ms', round(r.get('kernel_ms',0),4), 'GB/s', round(r['achieved']), 'frac', round(r['frac'],3), 'e2e ms', round(d['e2e'].get('ms_per_step',0),4), 'MHz', d['clocks']['sm_mhz'], d['clocks']['reasons'], 'W', d['clocks']['power_w_max'])
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
P
bash scripts/gpu_trace.sh $TAG C3 | grep -E "==|chunk65|steady_us"
