# Final artefacts of a round: GPU tests, smoke, both bench arms, bench sweep of every config and the per-rank
# shard shapes, launch list + ncu --set full of decode_kernel on C2 and C3.   usage: gpu_final.sh <tag>
TAG=${1:-r02}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/${TAG}_tests.log 2>&1; tail -n 3 gpurun_out/${TAG}_tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference_arm.json 2>/dev/null
timeout 600 python bench.py > gpurun_out/${TAG}_bench_C2_default.json 2> gpurun_out/${TAG}_bench_C2_default.err
for cfgs in "C1:--config C1" "C3:--config C3" "C4:--config C4" "C5:--config C5" "C5p64:--config C5 --page-size 64" "C2p16:--config C2 --page-size 16" \
            "C2tp8strong:--config C2 --tp-sim 8 --scaling strong" "C2tp2:--config C2 --tp-sim 2" "C2tp4:--config C2 --tp-sim 4" "C2tp8:--config C2 --tp-sim 8" \
            "C3tp2:--config C3 --tp-sim 2" "C3tp4:--config C3 --tp-sim 4" "C3tp8:--config C3 --tp-sim 8" "C3tp8p16:--config C3 --tp-sim 8 --page-size 16" \
            "C4tp8:--config C4 --tp-sim 8" "C5tp8:--config C5 --tp-sim 8"; do
  name=${cfgs%%:*}; a=${cfgs#*:}
  timeout 300 python bench.py --no-cpu-baseline --sustain 0 $a > gpurun_out/${TAG}_bench_$name.json 2> gpurun_out/${TAG}_bench_$name.err
done
python - <<P
import json,glob
for f in sorted(glob.glob('gpurun_out/${TAG}_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        if d.get('impl')=='reference':
            print('reference arm', round(d['value']), d['cpu_baseline']['cores'], 'cores'); continue
        r=d['roofline']
        print(f.split('_bench_')[1][:-5].ljust(13), 'step ms', round(d['ms_per_step'],4), 'tok/s', round(d['value']), 'K2 ms', round(r.get('kernel_ms',0),4), 'GB/s', round(r['achieved']), 'frac', round(r['frac'],3), 'e2e ms', round(d['e2e'].get('ms_per_step',0),4), 'MHz', d['clocks']['sm_mhz'], d['clocks']['reasons'], 'W', round(d['clocks']['power_w_max'] or 0))
    except Exception as e:
        print(f, 'ERR', e)
P
bash scripts/gpu_ncu.sh $TAG C2 --config C2
bash scripts/gpu_ncu.sh $TAG C3 --config C3
