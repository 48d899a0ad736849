# page-size sweep on the small-row shapes (is page contiguity already exploited by the per-row copies?)
TAG=$1
mkdir -p gpurun_out
for cfgs in "C3tp8_p1:--config C3 --tp-sim 8" "C3tp8_p16:--config C3 --tp-sim 8 --page-size 16" "C3tp8_p64:--config C3 --tp-sim 8 --page-size 64" "C3tp4_p16:--config C3 --tp-sim 4 --page-size 16" "C2tp8_p16:--config C2 --tp-sim 8 --page-size 16" "C5_p16:--config C5" "C5_p64:--config C5 --page-size 64" "C5tp8_p16:--config C5 --tp-sim 8" "C5tp8_p64:--config C5 --tp-sim 8 --page-size 64"; do
  name=${cfgs%%:*}; a=${cfgs#*:}
  timeout 300 python bench.py --no-cpu-baseline --sustain 0 $a > gpurun_out/${TAG}_pages_$name.json 2> gpurun_out/${TAG}_pages_$name.err
done
python - <<P
import json,glob
for f in sorted(glob.glob('gpurun_out/${TAG}_pages_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); r=d['roofline']
        print(f.split('_pages_')[1][:-5].ljust(12), 'step ms', round(d['ms_per_step'],4), 'K2 ms', round(r.get('kernel_ms',0),4), 'GB/s', round(r['achieved']), 'frac', round(r['frac'],3), 'unique GB/s', round(r.get('achieved_unique',0)), 'MHz', d['clocks']['sm_mhz'], d['clocks']['reasons'])
    except Exception as e:
        print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-300:])
P
