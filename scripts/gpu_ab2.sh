# Same-box A/B on chosen cases: this tree against _ab/<name>.   usage: gpu_ab2.sh <tag> <name> "<case>:<bench args>" ...
TAG=$1; OLD=$2; shift 2
ROOT=$PWD
mkdir -p gpurun_out
for rep in 1 2; do
for cfgs in "$@"; do
  name=${cfgs%%:*}; a=${cfgs#*:}
  (cd $ROOT && timeout 300 python bench.py --no-cpu-baseline --sustain 0 $a > gpurun_out/${TAG}_new_${name}_$rep.json 2>/dev/null)
  (cd $ROOT/_ab/$OLD && timeout 300 python bench.py --no-cpu-baseline --sustain 0 $a > $ROOT/gpurun_out/${TAG}_${OLD}_${name}_$rep.json 2>/dev/null)
done
done
python - <<P
import json,glob
for f in sorted(glob.glob('gpurun_out/${TAG}_*_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); r=d['roofline']
        print(f.split('/')[-1][:-5].ljust(30), 'step ms', round(d['ms_per_step'],4), 'K2 ms', round(r['kernel_ms'],4), 'GB/s', round(r['achieved']), 'MHz', d['clocks']['sm_mhz'], d['clocks']['reasons'], round(d['clocks']['power_w_max'] or 0))
    except Exception as e:
        print(f, 'ERR', e)
P
