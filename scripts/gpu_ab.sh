# Same-box A/B of decode_kernel: this tree against a snapshot of another commit exported to _ab/<name>
# (git archive <commit> | tar -x -C _ab/<name>; make -C _ab/<name>/ltp-sglang_b200/csrc).
# usage: gpu_ab.sh <tag> [<name>]      writes gpurun_out/<tag>_{new,<name>}_<case>.json
TAG=${1:-ab}
OLD=${2:-r1}
ROOT=$PWD
mkdir -p gpurun_out
run_case() {  # dir label case-name args...
  local dir=$1 label=$2 name=$3; shift 3
  (cd $dir && timeout 300 python bench.py --no-cpu-baseline "$@" > $ROOT/gpurun_out/${TAG}_${label}_${name}.json 2> $ROOT/gpurun_out/${TAG}_${label}_${name}.err)
}
for rep in 1 2; do
for side in new $OLD; do
  if [ $side = new ]; then dir=$ROOT; extra="--sustain 0"; else dir=$ROOT/_ab/$OLD; extra=""; fi
  [ -d $dir ] || continue
  run_case $dir $side C1_$rep --config C1 $extra
  run_case $dir $side C2_$rep --config C2 $extra
  run_case $dir $side C3_$rep --config C3 $extra
  run_case $dir $side C2tp8strong_$rep --config C2 --tp-sim 8 --scaling strong $extra
  run_case $dir $side C3tp8_$rep --config C3 --tp-sim 8 $extra
done
done
python - <<P
import json,glob
for f in sorted(glob.glob('gpurun_out/${TAG}_*_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        r=d['roofline']
        print(f.split('/')[-1][:-5].ljust(34), 'step ms', round(d['ms_per_step'],4), 'K2 ms', round(r.get('kernel_ms',0),4), 'GB/s', round(r['achieved']), 'frac', round(r['frac'],3), 'e2e ms', round(d['e2e'].get('ms_per_step',0),4))
    except Exception as e:
        print(f, 'ERR', e)
P
