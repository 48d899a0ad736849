# bench_extend (P1, P2, P1d64) on the product library and on every variant library under _ab/libs/ (same box).
# usage: gpu_extend_libs.sh <tag> [configs]
TAG=$1; CFGS=${2:-P1 P2 P1d64}
mkdir -p gpurun_out
for lib in product _ab/libs/*.so; do
  name=$(basename $lib .so)
  for c in $CFGS; do
    if [ $lib = product ]; then timeout 300 python bench_extend.py --config $c >> gpurun_out/${TAG}_$name.jsonl 2>> gpurun_out/${TAG}_$name.err
    else B200ATTN_LIB=$lib timeout 300 python bench_extend.py --config $c >> gpurun_out/${TAG}_$name.jsonl 2>> gpurun_out/${TAG}_$name.err; fi
  done
done
python - <<P
import json,glob
for f in sorted(glob.glob('gpurun_out/${TAG}_*.jsonl')):
    for l in open(f):
        try: d=json.loads(l)
        except Exception: continue
        r=d.get('roofline',{})
        print(f.split('/')[-1][:-6].ljust(28), str(d.get('config',{}).get('workload',''))[:40].ljust(40), 'ms', round(d.get('ms_per_step',0),4), round(r.get('achieved',0)), r.get('unit'))
P
