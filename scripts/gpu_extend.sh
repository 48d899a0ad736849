# extend path: tests, then bench_extend (this tree) and, where _ab/<name> exists, the same bench on that tree.  usage: gpu_extend.sh <tag> [<name>]
TAG=$1; OLD=${2:-}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extend.py tests/test_gpu_spec_graph.py -m gpu -q --timeout 600 > gpurun_out/${TAG}_ext_tests.log 2>&1; tail -n 3 gpurun_out/${TAG}_ext_tests.log
timeout 600 python bench_extend.py > gpurun_out/${TAG}_extend_new.jsonl 2> gpurun_out/${TAG}_extend_new.err
if [ -n "$OLD" ] && [ -d _ab/$OLD ]; then (cd _ab/$OLD && timeout 600 python bench_extend.py > ../../gpurun_out/${TAG}_extend_$OLD.jsonl 2>/dev/null); fi
python - <<P
import json,glob
for f in sorted(glob.glob('gpurun_out/${TAG}_extend_*.jsonl')):
    print(f)
    for l in open(f):
        try: d=json.loads(l)
        except Exception: continue
        r=d.get('roofline',{})
        print('  ', str(d.get('config',{}).get('workload',''))[:60].ljust(60), 'ms', round(d.get('ms_per_step',0),4), r.get('bound'), round(r.get('achieved',0)), r.get('unit'), 'frac', round(r.get('frac',0),3))
P
