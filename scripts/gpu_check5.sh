# cp.async producer path: parity, then K2 bandwidth vs the row-size threshold.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu5.log 2>&1; tail -5 gpurun_out/t_gpu5.log
run() { # name, env threshold, args...
  n=$1; thr=$2; shift 2
  B200ATTN_CPASYNC_MAX_ROW=$thr python bench.py --steps 20 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/b5_${n}_thr$thr.json 2> gpurun_out/b5_${n}_thr$thr.err || tail -3 gpurun_out/b5_${n}_thr$thr.err
}
for thr in 0 512 4096; do
  run C2 $thr --config C2
  run C3 $thr --config C3
  run C1 $thr --config C1
  run C2tp4 $thr --config C2 --tp-sim 4 --scaling strong
  run C2tp8 $thr --config C2 --tp-sim 8 --scaling strong
  run C3tp2 $thr --config C3 --tp-sim 2 --scaling strong
  run C3tp8 $thr --config C3 --tp-sim 8 --scaling strong
  run C4tp8 $thr --config C4 --tp-sim 8 --scaling strong
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b5_*.json')):
    for l in open(f):
        if not l.startswith('{'): continue
        d=json.loads(l); r=d['roofline']; e=d['e2e']
        print(f, 'value %.0f ms %.4f | K2 %.4f ms %.0f GB/s frac %.3f | e2e %.0f'%(d['value'],d['ms_per_step'],r['kernel_ms'],r['achieved'],r['frac'],e['value']))
P
