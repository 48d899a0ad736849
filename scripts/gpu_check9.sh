set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu9.log 2>&1; tail -3 gpurun_out/t_gpu9.log
run() { n=$1; shift; env "$@" timeout 300 python bench.py --steps 20 --warmup 3 --config C2 --no-cpu-baseline > gpurun_out/b9_$n.json 2> gpurun_out/b9_$n.err || tail -3 gpurun_out/b9_$n.err; }
run dynoff B200ATTN_DYN_PCT=-1
run dyn20_512k B200ATTN_DYN_PCT=20
run dyn20_1m B200ATTN_DYN_PCT=20 B200ATTN_DYN_BYTES=1048576
run dyn10_1m B200ATTN_DYN_PCT=10 B200ATTN_DYN_BYTES=1048576
run dyn10_2m B200ATTN_DYN_PCT=10 B200ATTN_DYN_BYTES=2097152
run dyn5_1m B200ATTN_DYN_PCT=5 B200ATTN_DYN_BYTES=1048576
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b9_*.json')):
    for l in open(f):
        if not l.startswith('{'): continue
        d=json.loads(l); r=d['roofline']; e=d['e2e']
        print(f, 'value %.0f ms %.4f | K2 %.4f ms %.0f GB/s frac %.3f | e2e %.0f'%(d['value'],d['ms_per_step'],r['kernel_ms'],r['achieved'],r['frac'],e['value']))
P
