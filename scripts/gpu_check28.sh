set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 180 > gpurun_out/t_gpu28.log 2>&1; tail -3 gpurun_out/t_gpu28.log
B="python bench.py --steps 20 --warmup 3 --scaling weak --no-cpu-baseline"
timeout 200 $B --tp-sim 8 > gpurun_out/b28_c2_tp8.json 2>/dev/null
timeout 200 $B --tp-sim 8 --config C4 > gpurun_out/b28_c4_tp8.json 2>/dev/null
timeout 200 $B --tp-sim 8 --config C3 > gpurun_out/b28_c3_tp8.json 2>/dev/null
timeout 200 $B --tp-sim 4 --config C3 > gpurun_out/b28_c3_tp4.json 2>/dev/null
B200ATTN_CPASYNC_MAX_ROW=512 timeout 200 $B --tp-sim 4 > gpurun_out/b28_c2_tp4_cpasync.json 2>/dev/null
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b28_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['value']), d['ms_per_step'], round(d['roofline']['achieved']), d['roofline']['kernel_ms'])
    except Exception as e: print(f,'ERR',e)
P
