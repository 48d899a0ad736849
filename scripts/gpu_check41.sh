mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 180 2>&1 | tail -3
B="python bench.py --steps 20 --warmup 3 --scaling weak --no-cpu-baseline"
for tp in 1 2 4 8; do timeout 200 $B --config C3 --tp-sim $tp > gpurun_out/b41_C3_tp$tp.json 2>/dev/null; done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b41_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['value']), round(d['ms_per_step'],4), round(d['roofline']['achieved']), round(d['roofline']['kernel_ms'],4))
    except Exception as e: print(f,'ERR',e)
P
