mkdir -p gpurun_out
export B200ATTN_CPASYNC_MAX_ROW=2048
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 180 2>&1 | tail -2
B="python bench.py --steps 20 --warmup 3 --scaling weak --no-cpu-baseline"
for cfg in C2 C3 C4 C5 C1; do timeout 200 $B --config $cfg > gpurun_out/b40_${cfg}_tp1.json 2>/dev/null; done
timeout 200 $B --config C2 --tp-sim 2 > gpurun_out/b40_C2_tp2.json 2>/dev/null
timeout 200 $B --config C3 --tp-sim 2 > gpurun_out/b40_C3_tp2.json 2>/dev/null
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b40_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['value']), round(d['ms_per_step'],4), round(d['roofline']['achieved']), round(d['roofline']['kernel_ms'],4))
    except Exception as e: print(f,'ERR',e)
P
