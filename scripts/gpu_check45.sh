mkdir -p gpurun_out
B="python bench.py --steps 20 --warmup 3 --scaling weak --no-cpu-baseline"
timeout 200 $B --config C2 --tp-sim 8 > gpurun_out/b45_C2_tp8.json 2>/dev/null
timeout 200 $B --config C2 --tp-sim 4 > gpurun_out/b45_C2_tp4.json 2>/dev/null
timeout 200 $B --config C3 --tp-sim 2 > gpurun_out/b45_C3_tp2.json 2>/dev/null
timeout 200 $B --config C4 --tp-sim 8 > gpurun_out/b45_C4_tp8.json 2>/dev/null
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b45_*.json')):
    d=json.load(open(f)); print(f, round(d['value']), round(d['ms_per_step'],4), round(d['roofline']['achieved']), round(d['roofline']['kernel_ms'],4))
P
