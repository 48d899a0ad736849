mkdir -p gpurun_out
B="python bench.py --steps 20 --warmup 3 --scaling weak --no-cpu-baseline"
B200ATTN_CPASYNC_MAX_ROW=256 timeout 200 $B --tp-sim 2 --config C3 > gpurun_out/b36_c3_tp2_bulk512.json 2>/dev/null
B200ATTN_CPASYNC_MAX_ROW=128 timeout 200 $B --tp-sim 4 --config C3 > gpurun_out/b36_c3_tp4_bulk256.json 2>/dev/null
B200ATTN_NO_XCH=1 timeout 200 $B --tp-sim 2 --config C3 > gpurun_out/b36_c3_tp2_noxch.json 2>/dev/null
B200ATTN_NO_SWIZZLE=1 timeout 200 $B --tp-sim 2 --config C3 > gpurun_out/b36_c3_tp2_noswz.json 2>/dev/null
timeout 200 $B --tp-sim 2 --config C3 > gpurun_out/b36_c3_tp2.json 2>/dev/null
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b36_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['value']), d['ms_per_step'], round(d['roofline']['achieved']), d['roofline']['kernel_ms'])
    except Exception as e: print(f,'ERR',e)
P
