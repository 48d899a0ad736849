set -x
mkdir -p gpurun_out
B="python bench.py --steps 20 --warmup 3 --scaling weak --no-cpu-baseline"
B200ATTN_CPASYNC_MAX_ROW=0 timeout 200 $B --tp-sim 8 > gpurun_out/b26_tp8_bulk.json 2>/dev/null
B200ATTN_CPASYNC_MAX_ROW=512 timeout 200 $B --tp-sim 4 > gpurun_out/b26_tp4_cpasync.json 2>/dev/null
timeout 200 $B --tp-sim 8 --config C4 > gpurun_out/b26_c4_tp8.json 2>/dev/null
timeout 200 $B --tp-sim 8 --config C3 > gpurun_out/b26_c3_tp8.json 2>/dev/null
for tp in 8 4; do
timeout 300 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 6 -c 1 -o gpurun_out/ncu26_decode_tp$tp -f python bench.py --steps 3 --warmup 3 --scaling weak --no-cpu-baseline --tp-sim $tp > gpurun_out/ncu26_tp$tp.log 2>&1; tail -2 gpurun_out/ncu26_tp$tp.log
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b26_*.json')):
    try:
        d=json.load(open(f)); print(f, round(d['value']), d['ms_per_step'], round(d['roofline']['achieved']), d['roofline']['kernel_ms'])
    except Exception as e: print(f,'ERR',e)
P
