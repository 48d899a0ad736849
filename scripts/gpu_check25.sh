set -x
mkdir -p gpurun_out
for tp in 2 4 8; do
timeout 300 python bench.py --steps 20 --warmup 3 --tp-sim $tp --scaling weak --no-cpu-baseline > gpurun_out/bench25_weak_tp$tp.json 2> gpurun_out/bench25_weak_tp$tp.err; tail -1 gpurun_out/bench25_weak_tp$tp.err
done
python - <<'P'
import json
for tp in (2,4,8):
    try:
        d=json.load(open(f'gpurun_out/bench25_weak_tp{tp}.json'))
        print(tp, d['value'], d['ms_per_step'], d['roofline']['achieved'], d['roofline']['kernel_ms'], d['e2e']['value'], d['config']['workload'])
    except Exception as e: print(tp, 'ERR', e)
P
