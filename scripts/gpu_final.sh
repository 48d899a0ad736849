set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --timeout 600 > gpurun_out/t_gpu_final.log 2>&1; tail -3 gpurun_out/t_gpu_final.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_final_ref.json 2>/dev/null
timeout 400 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err
python - <<'P'
import json
d=json.load(open('gpurun_out/bench_final.json')); r=json.load(open('gpurun_out/bench_final_ref.json'))
print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'frac', d['roofline']['frac'], 'cpu', d['cpu_baseline']['value'], 'ref arm', r['value'], 'launches', d['gpu_launches'], d['clocks'])
P
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_final.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1; wc -l gpurun_out/launches_final.csv
