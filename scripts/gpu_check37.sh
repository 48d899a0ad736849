set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --timeout 600 > gpurun_out/t_gpu37.log 2>&1; tail -3 gpurun_out/t_gpu37.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/bench37.json 2> gpurun_out/bench37.err; python -c "
import json; d=json.load(open('gpurun_out/bench37.json')); print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['cpu_baseline']['value'], d['gpu_launches'])"
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench37_ref.json 2>/dev/null; cut -c1-300 gpurun_out/bench37_ref.json
