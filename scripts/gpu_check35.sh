set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --timeout 600 > gpurun_out/t_gpu35.log 2>&1; tail -3 gpurun_out/t_gpu35.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 400 python bench_extend.py --steps 20 --warmup 3 > gpurun_out/bext35.json 2>gpurun_out/bext35.err
timeout 300 python bench_layer.py --steps 20 --warmup 5 > gpurun_out/blayer35.json 2>gpurun_out/blayer35.err; tail -c 400 gpurun_out/blayer35.json
