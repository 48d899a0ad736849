set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_extend.py -m gpu -x -q --timeout 120 > gpurun_out/t_ext32.log 2>&1; tail -5 gpurun_out/t_ext32.log
timeout 300 python bench_extend.py --steps 20 --warmup 3 --config P3 > gpurun_out/bext32_P3.json 2>gpurun_out/bext32_P3.err; tail -c 600 gpurun_out/bext32_P3.json
timeout 300 python bench_extend.py --steps 20 --warmup 3 --config P3f > gpurun_out/bext32_P3f.json 2>gpurun_out/bext32_P3f.err; tail -c 600 gpurun_out/bext32_P3f.json
B200ATTN_SMALLQ=oneshot timeout 300 python bench_extend.py --steps 20 --warmup 3 --config P3 > gpurun_out/bext32_P3_oneshot.json 2>/dev/null; tail -c 600 gpurun_out/bext32_P3_oneshot.json
