# round 2, call 9: transposed decode consumer (tokens in M): tests, then A/B against the round-1 tree
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2_t9.log 2>&1; tail -n 15 gpurun_out/r2_t9.log
bash scripts/gpu_ab.sh r2i r1
