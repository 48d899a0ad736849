set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu12.log 2>&1; tail -8 gpurun_out/t_gpu12.log
for c in C2 C3; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 6 -c 1 -o gpurun_out/ncu_full_v3_$c -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --config $c > gpurun_out/ncu_full_v3_$c.log 2>&1
done
