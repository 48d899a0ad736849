set -x
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:extend_tc -s 3 -c 1 -o gpurun_out/ncu_ext_tc_P1 -f python bench_extend.py --config P1 --steps 3 --warmup 3 > gpurun_out/ncu14.log 2>&1; tail -3 gpurun_out/ncu14.log
