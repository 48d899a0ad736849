set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 180 > gpurun_out/t_gpu24.log 2>&1; tail -3 gpurun_out/t_gpu24.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench24.json 2> gpurun_out/bench24.err; tail -2 gpurun_out/bench24.err
timeout 400 python bench_extend.py --steps 20 --warmup 3 --sdpa > gpurun_out/bext24.json 2>gpurun_out/bext24.err
for c in P1 P3; do
timeout 300 ncu --set full --clock-control none --import-source on -k regex:extend_tc -s 3 -c 1 -o gpurun_out/ncu_ext_tc_${c}_v5 -f python bench_extend.py --config $c --steps 3 --warmup 3 > gpurun_out/ncu24_$c.log 2>&1; tail -2 gpurun_out/ncu24_$c.log
done
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_ext_P1_v5.csv python bench_extend.py --config P1 --steps 3 --warmup 3 > /dev/null 2>&1
