set -x
mkdir -p gpurun_out
B="python bench.py --steps 20 --warmup 3 --scaling weak --no-cpu-baseline"
for cfg in C2 C3 C4; do for tp in 1 2 4 8; do
timeout 200 $B --tp-sim $tp --config $cfg > gpurun_out/b31_${cfg}_tp$tp.json 2>/dev/null
done; done
timeout 200 $B --config C1 > gpurun_out/b31_C1_tp1.json 2>/dev/null
timeout 200 $B --config C5 > gpurun_out/b31_C5_tp1.json 2>/dev/null
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/bench31.json 2> gpurun_out/bench31.err
timeout 300 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 6 -c 1 -o gpurun_out/ncu31_decode_tp8 -f python bench.py --steps 3 --warmup 3 --scaling weak --no-cpu-baseline --tp-sim 8 > gpurun_out/ncu31_tp8.log 2>&1; tail -2 gpurun_out/ncu31_tp8.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches31_C2.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
