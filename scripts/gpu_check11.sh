set -x
mkdir -p gpurun_out
for c in C2 C1; do
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches11_$c.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline --config $c > gpurun_out/ncu11_$c.log 2>&1
done
