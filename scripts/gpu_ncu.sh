# usage: gpu_ncu.sh <tag> <name> <bench args...>: plain run, then launch list + ncu --set full of decode_kernel for one bench command line
TAG=$1; NAME=$2; shift 2
mkdir -p gpurun_out
CMD="python bench.py --no-cpu-baseline --sustain 0 --steps 3 --warmup 3 $*"
$CMD > gpurun_out/${TAG}_plain_$NAME.json 2> gpurun_out/${TAG}_plain_$NAME.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_$NAME.csv $CMD > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 30 -c 2 -f -o gpurun_out/${TAG}_ncu_$NAME $CMD > gpurun_out/${TAG}_ncu_$NAME.log 2>&1
tail -n 3 gpurun_out/${TAG}_ncu_$NAME.log; ls -la gpurun_out/${TAG}_ncu_$NAME.ncu-rep
