# usage: gpu_ncu_extend.sh <tag> <config>: plain bench_extend run, then ncu --set full of the extend kernel of that config
TAG=$1; CFG=${2:-P1}
mkdir -p gpurun_out
CMD="python bench_extend.py --config $CFG --steps 3 --warmup 2"
$CMD > gpurun_out/${TAG}_plain_extend_$CFG.json 2> gpurun_out/${TAG}_plain_extend_$CFG.err && cat gpurun_out/${TAG}_plain_extend_$CFG.json | cut -c1-300 &&
ncu --set full --clock-control none --import-source on -k regex:extend_tc -s 2 -c 1 -f -o gpurun_out/${TAG}_ncu_extend_$CFG $CMD > gpurun_out/${TAG}_ncu_extend_$CFG.log 2>&1
tail -n 3 gpurun_out/${TAG}_ncu_extend_$CFG.log; ls -la gpurun_out/${TAG}_ncu_extend_$CFG.ncu-rep
