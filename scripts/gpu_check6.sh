# rope kernel parity, gather roofline, layer bench at N=1 (and N=2 when two GPUs are visible)
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu6.log 2>&1; tail -5 gpurun_out/t_gpu6.log
tools/bin/gather_roofline > gpurun_out/gather_roofline.jsonl 2> gpurun_out/gather_roofline.err; cat gpurun_out/gather_roofline.jsonl
NG=$(nvidia-smi -L | wc -l)
python bench_layer.py --steps 20 > gpurun_out/layer_tp1.json 2> gpurun_out/layer_tp1.err; tail -3 gpurun_out/layer_tp1.err; cat gpurun_out/layer_tp1.json
python bench_layer.py --steps 20 --unfused > gpurun_out/layer_tp1_unfused.json 2> gpurun_out/layer_tp1_unfused.err; cat gpurun_out/layer_tp1_unfused.json
if [ "$NG" -ge 2 ]; then
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench_layer.py --gpus 2 --steps 20 > gpurun_out/layer_tp2.json 2> gpurun_out/layer_tp2.err; tail -3 gpurun_out/layer_tp2.err; cat gpurun_out/layer_tp2.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/b6_C2_n2.json 2> gpurun_out/b6_C2_n2.err; tail -3 gpurun_out/b6_C2_n2.err; cut -c1-1500 gpurun_out/b6_C2_n2.json
fi
