# 2-GPU: bench.py and bench_layer.py under torchrun, every command under its own timeout
set -x
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/b7_C2_n2.json 2> gpurun_out/b7_C2_n2.err; echo rc=$?; tail -3 gpurun_out/b7_C2_n2.err; cut -c1-600 gpurun_out/b7_C2_n2.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench_layer.py --gpus 2 --steps 20 > gpurun_out/layer_tp2.json 2> gpurun_out/layer_tp2.err; echo rc=$?; tail -3 gpurun_out/layer_tp2.err
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/b7_ref_n2.json 2> gpurun_out/b7_ref_n2.err; echo rc=$?; cut -c1-300 gpurun_out/b7_ref_n2.json
