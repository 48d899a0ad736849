# usage: gpu_multi.sh N   (run under gpurun --gpus N)
N=$1
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; tail -2 gpurun_out/scale_n$N.err | cut -c1-300
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench_layer.py --gpus $N --steps 20 --warmup 5 > gpurun_out/layer_n$N.json 2> gpurun_out/layer_n$N.err; tail -2 gpurun_out/layer_n$N.err | cut -c1-300
python - <<P
import json
for f in ('gpurun_out/scale_n$N.json','gpurun_out/layer_n$N.json'):
    for l in open(f):
        l=l.strip()
        if l.startswith('{'):
            d=json.loads(l); print(f, d.get('n_gpus'), round(d['value']), d['ms_per_step'], d.get('roofline',{}).get('achieved'), d.get('e2e',{}).get('value'), d.get('breakdown_ms'), (d.get('strong_scaling') or {}).get('value'))
P
