mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extend.py -m gpu -x -q --timeout 600 2>&1 | tail -2
timeout 400 python bench_extend.py --steps 20 --warmup 3 > gpurun_out/bext43.json 2>gpurun_out/bext43.err
python - <<'P'
import json
for l in open('gpurun_out/bext43.json'):
    l=l.strip()
    if not l.startswith('{'): continue
    d=json.loads(l); r=d.get('roofline',{})
    print(d['config']['workload'][:30], r.get('kernel_ms'), round(r.get('tflops',0)), round(r.get('gb_per_s',0)))
P
echo "== persistent bf16"; B200ATTN_SMALLQ=persistent timeout 200 python tools/time_smallq.py bf16 2>&1 | tail -7
