timeout 120 python -m pytest tests/test_gpu_extend.py -x -q --timeout 40 2>&1 | tail -2
B200ATTN_LIB=tools/bin/libb200attn_trace.so timeout 60 python tools/trace_tc.py 2>&1 | grep -E "^n= 3[01]|steady|st= 6"
timeout 90 python bench_extend.py --steps 10 2>/dev/null > gpurun_out/bext_last.json
python - <<'P'
import json
for l in open('gpurun_out/bext_last.json'):
    d=json.loads(l); r=d["roofline"]; print(d["config"]["workload"][:3], round(r["kernel_ms"],3), "ms", round(r["tflops"]), "TFLOP/s", round(r["gb_per_s"]), "GB/s")
P
