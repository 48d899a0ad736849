# round 2, call 5 (after the machine was replaced): tests, traces, bench of every config, same-box A/B against the round-1 tree
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -x > gpurun_out/r2_t5.log 2>&1; tail -n 4 gpurun_out/r2_t5.log
(
export B200ATTN_LIB=$PWD/ltp-sglang_b200/b200attn/libb200attn_trace.so
for c in C2 C3 C1; do
fl=""; [ $c = C1 ] && fl="--flush"
timeout 300 python tools/trace_decode.py --config $c $fl --repeat 6 --out gpurun_out/r2e_trace_$c.json > /dev/null 2>> gpurun_out/r2e_trace.err
done
timeout 300 python tools/trace_decode.py --config C2 --tp-sim 8 --repeat 6 --out gpurun_out/r2e_trace_C2_tp8_b128.json > /dev/null 2>> gpurun_out/r2e_trace.err
)
timeout 600 python bench.py > gpurun_out/r2e_bench_C2_full.json 2> gpurun_out/r2e_bench_C2_full.err
timeout 300 python bench.py --config C4 --no-cpu-baseline --sustain 0 > gpurun_out/r2e_bench_C4.json 2>/dev/null
timeout 300 python bench.py --config C5 --no-cpu-baseline --sustain 0 > gpurun_out/r2e_bench_C5.json 2>/dev/null
B200ATTN_NO_BALANCE=1 timeout 300 python bench.py --config C3 --no-cpu-baseline --sustain 0 > gpurun_out/r2e_bench_C3_nobalance.json 2>/dev/null
B200ATTN_NO_PDL=1 timeout 300 python bench.py --config C3 --no-cpu-baseline --sustain 0 > gpurun_out/r2e_bench_C3_nopdl.json 2>/dev/null
set +x
bash scripts/gpu_ab.sh r2e r1
