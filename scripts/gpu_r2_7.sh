# round 2, call 7: setmaxnreg (no spills), header in registers, L2 prefetch + batched loads in the merge shortcut, wide merge for many partials
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -x > gpurun_out/r2_t7.log 2>&1; tail -n 4 gpurun_out/r2_t7.log
bash scripts/gpu_ab.sh r2g r1
export B200ATTN_LIB=$PWD/ltp-sglang_b200/b200attn/libb200attn_trace.so
for c in C2 C3; do
timeout 300 python tools/trace_decode.py --config $c --repeat 8 --out gpurun_out/r2g_trace_$c.json > /dev/null 2>> gpurun_out/r2g_trace.err
done
timeout 300 python tools/trace_decode.py --config C1 --flush --repeat 8 --out gpurun_out/r2g_trace_C1.json > /dev/null 2>> gpurun_out/r2g_trace.err
timeout 300 python tools/trace_decode.py --config C2 --tp-sim 8 --repeat 8 --out gpurun_out/r2g_trace_C2_tp8_b128.json > /dev/null 2>> gpurun_out/r2g_trace.err
tail -n 3 gpurun_out/r2g_trace.err
