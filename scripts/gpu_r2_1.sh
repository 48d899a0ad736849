# round 2, call 1: all GPU tests (new: spec/graph, full-size configs), then decode-kernel timelines
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -x > gpurun_out/r2_t1.log 2>&1; tail -15 gpurun_out/r2_t1.log
export B200ATTN_LIB=$PWD/ltp-sglang_b200/b200attn/libb200attn_trace.so
timeout 300 python tools/trace_decode.py --config C2 --out gpurun_out/r2_trace_C2.json > /dev/null 2> gpurun_out/r2_trace_C2.err
timeout 300 python tools/trace_decode.py --config C2 --tp-sim 8 --out gpurun_out/r2_trace_C2_tp8_b128.json > /dev/null 2> gpurun_out/r2_trace_C2_tp8.err
timeout 300 python tools/trace_decode.py --config C1 --flush --out gpurun_out/r2_trace_C1.json > /dev/null 2> gpurun_out/r2_trace_C1.err
timeout 300 python tools/trace_decode.py --config C3 --out gpurun_out/r2_trace_C3.json > /dev/null 2> gpurun_out/r2_trace_C3.err
tail -3 gpurun_out/r2_trace_*.err
