# per-SM steady-state rates of decode_kernel across launches (is a slow SM always slow?)
mkdir -p gpurun_out
export B200ATTN_LIB=$PWD/ltp-sglang_b200/b200attn/libb200attn_trace.so
timeout 300 python tools/trace_decode.py --config C2 --per-cta --repeat 4 --out gpurun_out/r2c_trace_C2_percta.json 2>&1 | tail -n 5
timeout 300 python tools/trace_decode.py --config C4 --per-cta --repeat 3 --out gpurun_out/r2c_trace_C4_percta.json 2>&1 | tail -n 5
