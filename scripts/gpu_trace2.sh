mkdir -p gpurun_out
export B200ATTN_LIB=$PWD/tools/bin/libb200attn_trace.so
timeout 120 python tools/trace_tc_verify.py > gpurun_out/trace_verify.txt 2>&1
head -4 gpurun_out/trace_verify.txt; grep -E "n= 3[012]|^CTAs|^st=  [0-3]" gpurun_out/trace_verify.txt
