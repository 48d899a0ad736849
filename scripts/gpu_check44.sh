mkdir -p gpurun_out
B200ATTN_SMALLQ=persistent timeout 300 ncu --set full --clock-control none --import-source on -k regex:smallq -s 3 -c 1 -o gpurun_out/ncu_smallq_P3 -f python bench_extend.py --config P3 --steps 3 --warmup 3 > gpurun_out/ncu44_a.log 2>&1; tail -1 gpurun_out/ncu44_a.log
timeout 300 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 6 -c 1 -o gpurun_out/ncu_decode_C3_tp2 -f python bench.py --config C3 --tp-sim 2 --scaling weak --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu44_b.log 2>&1; tail -1 gpurun_out/ncu44_b.log
