mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_extend.py -m gpu -x -q --timeout 600 2>&1 | tail -2
echo "== persistent bf16"; B200ATTN_SMALLQ=persistent timeout 200 python tools/time_smallq.py bf16 2>&1 | tail -7
echo "== persistent fp8"; B200ATTN_SMALLQ=persistent timeout 200 python tools/time_smallq.py fp8_e4m3 2>&1 | tail -7
