mkdir -p gpurun_out
echo "== extend on mma.sync kernel"; B200ATTN_EXTEND_IMPL=mma timeout 600 python -m pytest tests/test_gpu_extend.py -m gpu -x -q --timeout 300 -k "not forced" 2>&1 | tail -2
echo "== decode, padded cp.async rows"; B200ATTN_NO_SWIZZLE=1 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 300 2>&1 | tail -2
echo "== decode, no in-CTA slice merge"; B200ATTN_NO_XCH=1 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 300 2>&1 | tail -2
echo "== decode, bulk copies only"; B200ATTN_CPASYNC_MAX_ROW=0 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 300 2>&1 | tail -2
echo "== decode, cp.async everywhere"; B200ATTN_CPASYNC_MAX_ROW=2048 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 300 2>&1 | tail -2
