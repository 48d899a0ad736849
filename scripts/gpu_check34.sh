mkdir -p gpurun_out
for kv in bf16 fp8_e4m3; do
echo "== persistent $kv"; timeout 200 python tools/time_smallq.py $kv 2>&1 | tail -7
echo "== oneshot $kv"; B200ATTN_SMALLQ=oneshot timeout 200 python tools/time_smallq.py $kv 2>&1 | tail -7
done > gpurun_out/time_smallq.txt 2>&1
cat gpurun_out/time_smallq.txt
