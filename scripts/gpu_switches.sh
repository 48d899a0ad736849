# GPU suites under every tuning / test switch (each must be green).  usage: gpu_switches.sh <out file>
OUT=${1:-gpurun_out/r02_tests_under_tuning_switches.txt}
mkdir -p gpurun_out; : > $OUT
for sw in "B200ATTN_NO_NARROW=1" "B200ATTN_NO_SUBS=1" "B200ATTN_NO_BALANCE=1" "B200ATTN_NO_XCH=1" "B200ATTN_NO_SWIZZLE=1" "B200ATTN_NO_PDL=1" "B200ATTN_CPASYNC_MAX_ROW=0" "B200ATTN_EXTEND_IMPL=mma" "B200ATTN_SMALLQ=oneshot" "B200ATTN_SMALLQ=persistent" "B200ATTN_DYN_PCT=15" "B200ATTN_TC_LAYOUT=two"; do
  echo "== $sw" >> $OUT
  env $sw timeout 900 python -m pytest tests -m gpu -q --timeout 600 2>&1 | tail -n 2 >> $OUT
done
cat $OUT
