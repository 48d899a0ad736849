# usage: gpu_trace.sh <tag> [configs...]   traces of decode_kernel (trace build) into gpurun_out/<tag>_trace_<cfg>.json
TAG=$1; shift
mkdir -p gpurun_out
export B200ATTN_LIB=$PWD/ltp-sglang_b200/b200attn/libb200attn_trace.so
for c in "$@"; do
  case $c in
    C1) a="--config C1 --flush";;
    C2tp8) a="--config C2 --tp-sim 8";;
    C3tp8) a="--config C3 --tp-sim 8 --batch 256";;
    C3tp2) a="--config C3 --tp-sim 2 --batch 256";;
    *) a="--config $c";;
  esac
  timeout 300 python tools/trace_decode.py $a --repeat 8 --out gpurun_out/${TAG}_trace_$c.json > /dev/null 2>> gpurun_out/${TAG}_trace.err
done
tail -n 3 gpurun_out/${TAG}_trace.err
python - <<P
import json
for c in "$*".split():
    try:
        d=json.load(open(f'gpurun_out/${TAG}_trace_{c}.json'))
    except Exception as e:
        print(c, 'ERR', e); continue
    print('==',c, d['config'], 'kernel', d['kernel_us_globaltimer'], 'ideal', d['ideal_us_at_6545GBs'], d['geometry'])
    print("  chunk65_clk", d.get("chunk65_clk"))
    for k in ['entry_to_first_copies_issued_us','entry_to_first_chunk_consumed_us','chunks_per_cta','steady_us_per_chunk','steady_gb_per_s_per_sm','stream_us_first_to_last_chunk','last_chunk_to_last_epilogue_us','cta_lifetime_us','idle_after_exit_us']:
        print('  ',k,d[k])
P
