# 1-GPU: ncu --set full of the decode kernel (C2, C3), tp-sim dry runs; outputs under gpurun_out/.
set -x
mkdir -p gpurun_out
for c in C2 C3; do
ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 6 -c 1 -o gpurun_out/ncu_full_$c -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --config $c > gpurun_out/ncu_full_$c.log 2>&1
done
for t in 2 4 8; do
python bench.py --steps 20 --warmup 3 --tp-sim $t --scaling strong --no-cpu-baseline > gpurun_out/b4_C2_tpsim$t.json 2> gpurun_out/b4_C2_tpsim$t.err; tail -3 gpurun_out/b4_C2_tpsim$t.err
python bench.py --steps 20 --warmup 3 --tp-sim $t --config C4 --scaling strong --no-cpu-baseline > gpurun_out/b4_C4_tpsim$t.json 2> gpurun_out/b4_C4_tpsim$t.err; tail -3 gpurun_out/b4_C4_tpsim$t.err
python bench.py --steps 20 --warmup 3 --tp-sim $t --config C3 --scaling strong --no-cpu-baseline > gpurun_out/b4_C3_tpsim$t.json 2> gpurun_out/b4_C3_tpsim$t.err; tail -3 gpurun_out/b4_C3_tpsim$t.err
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b4_*.json')):
    for l in open(f):
        if not l.startswith('{'): continue
        d=json.loads(l); r=d['roofline']; e=d['e2e']
        print(f, 'value %.0f ms %.4f | K2 %.4f ms %.0f GB/s frac %.3f | e2e %.0f'%(d['value'],d['ms_per_step'],r['kernel_ms'],r['achieved'],r['frac'],e['value']))
P
