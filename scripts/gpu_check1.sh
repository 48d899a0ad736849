# 1-GPU check: parity tests, smoke, default bench, small config, reference arm.  Outputs under gpurun_out/.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu.log 2>&1; tail -5 gpurun_out/t_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
python bench.py --steps 20 --warmup 3 > gpurun_out/b1.json 2> gpurun_out/b1.err; tail -3 gpurun_out/b1.err
for c in C1 C3 C4 C5; do
python bench.py --steps 20 --warmup 3 --config $c --no-cpu-baseline > gpurun_out/b1_$c.json 2> gpurun_out/b1_$c.err; tail -3 gpurun_out/b1_$c.err
done
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/bref.json 2> gpurun_out/bref.err; tail -3 gpurun_out/bref.err
cat gpurun_out/b1.json | cut -c1-3000
nproc; lscpu | grep -i 'model name'
