mkdir -p gpurun_out
timeout 1100 python tools/bench_competitors.py C2 C3 C4 C2/2 C2/4 C2/8 C4/8 C3/8 > gpurun_out/competitors.jsonl 2> gpurun_out/competitors.err
cat gpurun_out/competitors.jsonl | cut -c1-900; tail -3 gpurun_out/competitors.err | cut -c1-400
