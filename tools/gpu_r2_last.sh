#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/final_bench_n1.json 2> gpurun_out/final_bench.err; echo "bench exit $?"; tail -2 gpurun_out/final_bench.err
python - <<P
import json
d=json.load(open('gpurun_out/final_bench_n1.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e_with_image_readback']['value'], d['gpu_launches'])
print(d['stages_ms'], sum(d['stages_ms'].values()))
r=d['roofline']; print({k: r.get(k) for k in ('kernel','bound','achieved','peak','frac','traffic','issue','ncu')})
P
timeout 300 python tools/bench_configs.py sweep 3000000 > gpurun_out/cfg3m.json 2>/dev/null; cut -c1-400 gpurun_out/cfg3m.json
