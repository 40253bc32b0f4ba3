#!/bin/bash
# Round 2: hand-written binning + canonical exp on the GPU for the first time.
mkdir -p gpurun_out; ./tools/b200_peaks > gpurun_out/peaks.json 2>&1; cat gpurun_out/peaks.json
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider --durations=8 > gpurun_out/tests.log 2>&1; echo "tests exit $?"; tail -45 gpurun_out/tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
timeout 400 python bench.py --steps 50 --warmup 10 --no-train-iter > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; python - <<'P'
import json
try:
    d=json.load(open('gpurun_out/bench.json'))
    print(d['value'], d['ms_per_step'], d['e2e']['value'], d['stages_ms']); print(d['roofline']['kernel'], d['roofline']['bound'], d['roofline']['frac'], d['roofline'].get('work')); print(d.get('cpu_baseline'))
except Exception as e: print('bench parse', e)
P
tail -5 gpurun_out/bench.err
bash tools/sanitize.sh memcheck
