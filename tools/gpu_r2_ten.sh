#!/bin/bash
# parity of the new build (clean backward scratch, statistics off the critical path, lean scatter, adaptive 3-pass depth order),
# then the bench line of the product library and of two variants built with the old settings
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider -x > gpurun_out/tests.log 2>&1; echo "tests exit $?"; tail -6 gpurun_out/tests.log
for v in "" lean4 old; do
  if [ -n "$v" ]; then export B200SPLAT_LIB=$PWD/3dgs_study_b200/variants/libb200splat_$v.so; fi
  timeout 300 python bench.py --steps 60 --warmup 10 --no-train-iter --no-cpu-baseline > gpurun_out/bench_v_$v.json 2> gpurun_out/bench_v_$v.err; echo "variant=[$v] bench exit $?"; tail -2 gpurun_out/bench_v_$v.err
  python - <<P
import json
try:
    d=json.load(open('gpurun_out/bench_v_$v.json'))
    print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e_with_image_readback']['value'], d['stages_ms'])
except Exception as e: print('bench parse', e)
P
done
