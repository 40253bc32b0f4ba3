#!/bin/bash
# Final evidence of round 2 on one B200: GPU suite, smoke, the bench line (both arms), configs 3 / 5 / 4 at N = 1, translucent
# preset, then the ncu launch list of the bench command and one --set full capture (each after its command exited 0 without ncu).
mkdir -p gpurun_out; rm -f gpurun_out/widen_observed.jsonl
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider -x > gpurun_out/final_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/final_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/final_smoke.log
timeout 600 python bench.py > gpurun_out/final_bench_n1.json 2> gpurun_out/final_bench.err; echo "bench exit $?"; tail -2 gpurun_out/final_bench.err; cut -c1-400 gpurun_out/final_bench_n1.json
timeout 600 python bench.py --impl reference > gpurun_out/final_ref_n1.json 2> gpurun_out/final_ref.err; echo "ref exit $?"; cut -c1-200 gpurun_out/final_ref_n1.json
timeout 300 python bench.py --preset translucent --steps 40 --warmup 8 --no-train-iter --no-cpu-baseline > gpurun_out/final_bench_translucent.json 2> gpurun_out/final_tr.err; echo "translucent exit $?"; cut -c1-200 gpurun_out/final_bench_translucent.json
timeout 300 python tools/bench_config4.py > gpurun_out/final_c4_n1.json 2> gpurun_out/final_c4.err; echo "c4 exit $?"; cat gpurun_out/final_c4_n1.json
timeout 600 python tools/bench_configs.py > gpurun_out/final_configs.jsonl 2> gpurun_out/final_configs.err; echo "configs exit $?"; tail -2 gpurun_out/final_configs.jsonl | cut -c1-250
BENCH="python bench.py --steps 3 --warmup 3 --no-train-iter --no-cpu-baseline"
timeout 300 $BENCH > gpurun_out/ev_bench_plain.json 2> gpurun_out/ev_bench_plain.err || { echo "plain bench failed"; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 \
    --csv --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launches.log 2>&1
echo "launch list exit $?"; wc -l gpurun_out/launches.csv
CMD="python tools/one_view.py 3"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
tail -1 gpurun_out/plain.log
timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:'radix_scatter|radix_count|radix_colscan|render_bwd|render_fwd|emit_keys|preprocess_fwd|preprocess_bwd|scan_spine|compact_active|accumulate_packed|clear_accumulated|densify_stats' -s 30 -c 30 \
    -o gpurun_out/prof_r02_final $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture exit $?"; ls -la gpurun_out/*.ncu-rep 2>/dev/null
