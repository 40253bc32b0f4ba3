#!/bin/bash
# compute-sanitizer over every kernel of libb200splat.so at small sizes.  ONE tool per call
# (B200_PROFILING.md: several sanitizer tools in one GPU session have wedged the GPU on this pool):
#   bash tools/sanitize.sh memcheck | racecheck | initcheck | synccheck
# racecheck covers the shared-memory hazards (cp.async staging, strip bitmaps, radix ranking); initcheck the reads of
# never-written global memory (scratch buffers, the opaque state).  Kernels of other libraries are not instrumented.
# Output: gpurun_out/sanitize_<tool>.txt; the summaries are copied to profiles/ by hand.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
tool=${1:-memcheck}
timeout 1200 compute-sanitizer --tool $tool --kernel-regex kns=b200splat --print-limit 40 \
    --log-file gpurun_out/sanitize_$tool.txt python tools/sanitize_case.py > gpurun_out/sanitize_$tool.stdout 2>&1
echo "== $tool rc=$? $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/sanitize_$tool.txt | tail -1)"
