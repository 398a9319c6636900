#!/bin/bash
# compute-sanitizer over tools/sanitize_run.py: memcheck, racecheck (shared-memory hazards of the warp queues /
# mbarrier + TMA staging), synccheck and initcheck.  Summaries -> gpurun_out/r2_sanitizer_<tool>.log
# usage (GPU box): bash tools/sanitize.sh [tools...]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TOOLS=${@:-memcheck racecheck synccheck initcheck}
for t in $TOOLS; do
    timeout ${SAN_TIMEOUT:-500} compute-sanitizer --tool $t --print-limit 20 python tools/sanitize_run.py > gpurun_out/r2_sanitizer_$t.log 2>&1
    echo "== $t rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/r2_sanitizer_$t.log | tail -1)"
done
