#!/bin/bash
# From the .ncu-rep files of tools/evidence.sh (gpurun_out/, scratch) to the committed summaries under profiles/.
# usage (here, after gpurun merged gpurun_out/): bash tools/evidence_derive.sh <tag>
tag=${1:-x}
cd "$(dirname "$0")/.."
g=gpurun_out; p=profiles
cp $g/r2_pytest_$tag.log $g/r2_bench_$tag.json $g/r2_bench_${tag}_ref.json $g/r2_launches_$tag.csv $g/r2_fuzz_extractor_$tag.log $g/r2_fuzz_matcher_$tag.log $p/ 2>/dev/null
python tools/ncu_traffic.py $g/r2_step_full_$tag.ncu-rep "ORBX_LANES=1 ORBX_FORK_BLUR=0 ORBX_GRAPH=0 ncu --set full: one 256-frame extraction step + stereo (tools/evidence.sh $tag)" > $p/r2_traffic.json
python tools/ncu_traffic.py $g/r2_matcher_full_$tag.ncu-rep "ncu --set full: one call of every matcher kernel, 8 problems each (tools/evidence.sh $tag)" > $p/r2_traffic_matcher.json
python tools/ncu_traffic.py $g/r2_knn_full_$tag.ncu-rep "ncu --set full: k_knn2 65536 queries x 1 M rows (tools/evidence.sh $tag)" > $p/r2_traffic_knn.json
for k in k_fast_units k_resize_band k_blur7 k_orient_describe k_octree; do
  { echo "== $k: instructions / stall samples by source line (>= 0.8 %), then by phase between barriers"
    ncu -i $g/r2_step_full_$tag.ncu-rep --page source --csv --print-source cuda,sass --kernel-name regex:$k 2>/dev/null | python tools/ncu_line_summary.py 0.8
    ncu -i $g/r2_step_full_$tag.ncu-rep --page source --csv --kernel-name regex:$k 2>/dev/null | python tools/ncu_phase_summary.py
  } > $p/r2_lines_$k.txt
done
ls -la $p | grep r2_
