#!/bin/bash
# The matcher / pre-processing part of tools/evidence.sh alone (the extraction kernels' captures stay valid when only
# csrc/matcher.cu or csrc/preproc.cu changed):  gpurun --timeout 1500 -- 'bash tools/evidence_matcher.sh f'
# then here:  python tools/ncu_traffic.py gpurun_out/r2_matcher_full_f.ncu-rep "<source line>" > profiles/r2_traffic_matcher.json
tag=${1:-x}
out=gpurun_out
mkdir -p $out
timeout 900 ncu --set full --clock-control none -k 'regex:k_search|k_fuse|k_knn|knn2|k_distinctive|k_mean_value|k_assign_grid|k_features_in_area|k_remap|k_resize_gray' -c 20 -f \
  -o $out/r2_matcher_full_$tag python tools/profile_step.py --steps 1 --knn 4096 --matcher > $out/ncu_matcher_$tag.log 2>&1; tail -1 $out/ncu_matcher_$tag.log
