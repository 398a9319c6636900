#!/bin/bash
# One gpurun call's worth of round evidence, written under gpurun_out/ with tag $1:
# GPU test suite, bench lines (product + reference arm), ncu launch list of the bench command,
# --set full captures of one extraction+stereo step (with source), of the matcher kernels and of the 2-NN at the bench shape.
# Run ON the GPU box:  gpurun --timeout 2400 -- 'bash tools/evidence.sh a'
# The .ncu-rep files stay in gpurun_out/ (scratch); tools/evidence_derive.sh turns them into the json / csv / txt
# summaries that are committed under profiles/.
tag=${1:-x}
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/r2_pytest_$tag.log 2>&1; tail -1 $out/r2_pytest_$tag.log
# random configurations against the oracle, logs kept: extractor (sizes, features, scale, levels, thresholds, lapping; pyramid,
# blur, candidates, keypoints, descriptors) and every matcher entry point (sizes, radii, ratios, flags, masks)
timeout 900 python tools/fuzz_parity.py 150 21 > $out/r2_fuzz_extractor_$tag.log 2>&1; tail -1 $out/r2_fuzz_extractor_$tag.log
timeout 600 python tools/fuzz_matcher.py 800 22 > $out/r2_fuzz_matcher_$tag.log 2>&1; tail -1 $out/r2_fuzz_matcher_$tag.log
timeout 600 python bench.py 2> $out/bench_$tag.err | tail -1 > $out/r2_bench_$tag.json; cut -c1-300 $out/r2_bench_$tag.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2>> $out/bench_$tag.err | tail -1 > $out/r2_bench_${tag}_ref.json; cut -c1-200 $out/r2_bench_${tag}_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/r2_launches_$tag.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-matching > $out/ncu_launches_$tag.log 2>&1
# whole-batch kernels on one stream (ORBX_LANES=1 ORBX_FORK_BLUR=0): 14 extraction launches + 2 stereo launches per step
ORBX_LANES=1 ORBX_FORK_BLUR=0 ORBX_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -s 16 -c 16 -f -o $out/r2_step_full_$tag \
  python tools/profile_step.py --steps 2 --stereo > $out/ncu_full_$tag.log 2>&1; tail -1 $out/ncu_full_$tag.log
timeout 900 ncu --set full --clock-control none -k 'regex:k_search|k_fuse|k_knn|knn2|k_distinctive|k_mean_value|k_assign_grid|k_features_in_area|k_remap|k_resize_gray' -c 20 -f \
  -o $out/r2_matcher_full_$tag python tools/profile_step.py --steps 1 --knn 4096 --matcher > $out/ncu_matcher_$tag.log 2>&1; tail -1 $out/ncu_matcher_$tag.log
# the 2-NN kernel at a shape that fills the GPU (65536 queries x 1 M rows: 64 x 1 CTAs of 1024 queries)
timeout 900 ncu --set full --clock-control none -k 'regex:k_knn2' -c 2 -f -o $out/r2_knn_full_$tag python tools/profile_step.py --steps 1 --knn 65536 --knn-rows 1048576 > $out/ncu_knn_$tag.log 2>&1; tail -1 $out/ncu_knn_$tag.log
ls -la $out | grep "_$tag"
