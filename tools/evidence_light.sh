#!/bin/bash
# The cheap part of tools/evidence.sh for a library whose kernels did not change since the last ncu captures:
# GPU test suite, fuzz logs, both bench arms.  gpurun --timeout 1800 -- 'bash tools/evidence_light.sh e'
tag=${1:-x}
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/r2_pytest_$tag.log 2>&1; tail -1 $out/r2_pytest_$tag.log
timeout 900 python tools/fuzz_parity.py 150 21 > $out/r2_fuzz_extractor_$tag.log 2>&1; tail -1 $out/r2_fuzz_extractor_$tag.log
timeout 600 python tools/fuzz_matcher.py 800 22 > $out/r2_fuzz_matcher_$tag.log 2>&1; tail -1 $out/r2_fuzz_matcher_$tag.log
timeout 600 python bench.py 2> $out/bench_$tag.err | tail -1 > $out/r2_bench_$tag.json; cut -c1-300 $out/r2_bench_$tag.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2>> $out/bench_$tag.err | tail -1 > $out/r2_bench_${tag}_ref.json; cut -c1-200 $out/r2_bench_${tag}_ref.json
