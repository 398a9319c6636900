#!/bin/bash
# On the GPU box: back-to-back A/B of library variants / tuning variables with tools/stage_bench.py.
# usage: tools/ab2.sh "label|libdir-or-minus|ENV=.. ENV2=.." ...     (libdir relative to orb_slam3_v1.0_b200/, "-" = lib)
cd "$(dirname "$0")/.."
for spec in "$@"; do
  IFS='|' read -r label dir envs <<< "$spec"
  if [ "$dir" != "-" ]; then export ORBX_LIB_DIR=$PWD/orb_slam3_v1.0_b200/$dir; else unset ORBX_LIB_DIR; fi
  env $envs timeout 300 python tools/stage_bench.py 20 "$label" 2>&1 | tail -1
done
