#!/bin/bash
# A/B of library builds on the GPU box: bash tools/ab_lib.sh "lib lib_x lib_y"   (directories under orb_slam3_v1.0_b200/)
# prints device-resident frames/s, ms per step and the stage times for each build (2 repeats, interleaved)
for rep in 1 2; do
  for n in $1; do
    ORBX_LIB_DIR=$PWD/orb_slam3_v1.0_b200/$n python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-matching 2>/dev/null | grep '^{"metric"' | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$n', round(d['value']), round(d['ms_per_step'],4), {k:v['ms'] for k,v in d['stages'].items()})"
  done
done
