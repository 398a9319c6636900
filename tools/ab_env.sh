#!/bin/bash
# A/B of an environment knob on the GPU box: bash tools/ab_env.sh VAR "0 1" [bench args]
# prints device-resident frames/s, ms per step and the stage times for each value (3 repeats each, interleaved)
var=$1; vals=$2; shift 2
for rep in 1 2 3; do
  for v in $vals; do
    env $var=$v python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-matching "$@" 2>/dev/null | grep '^{"metric"' | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$var=$v', round(d['value']), round(d['ms_per_step'],4), {k:v['ms'] for k,v in d['stages'].items()})"
  done
done
