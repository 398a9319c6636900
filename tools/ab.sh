#!/bin/bash
# On the GPU box: time library variants / environment settings back to back.
# usage: tools/ab.sh "label|variant-so-name|ENV=.. ENV2=.." ...   (variant "-" = lib/liborbx.so as shipped)
cp orb_slam3_v1.0_b200/lib/liborbx.so /tmp/liborbx_shipped.so
for spec in "$@"; do
  IFS='|' read -r label var envs <<< "$spec"
  if [ "$var" != "-" ]; then cp orb_slam3_v1.0_b200/lib/variants/liborbx_$var.so orb_slam3_v1.0_b200/lib/liborbx.so; else cp /tmp/liborbx_shipped.so orb_slam3_v1.0_b200/lib/liborbx.so; fi
  echo "== $label"
  env $envs timeout 300 python -m pytest tests/test_gpu_extractor.py -x -q 2>&1 | tail -1
  env $envs timeout 120 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-matching 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['e2e']['value']), {k:v['ms'] for k,v in d['stages'].items()})"
done
cp /tmp/liborbx_shipped.so orb_slam3_v1.0_b200/lib/liborbx.so
