#!/bin/bash
# On the GPU box: e2e sensitivity -- library variant x environment, bench only (no parity run).
cp orb_slam3_v1.0_b200/lib/liborbx.so /tmp/liborbx_shipped.so
for spec in "$@"; do
  IFS='|' read -r label var envs <<< "$spec"
  if [ "$var" != "-" ]; then cp orb_slam3_v1.0_b200/lib/variants/liborbx_$var.so orb_slam3_v1.0_b200/lib/liborbx.so; else cp /tmp/liborbx_shipped.so orb_slam3_v1.0_b200/lib/liborbx.so; fi
  echo "== $label"
  env $envs timeout 120 python bench.py --steps 40 --warmup 3 --no-cpu-baseline --no-matching 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['e2e']['value']), d['e2e'])"
done
cp /tmp/liborbx_shipped.so orb_slam3_v1.0_b200/lib/liborbx.so
