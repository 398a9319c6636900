#!/bin/bash
# Build liborbx variants with extra -D flags (here, no GPU) and time each on a B200 in one gpurun call.
# usage: tools/variant_bench.sh "name1:-DRS_RPT=8" "name2:-DFOO=1 -DBAR=2" ...
# Each variant is checked with the extractor parity tests before it is timed.
set -e
cd "$(dirname "$0")/.."
mkdir -p orb_slam3_v1.0_b200/lib/variants
names=""
for spec in "$@"; do
  name="${spec%%:*}"; flags="${spec#*:}"
  (cd orb_slam3_v1.0_b200/csrc && /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false \
     -Xcompiler -fPIC $flags -shared -o ../lib/variants/liborbx_$name.so extractor.cu matcher.cu stereo.cu 2>/dev/null)
  names="$names $name"
done
/usr/local/graft/bin/gpurun --timeout 900 -- "for v in $names; do cp orb_slam3_v1.0_b200/lib/variants/liborbx_\$v.so orb_slam3_v1.0_b200/lib/liborbx.so; echo \"== \$v\"; timeout 200 python -m pytest tests/test_gpu_extractor.py -x -q 2>&1 | tail -1; timeout 120 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-matching 2>/dev/null | tail -1 | python -c \"import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), {k:v['ms'] for k,v in d['stages'].items()})\"; done" 2>&1 | grep -v "^\[gpurun\] sending"
