#!/bin/bash
# Builds liborbx.so with -DORBX_CHECKS (device-side invariant traps, orbx_internal.h) into orb_slam3_v1.0_b200/lib_checked
# and runs the one-call-per-kernel workload and the GPU parity tests against it (ORBX_LIB_DIR selects the library).
# compute-sanitizer is closed on this GPU pool; this is the substitute evidence.  usage (GPU box): bash tools/checked_run.sh
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
make -s -j4 -C orb_slam3_v1.0_b200/csrc LIBDIR=../lib_checked EXTRA=-DORBX_CHECKS || exit 1
export ORBX_LIB_DIR=$PWD/orb_slam3_v1.0_b200/lib_checked
{
  echo "== library: $ORBX_LIB_DIR/liborbx.so (-DORBX_CHECKS), $(cuobjdump -sass $ORBX_LIB_DIR/liborbx.so | grep -c 'BPT.TRAP') trap sites in SASS"
  python tools/sanitize_run.py 2>&1 | tail -4
  python -m pytest tests/test_gpu_extractor.py tests/test_gpu_matcher.py tests/test_gpu_preproc.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -4
  echo "== ORBX_DCHECK failures printed: $(grep -c 'ORBX_DCHECK failed' gpurun_out/r2_checked_run.log 2>/dev/null || echo 0)"
} 2>&1 | tee gpurun_out/r2_checked_run.log
