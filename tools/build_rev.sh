#!/bin/bash
# Build the library as of git revision $1 into orb_slam3_v1.0_b200/$2 (for tools/ab2.sh A/B runs on the GPU box).
# usage: tools/build_rev.sh <rev> <libdir-name> [EXTRA nvcc flags]
set -e
cd "$(dirname "$0")/.."
T=$(mktemp -d)
git archive "$1" orb_slam3_v1.0_b200/csrc include tools/popc_peak.cu | tar -x -C "$T"
make -s -j4 -C "$T/orb_slam3_v1.0_b200/csrc" LIBDIR="$PWD/orb_slam3_v1.0_b200/$2" EXTRA="$3"
rm -rf "$T"
echo "built $1 -> orb_slam3_v1.0_b200/$2"
