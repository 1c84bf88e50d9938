#!/bin/bash
# Checks patches/solver_batched.patch against the reference tree (build container only):
#   1. the unified diff applies cleanly to pristine copies of the files it touches;
#   2. the new translation unit it adds (src/optimization/csm_batch.cc) parses and
#      type-checks against the reference's REAL headers (slam_types.h, slam_util.h,
#      data_structures.h, kdtree.h, math_util.h, feature_extracter.h) and the product's
#      include/nautilus_csm.h, with stand-ins only for the un-vendored third parties
#      (Eigen, glog, ROS message types, the Ceres names data_structures.h mentions:
#      oracle/shim).
# solver.cc itself needs Ceres, ROS, Boost.Math and vector_maps to parse and is not
# compiled here; its hunks are thin calls into csm_batch.
set -e
REF=${1:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
TMP=$(mktemp -d)
trap 'rm -rf "$TMP"' EXIT
mkdir -p "$TMP/src" "$TMP/config"
cp -r "$REF/src/optimization" "$REF/src/util" "$REF/src/input" "$TMP/src/"
cp "$REF/config/default_config.lua" "$TMP/config/"
cp "$REF/CMakeLists.txt" "$TMP/"
(cd "$TMP" && patch -p1 --quiet < "$HERE/solver_batched.patch")
echo "patch applies"
g++ -std=c++17 -fsyntax-only \
  -I"$HERE/../oracle/shim" -I"$HERE/../include" \
  "$TMP/src/optimization/csm_batch.cc"
echo "csm_batch.cc type-checks against the reference headers"
