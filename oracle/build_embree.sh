#!/usr/bin/env bash
# Build the reference's vendored Embree 3.8.0 (ext/embree) out-of-tree into oracle/_ref/.
#
# TEST INFRASTRUCTURE ONLY. The product never links or loads this library; it is the
# reference's own CPU BVH traversal (src/librender/scene_embree.inl:54-67,272-286 call
# rtcIntersect1 / rtcOccluded1 on it), used (a) to cross-check the oracle's first hits and
# (b) as the timed CPU baseline ("cpu_baseline.kind": "reference" for the traversal part).
#
# Options mirror /root/reference/ext/CMakeLists.txt:41-73 (no ISPC, no tutorials, internal
# tasking, AVX2 max ISA).  Sources are compiled where they lie under /root/reference; nothing
# is copied into this repository.  Outputs: oracle/_ref/libembree3.so* (git-ignored).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${BF_REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/_ref"
BUILD="${BF_EMBREE_BUILD_DIR:-/tmp/bf_embree_build}"
if [ -f "$OUT/libembree3.so" ]; then echo "embree already built: $OUT/libembree3.so"; exit 0; fi
if [ ! -d "$REF/ext/embree" ]; then echo "reference not present ($REF); skipping embree"; exit 0; fi
mkdir -p "$OUT" "$BUILD"
cmake -S "$REF/ext/embree" -B "$BUILD" -G Ninja \
  -DCMAKE_BUILD_TYPE=Release \
  -DEMBREE_ISPC_SUPPORT=OFF -DEMBREE_TUTORIALS=OFF -DEMBREE_TASKING_SYSTEM=INTERNAL \
  -DEMBREE_MAX_ISA=AVX2 -DEMBREE_STATIC_LIB=OFF -DEMBREE_GEOMETRY_QUAD=OFF \
  -DEMBREE_GEOMETRY_CURVE=OFF -DEMBREE_GEOMETRY_SUBDIVISION=OFF -DEMBREE_GEOMETRY_INSTANCE=OFF \
  -DEMBREE_GEOMETRY_USER=OFF -DEMBREE_GEOMETRY_POINT=OFF \
  -DCMAKE_POLICY_VERSION_MINIMUM=3.5 > "$BUILD/configure.log" 2>&1
ninja -C "$BUILD" embree > "$BUILD/build.log" 2>&1
cp -L "$BUILD"/libembree3.so "$OUT/libembree3.so"
echo "built $OUT/libembree3.so"
