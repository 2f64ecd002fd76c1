#!/usr/bin/env bash
# TEST INFRASTRUCTURE — builds the real reference (CulDeVu/CUDAPathTracer) for HOST execution.
#
# Recipe: the reference's own source files are streamed through `sed` from where
# they lie under /root/reference into a throw-away directory, compiled by nvcc's host pass with FMA
# contraction off, and linked with oracle/ref_driver.inc (our C driver, appended to the same
# translation unit because the reference's headers define globals and non-inline functions).
# Output: oracle/_ref/libcupt_ref.so ONLY (git-ignored, travels to the GPU box as a binary).
# No reference source is copied into the repository; the patched copies live in $TMP and are removed.
#
# Patch list (each is mechanical; numbers are reference file:line):
#   1 kernel.cu:9   <windows.h> include dropped
#   2 kernel.cu:23  "triple.h" include dropped (broken WIP header)
#   3 kernel.cu:14  <nvml.h> include dropped; kernel.cu:555-790 (getGPUTemp, main) cut
#   4 __device__ -> __host__ __device__ on free functions
#     kernel.cu:44,56,60,78,101,112,217,417  BVH.h:45,51  modelLoader.h:49  color.h:46,50,55,63  camera.h:77
#   5 limits.h:3    __device__ const float MAX_FLOAT -> const float MAX_FLOAT (host-visible)
#   6 kernel.cu:163-215 trace_shared cut (__shared__/__syncthreads, dead code)
#   7 QUALIFIERS override so cuRAND's XORWOW/Philox device headers run on the host (ref_prelude.h)
#   8 kernel.cu:33  NUM_BOUNCES -> run-time int; `curandState` -> OracleRng (switchable XORWOW / keyed Philox)
#   9 kernel.cu:114 one counter increment added at the top of trace() (rays are counted, not nominal)
#  10 nvcc host pass, -ffp-contract=off, -O2
set -euo pipefail
REF=${REF_DIR:-/root/reference}
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="$HERE/_ref"
if [ ! -d "$REF" ]; then
  echo "build_ref.sh: $REF not present (GPU box?) — keeping prebuilt $OUT" >&2
  exit 0
fi
TMP="$(mktemp -d /tmp/cupt_ref.XXXXXX)"
trap 'rm -rf "$TMP"' EXIT
mkdir -p "$OUT"

# --- sanity: the line numbers this recipe relies on ---
chk() { # file line expected-substring
  sed -n "${2}p" "$REF/$1" | grep -qF -- "$3" || { echo "build_ref.sh: $1:$2 does not contain '$3'" >&2; exit 1; }
}
chk kernel.cu 9 'windows.h'
chk kernel.cu 14 'nvml.h'
chk kernel.cu 23 'triple.h'
chk kernel.cu 33 '#define NUM_BOUNCES'
chk kernel.cu 114 'uint32_t stack[MAX_BVH_DEPTH];'
chk kernel.cu 163 'trace_shared'
chk kernel.cu 215 '}'
chk kernel.cu 555 'getGPUTemp'
chk limits.h 3 'MAX_FLOAT'

# --- headers: only the __device__ -> __host__ __device__ widening (patch 4) and MAX_FLOAT (patch 5) ---
for h in vec3.h AABBvec3.h ivec3.h color.h camera.h BVH.h modelLoader.h SIUnits.h sphere.h tiny_obj_loader.h; do
  sed -E 's/^([[:space:]]*)__device__ /\1__host__ __device__ /' "$REF/$h" > "$TMP/$h"
done
sed -E 's/^__device__ const float MAX_FLOAT/const float MAX_FLOAT/' "$REF/limits.h" > "$TMP/limits.h"

# --- kernel.cu (patches 1,2,3,4,6,8,9) ---
sed -E \
  -e '555,$d' \
  -e '163,215d' \
  -e '114s/$/ ORACLE_COUNT_RAY();/' \
  -e '33s/.*/#define NUM_BOUNCES ORACLE_NUM_BOUNCES/' \
  -e '23d' -e '14d' -e '9d' \
  -e 's/^__device__ /__host__ __device__ /' \
  "$REF/kernel.cu" > "$TMP/kernel_patched.cu"

cat > "$TMP/ref_tu.cu" <<EOF
#include "$HERE/ref_prelude.h"
#include "kernel_patched.cu"
#include "$HERE/ref_driver.inc"
EOF

NVCC=${NVCC:-nvcc}
# tinyobj is plain C++03; built by g++ as the survey verified
g++ -O2 -fPIC -ffp-contract=off -c "$REF/tiny_obj_loader.cc" -o "$TMP/tiny_obj_loader.o"
"$NVCC" -O2 -std=c++17 -fmad=false -w -Xcompiler -fPIC,-O2,-ffp-contract=off,-pthread \
  -c "$TMP/ref_tu.cu" -o "$TMP/ref_tu.o"
"$NVCC" -shared -Xcompiler -pthread "$TMP/ref_tu.o" "$TMP/tiny_obj_loader.o" -o "$OUT/libcupt_ref.so"
echo "built $OUT/libcupt_ref.so"

# ---- legacy GPU baseline: the original kernel.cu rebuilt for sm_100a (SURVEY.md §8d "Legacy GPU baseline") ----
# Compile patches only: 1 (windows.h), 2 (triple.h), 3 (nvml.h, getGPUTemp/main cut), the ray counter hook (9) and
# the image size (kernel.cu:28-29: 512 -> LEGACY_IMAGE_DIM, default 1024 = BASELINE config 2).  Device code is the
# reference's: cuRAND XORWOW, NUM_BOUNCES 3, radianceAlongSingleStep, 512-thread blocks, default nvcc flags
# (FMA contraction on, as compile.bat:4 had it).  Output: oracle/_ref/libcupt_legacy.so (runs only on a GPU box).
chk kernel.cu 28 '#define IMAGE_WIDTH 512'
chk kernel.cu 29 '#define IMAGE_HEIGHT 512'
mkdir -p "$TMP/legacy"
for h in vec3.h AABBvec3.h ivec3.h color.h camera.h BVH.h modelLoader.h SIUnits.h sphere.h tiny_obj_loader.h limits.h; do
  cp "$REF/$h" "$TMP/legacy/$h"
done
sed -E \
  -e '555,$d' \
  -e '114s/$/ ORACLE_COUNT_RAY();/' \
  -e '28s/512/LEGACY_IMAGE_DIM/' -e '29s/512/LEGACY_IMAGE_DIM/' \
  -e '23d' -e '14d' -e '9d' \
  "$REF/kernel.cu" > "$TMP/legacy/kernel_legacy.cu"
cat > "$TMP/legacy/legacy_tu.cu" <<EOF2
#include "$HERE/legacy_prelude.h"
#include "kernel_legacy.cu"
#include "$HERE/legacy_driver.inc"
EOF2
"$NVCC" -std=c++17 -w -gencode arch=compute_100a,code=sm_100a -DLEGACY_IMAGE_DIM=${LEGACY_IMAGE_DIM:-1024} -Xptxas -v \
  -Xcompiler -fPIC -c "$TMP/legacy/legacy_tu.cu" -o "$TMP/legacy/legacy_tu.o" 2> "$OUT/legacy_ptxas.log" || { cat "$OUT/legacy_ptxas.log"; exit 1; }
"$NVCC" -shared -gencode arch=compute_100a,code=sm_100a "$TMP/legacy/legacy_tu.o" "$TMP/tiny_obj_loader.o" -o "$OUT/libcupt_legacy.so"
grep -A1 "drawPixel" "$OUT/legacy_ptxas.log" | grep -E "Used|stack" || true
echo "built $OUT/libcupt_legacy.so"
