#!/bin/bash
# Developer tool: build libcupt.so variants with extra -D flags into variants/ (git-ignored, travels with gpurun).
#   tools/build_variants.sh name1 "-DFOO=1" name2 "-DBAR=0 -DBAZ=2" ...
set -e
cd "$(dirname "$0")/../cudapathtracer_b200/csrc"
mkdir -p ../../variants
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  ( make -s BUILD=build_$name OUT=../../variants/libcupt_$name.so EXTRA="$flags" -j8 2>&1 | grep -E "error" || true
    grep -A2 "k_extendILi0ELb0ELb[01]" build_$name/context.ptxas.log | grep -E "registers|spill" | tr '\n' ' '; echo " <- $name" ) &
done
wait
