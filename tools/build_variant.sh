#!/bin/bash
# Build a variant of libyoho_b200.so for A/B timing: recompiles ONE source with extra -D flags and links it with the
# objects of the regular build (build/obj).  usage: tools/build_variant.sh <name> <source.cu> <flags...>
#   -> build/variants/libyoho_b200_<name>.so   (select it with YOHO_B200_LIB=<path>)
set -euo pipefail
name=$1; src=$2; shift 2
root=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p "$root/build/variants"
obj="$root/build/variants/${src}.${name}.o"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr "$@" \
  -c "$root/yoho_b200/csrc/$src" -o "$obj"
objs=()
for o in "$root"/build/obj/*.o; do
  if [ "$(basename "$o")" = "${src}.o" ]; then objs+=("$obj"); else objs+=("$o"); fi
done
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o "$root/build/variants/libyoho_b200_${name}.so" "${objs[@]}" \
  -cudart static -Xlinker --exclude-libs=ALL -ldl
echo "$root/build/variants/libyoho_b200_${name}.so"
