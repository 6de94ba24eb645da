#!/bin/bash
# Build an A/B copy of the library in which one translation unit is replaced by another source and/or compiled with
# extra flags.  usage: tools/build_variant.sh <suffix> <unit: attention_sm100|gemm_sm100|...> <source> [nvcc flags...]
#   -> yoho_b200/lib/libyoho_b200_<suffix>.so   (select it with YOHO_B200_LIB; the in-tree build must exist: build/obj)
set -e
sfx=$1; unit=$2; src=$3; shift 3
mkdir -p build/ab
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr "$@" -I include -I yoho_b200/csrc -c $src -o build/ab/${unit}_$sfx.o
objs=$(ls build/obj/*.o | grep -v "/$unit\.")
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o yoho_b200/lib/libyoho_b200_$sfx.so $objs build/ab/${unit}_$sfx.o -cudart static -Xlinker --exclude-libs=ALL
echo yoho_b200/lib/libyoho_b200_$sfx.so
