#!/bin/bash
# Build an A/B copy of the library with another attention source / extra flags, next to the in-tree build.
# usage: tools/build_variant.sh <suffix> <attention source> [nvcc flags...]   -> yoho_b200/lib/libyoho_b200_<suffix>.so
set -e
sfx=$1; src=$2; shift 2
mkdir -p build/ab
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr "$@" -I include -I yoho_b200/csrc -c $src -o build/ab/attention_$sfx.o
objs=$(ls build/obj/*.o | grep -v attention_sm100)
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o yoho_b200/lib/libyoho_b200_$sfx.so $objs build/ab/attention_$sfx.o -cudart static -Xlinker --exclude-libs=ALL
echo yoho_b200/lib/libyoho_b200_$sfx.so
