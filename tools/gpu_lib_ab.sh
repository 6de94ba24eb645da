#!/bin/bash
# A/B of library variants (build/variants/libyoho_b200_<name>.so, tools/build_variant.sh) inside the benchmarked step:
#   tools/gpu_lib_ab.sh rounds name1 name2 ...     ("default" = the in-tree library)
rounds=$1; shift
mkdir -p gpurun_out
out=gpurun_out/lib_ab.txt
: > $out
for r in $(seq 1 $rounds); do
  for name in "$@"; do
    if [ "$name" = default ]; then unset YOHO_B200_LIB; else export YOHO_B200_LIB=$PWD/build/variants/libyoho_b200_$name.so; fi
    timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['rooflines']
print('$name round $r: ms/step', d['ms_per_step'], 'attention ms', r['attention_tcgen05']['ms_per_step'], 'gemm ms', r['gemm_tcgen05']['ms_per_step'], 'clock', d['clocks']['sm_mhz'], 'power', d['clocks'].get('power_w_max'))" >> $out
  done
done
cat $out
