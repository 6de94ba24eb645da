#!/bin/bash
# A/B of whole-step throughput between library variants.  usage: [WORKLOAD=base64] tools/ab_bench.sh <suffix> ...  ("" = in-tree build)
mkdir -p gpurun_out
for sfx in "$@"; do
  lib=$PWD/yoho_b200/lib/libyoho_b200${sfx:+_$sfx}.so
  echo "== ${sfx:-intree}"
  YOHO_B200_LIB=$lib timeout 300 python bench.py ${WORKLOAD:+--workload $WORKLOAD} --no-cpu-baseline --steps 6 --warmup 3 2>/dev/null | tee gpurun_out/ab_bench_${sfx:-intree}.json | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['clocks']['sm_mhz'], {k:(v['achieved'],v['ms_per_step']) for k,v in d['rooflines'].items() if k in ('gemm_tcgen05','attention_tcgen05','layernorm')})"
done
