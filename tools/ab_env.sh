#!/bin/bash
# A/B of whole-step throughput between environment settings of the same build.  usage: tools/ab_env.sh "VAR=a" "VAR=b" ...
mkdir -p gpurun_out
for kv in "$@"; do
  echo "== ${kv:-default}"
  env $kv timeout 300 python bench.py --no-cpu-baseline --steps 6 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['clocks']['sm_mhz'], {k:(v['achieved'],v['ms_per_step']) for k,v in d['rooflines'].items() if k in ('gemm_tcgen05','attention_tcgen05','layernorm')})"
done
