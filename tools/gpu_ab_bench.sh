#!/bin/bash
# A/B of an environment switch inside the benchmarked step: tools/gpu_ab_bench.sh VAR [rounds]   (alternates VAR=0 / VAR=1)
var=$1; rounds=${2:-2}
mkdir -p gpurun_out
out=gpurun_out/ab_$var.txt
: > $out
for r in $(seq 1 $rounds); do
  for v in 0 1; do
    env $var=$v timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/ab_${var}_${v}_$r.json 2>/dev/null
    python - <<PY >> $out
import json
d = json.loads(open("gpurun_out/ab_${var}_${v}_$r.json").read().strip().splitlines()[-1])
r = d["rooflines"]
print("$var=$v round $r: value", d["value"], "ms/step", d["ms_per_step"], "e2e", d["e2e"]["value"], "attention ms", r["attention_tcgen05"]["ms_per_step"], "gemm ms", r["gemm_tcgen05"]["ms_per_step"], "clock", d["clocks"]["sm_mhz"])
PY
  done
done
cat $out
