#!/bin/bash
# A/B timing of attention variants built by tools/build_variant.sh.  usage: tools/ab_attn.sh <suffix> ...   ("" = in-tree build)
mkdir -p gpurun_out
for sfx in "$@"; do
  lib=$PWD/yoho_b200/lib/libyoho_b200${sfx:+_$sfx}.so
  echo "== ${sfx:-intree}"
  YOHO_B200_LIB=$lib timeout 200 python tools/gpu_debug.py attnperf 2>&1 | grep attnperf | tee -a gpurun_out/ab_attn.log
done
