#!/bin/bash
# ncu --set full capture of one kernel of the prof_case workload.  usage: tools/gpu_ncu_one.sh <tag> <kernel regex> <skip> <count> [prof_case args]
tag=$1; rx=$2; skip=$3; cnt=$4; shift 4
mkdir -p gpurun_out
CMD="python tools/prof_case.py $*"
$CMD > gpurun_out/${tag}_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/${tag}_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c $cnt -f -o gpurun_out/${tag} $CMD > gpurun_out/${tag}_ncu.log 2>&1
echo "ncu exit $?"; tail -2 gpurun_out/${tag}_ncu.log
