#!/bin/bash
# ncu launch list of the bench command itself (B200_PROFILING.md): plain run first, then the gpu__time_duration pass over the
# first 400 launches (= the first warm-up steps of the 256-clip workload); tools/launch_shares.py turns it into class shares
tag=${1:-r02c}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/${tag}_ll_bench_plain.json 2> gpurun_out/${tag}_ll_bench_plain.err || { echo "plain run failed"; tail -20 gpurun_out/${tag}_ll_bench_plain.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_bench.csv $CMD > gpurun_out/${tag}_ll_bench_ncu.log 2>&1
echo "ncu exit $?"
python tools/launch_shares.py gpurun_out/${tag}_launches_bench.csv gpurun_out/${tag}_ll_bench_plain.json | tee gpurun_out/${tag}_launch_shares.txt
