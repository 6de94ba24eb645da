#!/bin/bash
# One GPU-box pass: parity tests, smoke, the headline bench and the ncu launch list of a short bench run.
# usage: tools/gpu_round.sh [tag]      logs -> gpurun_out/<tag>_*.log
tag=${1:-r01}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv > gpurun_out/${tag}_gpu.txt 2>&1
echo "===== pytest -m gpu ====="
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tee gpurun_out/${tag}_pytest.log | tail -15
echo "===== smoke ====="
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tee gpurun_out/${tag}_smoke.log | tail -5
echo "===== bench (default) ====="
timeout 900 python bench.py 2> gpurun_out/${tag}_bench.err | tee gpurun_out/${tag}_bench.json | tail -3
tail -5 gpurun_out/${tag}_bench.err
echo "===== bench base64 ====="
timeout 600 python bench.py --workload base64 --no-cpu-baseline 2> gpurun_out/${tag}_bench_base64.err | tee gpurun_out/${tag}_bench_base64.json | tail -3
echo "===== bench frontend ====="
timeout 600 python bench.py --workload frontend --no-cpu-baseline 2> gpurun_out/${tag}_bench_frontend.err | tee gpurun_out/${tag}_bench_frontend.json | tail -3
echo "===== ncu launch list ====="
CMD="python bench.py --workload base64 --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu.log 2>&1
echo "ncu exit $?"
tail -3 gpurun_out/${tag}_ncu.log
