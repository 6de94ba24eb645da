#!/bin/bash
# round 2, call A: full parity suite (new direct-oracle tests), default bench, single-clip graph bench
tag=r02a
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity.txt
echo "===== pytest -m gpu ====="
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tee gpurun_out/${tag}_pytest.log | tail -15
echo "===== smoke ====="
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tee gpurun_out/${tag}_smoke.log | tail -5
echo "===== bench (default) ====="
timeout 900 python bench.py 2> gpurun_out/${tag}_bench.err | tee gpurun_out/${tag}_bench.json | cut -c1-1500
tail -5 gpurun_out/${tag}_bench.err
echo "===== bench tiny1 (graph) ====="
timeout 300 python bench.py --workload tiny1 --no-cpu-baseline 2> gpurun_out/${tag}_bench_tiny1.err | tee gpurun_out/${tag}_bench_tiny1.json | cut -c1-1200
tail -3 gpurun_out/${tag}_bench_tiny1.err
echo "===== bench tiny1 (no graph) ====="
timeout 300 python bench.py --workload tiny1 --no-cpu-baseline --no-graph 2>> gpurun_out/${tag}_bench_tiny1.err | tee gpurun_out/${tag}_bench_tiny1_nograph.json | cut -c1-400
echo "===== logmelperf ====="
timeout 300 python tools/gpu_debug.py logmelperf 2>&1 | tee gpurun_out/${tag}_logmelperf.txt | tail -12
