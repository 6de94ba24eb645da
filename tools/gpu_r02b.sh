#!/bin/bash
# round 2, call B: full GPU suite + smoke + every BASELINE config's bench line on the final code
tag=${1:-r02d}
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity.txt
echo "===== pytest -m gpu ====="
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tee gpurun_out/${tag}_pytest.log | tail -6
echo "===== smoke ====="
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tee gpurun_out/${tag}_smoke.log | tail -3
echo "===== bench (default: small256) ====="
timeout 900 python bench.py 2> gpurun_out/${tag}_bench.err | tee gpurun_out/${tag}_bench_small256.json | cut -c1-300
tail -3 gpurun_out/${tag}_bench.err
for w in tiny1 base64 largev3; do
  echo "===== bench $w ====="
  timeout 600 python bench.py --workload $w --no-cpu-baseline 2> gpurun_out/${tag}_bench_$w.err | tee gpurun_out/${tag}_bench_$w.json | cut -c1-260
done
echo "===== bench frontend (configs[4]: 960 clips = 8 h of PCM per launch sequence, 50 back-to-back steps) ====="
timeout 600 python bench.py --workload frontend --steps 50 --no-cpu-baseline 2> gpurun_out/${tag}_bench_frontend.err | tee gpurun_out/${tag}_bench_frontend.json | cut -c1-260
echo "===== reference arm ====="
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2> gpurun_out/${tag}_bench_ref.err | tee gpurun_out/${tag}_bench_reference_arm.json | cut -c1-300
echo "===== logmelperf ====="
timeout 300 python tools/gpu_debug.py logmelperf 2>&1 | tee gpurun_out/${tag}_logmelperf.txt | tail -10
