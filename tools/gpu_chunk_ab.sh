for mb in 48 24 96 100000; do
  echo "== YOHO_LOGMEL_CHUNK_MB=$mb"
  YOHO_LOGMEL_CHUNK_MB=$mb timeout 200 python tools/gpu_debug.py logmelperf 2>&1 | grep "B=256 N=480000 M=80 f32 norm=reference"
  YOHO_LOGMEL_CHUNK_MB=$mb timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['rooflines']
print('bench ms', d['ms_per_step'], 'logmel', r['logmel']['ms_per_step'], r['logmel']['frac'], r['logmel']['launches_per_step'], 'normalize', r['normalize']['ms_per_step'], r['normalize']['frac'])"
done
