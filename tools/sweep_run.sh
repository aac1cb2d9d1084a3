#!/bin/bash
# run on the GPU box: bench every tuning build, kernel-only
cd "$(dirname "$0")/.."
for f in build/sweep/*.so; do
  r=$(TI_LIB=$PWD/$f timeout 200 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-e2e --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('%.2f ms  %.1f GB/s' % (d['ms_per_step'], d['value']))")
  echo "$(basename $f) $r"
done
