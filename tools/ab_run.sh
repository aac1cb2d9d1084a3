#!/bin/bash
# run on the GPU box: kernel-only bench of the in-tree library and of every build/sweep/*.so, config2 and config3
# (TI_BENCH_NOCHECK=1: tuning builds may produce deliberately wrong output)
cd "$(dirname "$0")/.."
run() {
  for wl in ${WLS:-config2 config3}; do
    r=$(TI_BENCH_NOCHECK=1 TI_LIB=$1 timeout 300 python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('%.3f ms (best %.3f)  %.1f GB/s  frac %.3f' % (d['ms_per_step'], d['ms_best'], d['value'], d['roofline']['frac']))")
    echo "$(basename $1) $wl $r"
  done
}
run $PWD/accel_sim_framework_distributed_b200/csrc/libtraceingest.so
for f in build/sweep/*.so; do [ -e "$f" ] && run $PWD/$f; done
