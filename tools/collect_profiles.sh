#!/bin/bash
# run on the GPU box (1 GPU): everything profiles/ quotes for the committed kernel, into gpurun_out/r2/
cd "$(dirname "$0")/.."
o=gpurun_out/r2; mkdir -p $o
python bench.py > $o/bench_n1.json 2> $o/bench_n1.err
python bench.py --impl reference --steps 2 --warmup 1 > $o/reference_arm_n1.json 2> $o/reference_arm_n1.err
for wl in config3 config4 config5; do python bench.py --workload $wl --steps 3 --warmup 3 > $o/bench_${wl}_n1.json 2> $o/bench_${wl}_n1.err; done
python bench.py --split byte-range --steps 5 --warmup 3 > $o/bench_split_n1.json 2> $o/bench_split_n1.err
# launch list of the default command (after it exited 0 without ncu)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/launches.csv python bench.py --steps 3 --warmup 3 --no-extras > $o/launches_run.log 2>&1
# one steady-state launch of the dominant kernel, full set
ncu --set full --clock-control none --import-source on -k regex:ti_decode_kernel -s 4 -c 1 -o $o/full -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-extras > $o/full_run.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ti_decode_kernel -s 4 -c 1 -o $o/full_config3 -f python bench.py --workload config3 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-extras > $o/full_config3_run.log 2>&1
TI_LIB=$PWD/build/phase.so python tools/phase_timing.py config2 > $o/phase_config2.txt 2>&1
TI_LIB=$PWD/build/phase.so python tools/phase_timing.py config3 > $o/phase_config3.txt 2>&1
python tools/e2e_diag.py > $o/e2e_diag.txt 2>&1
ls -la $o
