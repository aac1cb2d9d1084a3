#!/bin/bash
# run on the GPU box with N GPUs: tools/scale_run.sh N  -> gpurun_out/r2/scale_nN_*.json
cd "$(dirname "$0")/.."
N=$1; o=gpurun_out/r2; mkdir -p $o
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > $o/scale_n${N}_config2.json 2> $o/scale_n${N}_config2.err
$TR --master-port 29512 bench.py --gpus $N --split byte-range --steps 5 --warmup 3 > $o/scale_n${N}_split.json 2> $o/scale_n${N}_split.err
$TR --master-port 29513 bench.py --gpus $N --workload config4 --steps 3 --warmup 3 > $o/scale_n${N}_config4.json 2> $o/scale_n${N}_config4.err
$TR --master-port 29514 bench.py --gpus $N --workload config5 --steps 1 --warmup 1 > $o/scale_n${N}_config5.json 2> $o/scale_n${N}_config5.err
$TR --master-port 29515 bench.py --gpus $N --impl reference --steps 1 --warmup 0 > $o/scale_n${N}_reference.json 2> $o/scale_n${N}_reference.err
for f in $o/scale_n${N}_*.err; do echo "== $f"; tail -c 300 $f; done
