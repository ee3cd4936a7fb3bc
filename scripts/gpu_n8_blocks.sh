#!/bin/bash
# 8-GPU call: 2-D blocks against row strips (strong scaling grid)
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
: > gpurun_out/n8b.log
run() { echo "## $1" >> gpurun_out/n8b.log; shift; timeout 200 "$@" >> gpurun_out/n8b.log 2>&1; }
run "strong N=$N 8x1 strips" $TR --nproc-per-node $N --master-port 29504 bench.py --gpus $N --workload strong --steps 300 --warmup 8 --no-e2e
run "strong N=$N 4x2 blocks" $TR --nproc-per-node $N --master-port 29505 bench.py --gpus $N --workload strong --steps 300 --warmup 8 --no-e2e --cols 2
run "strong N=$N 2x4 blocks" $TR --nproc-per-node $N --master-port 29506 bench.py --gpus $N --workload strong --steps 300 --warmup 8 --no-e2e --cols 4
run "c2 N=$N default with e2e" $TR --nproc-per-node $N --master-port 29507 bench.py --gpus $N --steps 4000 --warmup 20
grep -E "^##|Error" gpurun_out/n8b.log | cut -c1-160
