#!/bin/bash
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 150 $TR --nproc-per-node $N --master-port 29501 tests/check_strips_torchrun.py > gpurun_out/check_strips.log 2>&1; echo "rc=$?" >> gpurun_out/check_strips.log
tail -n 4 gpurun_out/check_strips.log
: > gpurun_out/n8c.log
run() { echo "## $1" >> gpurun_out/n8c.log; shift; timeout 150 "$@" >> gpurun_out/n8c.log 2>&1; }
run "strong N=$N 2x4 blocks" $TR --nproc-per-node $N --master-port 29506 bench.py --gpus $N --workload strong --steps 300 --warmup 8 --no-e2e --cols 4
run "strong N=$N 4x2 blocks" $TR --nproc-per-node $N --master-port 29505 bench.py --gpus $N --workload strong --steps 300 --warmup 8 --no-e2e --cols 2
