#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi.log
tail -n 4 gpurun_out/pytest_multi.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
: > gpurun_out/n2b.log
run() { echo "## $1" >> gpurun_out/n2b.log; shift; timeout 200 "$@" >> gpurun_out/n2b.log 2>&1; }
run "strong N=2 2x1 strips" $TR --nproc-per-node 2 --master-port 29504 bench.py --gpus 2 --workload strong --steps 100 --warmup 4 --no-e2e
run "strong N=2 1x2 blocks" $TR --nproc-per-node 2 --master-port 29505 bench.py --gpus 2 --workload strong --steps 100 --warmup 4 --no-e2e --cols 2
