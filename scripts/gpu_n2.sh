#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
: > gpurun_out/n2.log
run() { echo "## $1" >> gpurun_out/n2.log; shift; timeout 400 "$@" >> gpurun_out/n2.log 2>&1; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run "c2 N=2 default" $TR --nproc-per-node 2 --master-port 29511 bench.py --gpus 2
run "c2 N=2 reference arm" $TR --nproc-per-node 2 --master-port 29512 bench.py --gpus 2 --impl reference --steps 200
run "c2 N=2 T=4" $TR --nproc-per-node 2 --master-port 29513 bench.py --gpus 2 --temporal-block 4
tail -c 1500 gpurun_out/n2.log
