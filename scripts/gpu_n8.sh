#!/bin/bash
# 8-GPU call (charged 8x): keep it short
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 $TR --nproc-per-node $N --master-port 29501 tests/check_strips_torchrun.py > gpurun_out/check_strips.log 2>&1; echo "rc=$?" >> gpurun_out/check_strips.log
tail -5 gpurun_out/check_strips.log
: > gpurun_out/n8.log
run() { echo "## $1" >> gpurun_out/n8.log; shift; timeout 240 "$@" >> gpurun_out/n8.log 2>&1; }
run "c2 N=$N T=1" $TR --nproc-per-node $N --master-port 29502 bench.py --gpus $N --steps 4000 --warmup 20
run "c2 N=4 T=1" $TR --nproc-per-node 4 --master-port 29508 bench.py --gpus 4 --steps 4000 --warmup 20
run "strong N=$N T=1" $TR --nproc-per-node $N --master-port 29504 bench.py --gpus $N --workload strong --steps 400 --warmup 8
run "strong N=4 T=1" $TR --nproc-per-node 4 --master-port 29505 bench.py --gpus 4 --workload strong --steps 200 --warmup 8
run "strong N=$N T=4" $TR --nproc-per-node $N --master-port 29503 bench.py --gpus $N --workload strong --steps 400 --warmup 8 --temporal-block 4
run "weak32k N=$N" $TR --nproc-per-node $N --master-port 29507 bench.py --gpus $N --workload weak32k --steps 400 --warmup 8
grep -E "^##|value" gpurun_out/n8.log | cut -c1-150
