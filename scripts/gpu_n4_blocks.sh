#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi.log
tail -n 6 gpurun_out/pytest_multi.log
N=$(nvidia-smi -L | wc -l)
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
: > gpurun_out/n4.log
run() { echo "## $1" >> gpurun_out/n4.log; shift; timeout 240 "$@" >> gpurun_out/n4.log 2>&1; }
run "strong N=$N strips" $TR --nproc-per-node $N --master-port 29504 bench.py --gpus $N --workload strong --steps 200 --warmup 8 --no-e2e
run "strong N=$N blocks cols=2" $TR --nproc-per-node $N --master-port 29505 bench.py --gpus $N --workload strong --steps 200 --warmup 8 --no-e2e --cols 2
run "c2 N=$N blocks cols=2" $TR --nproc-per-node $N --master-port 29506 bench.py --gpus $N --steps 2000 --warmup 20 --no-e2e --cols 2
run "c2 N=$N strips" $TR --nproc-per-node $N --master-port 29507 bench.py --gpus $N --steps 2000 --warmup 20 --no-e2e
grep -E "^##|Error|error" gpurun_out/n4.log | cut -c1-200
