#!/bin/bash
# round 2, eight GPUs, final library: the driver's command at N=8 and N=4
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
: > gpurun_out/n8f.log; : > gpurun_out/n8f.err
run() { echo "## $1" >> gpurun_out/n8f.log; shift; timeout 300 "$@" >> gpurun_out/n8f.log 2>>gpurun_out/n8f.err; echo "rc=$?" >> gpurun_out/n8f.log; }
run "strong N=8 (driver command)" $TR --nproc-per-node 8 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5
run "strong N=4 (driver command)" $TR --nproc-per-node 4 --master-port 29512 bench.py --gpus 4 --steps 20 --warmup 5
grep -o '^## .*\|^rc=.*\|"value": [0-9.]*, "unit": "GLUP/s", "n_gpus"\|"bit_exact": [a-z]*' gpurun_out/n8f.log | head -30
tail -n 3 gpurun_out/n8f.err
