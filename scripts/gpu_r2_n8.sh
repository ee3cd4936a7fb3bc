#!/bin/bash
# round 2, eight GPUs: the driver's bench command at N=8 and N=4, K3 evidence, 2-D blocks, cross-process parity
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
: > gpurun_out/n8.log; : > gpurun_out/n8.err
run() { echo "## $1" >> gpurun_out/n8.log; shift; timeout 300 "$@" >> gpurun_out/n8.log 2>>gpurun_out/n8.err; echo "rc=$?" >> gpurun_out/n8.log; }
run "strong N=8 (driver command)" $TR --nproc-per-node 8 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5
run "strong N=8 k3 profile" $TR --nproc-per-node 8 --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 5 --no-e2e --no-also --k3-profile
run "strong N=8 2x4 blocks" $TR --nproc-per-node 8 --master-port 29513 bench.py --gpus 8 --steps 20 --warmup 5 --no-e2e --no-also --cols 4
run "strong N=8 4x2 blocks" $TR --nproc-per-node 8 --master-port 29514 bench.py --gpus 8 --steps 20 --warmup 5 --no-e2e --no-also --cols 2
run "strong N=4 (driver command)" $TR --nproc-per-node 4 --master-port 29515 bench.py --gpus 4 --steps 20 --warmup 5
run "weak c2 N=8" $TR --nproc-per-node 8 --master-port 29516 bench.py --gpus 8 --steps 20 --warmup 5 --no-e2e --no-also --workload c2
timeout 600 python -m pytest tests/test_multi_gpu.py -x -q -m gpu -k "long_runs or blocks_one_process" > gpurun_out/pytest_multi8.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi8.log
tail -n 3 gpurun_out/pytest_multi8.log
grep -o '^## .*\|^rc=.*\|"value": [0-9.]*, "unit": "GLUP/s", "n_gpus"\|"bit_exact": [a-z]*' gpurun_out/n8.log | head -60
tail -n 5 gpurun_out/n8.err
nvidia-smi topo -m > gpurun_out/topo8.txt 2>&1
