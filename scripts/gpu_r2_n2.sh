#!/bin/bash
# round 2, two GPUs: the cross-process tests the 1-GPU lease skips, then the driver's bench command at N=2
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_multi.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi.log
tail -n 5 gpurun_out/pytest_multi.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
: > gpurun_out/n2.log
run() { echo "## $1" >> gpurun_out/n2.log; shift; timeout 400 "$@" >> gpurun_out/n2.log 2>gpurun_out/n2.err; echo "rc=$?" >> gpurun_out/n2.log; }
run "strong N=2 (driver command)" $TR --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5
run "strong N=2 k3 profile" $TR --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --k3-profile
run "strong N=2 1x2 blocks" $TR --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --cols 2
run "weak c2 N=2 strips" $TR --master-port 29514 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --workload c2
run "weak c2 N=2 1x2 blocks" $TR --master-port 29515 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --workload c2 --cols 2
run "weak32k N=2" $TR --master-port 29516 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --workload weak32k
run "strong N=2 sendrecv" $TR --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --halo sendrecv
grep -o '"value": [0-9.]*\|^## .*\|^rc=.*\|"bit_exact": [a-z]*' gpurun_out/n2.log | head -80
tail -n 5 gpurun_out/n2.err
