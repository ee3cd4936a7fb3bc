#!/bin/bash
# round 2, two GPUs, final library: whole GPU suite, the cross-process tests again with the staged kernel forced, driver command
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -q -m gpu > gpurun_out/pytest_gpu_n2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_n2.log
tail -n 4 gpurun_out/pytest_gpu_n2.log
HEAT_K1_VARIANT=bulk timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_gpu_parity.py -q -m gpu > gpurun_out/pytest_bulk_n2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_bulk_n2.log
tail -n 4 gpurun_out/pytest_bulk_n2.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
: > gpurun_out/n2b.log
run() { echo "## $1" >> gpurun_out/n2b.log; shift; timeout 400 "$@" >> gpurun_out/n2b.log 2>gpurun_out/n2b.err; echo "rc=$?" >> gpurun_out/n2b.log; }
run "strong N=2 (driver command)" $TR --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5
run "strong N=2 K1 forced" env HEAT_K1_VARIANT=vec $TR --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --no-also
run "c3-like 16384 strips N=2 check" $TR --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --no-also --workload c3
grep -o '^## .*\|^rc=.*\|"value": [0-9.]*, "unit": "GLUP/s", "n_gpus"\|"bit_exact": [a-z]*' gpurun_out/n2b.log | head -40
tail -n 3 gpurun_out/n2b.err
