#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "resident" > gpurun_out/pytest_k5.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_k5.log
tail -n 15 gpurun_out/pytest_k5.log
: > gpurun_out/quick.log
run() { echo "## $1" >> gpurun_out/quick.log; shift; timeout 300 "$@" >> gpurun_out/quick.log 2>&1; }
B="--no-e2e --no-cpu-baseline --no-also"
for G in 1 2 4 6 8; do run "c1 resident G=$G" env HEAT_RESIDENT_G=$G python bench.py --workload c1 --steps 8000 --warmup 32 $B; done
run "c1 noresident" env HEAT_RESIDENT=0 python bench.py --workload c1 --steps 8000 --warmup 32 $B
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 6 gpurun_out/pytest_gpu.log
