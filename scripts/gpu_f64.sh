#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/quick.log
run() { echo "## $1" >> gpurun_out/quick.log; shift; timeout 300 "$@" >> gpurun_out/quick.log 2>&1; }
B="--no-e2e --no-cpu-baseline --no-also --arith f64mix"
run "c2 f64mix" python bench.py --steps 4000 --warmup 32 $B
run "c3 f64mix" python bench.py --workload c3 --steps 400 --warmup 16 $B
run "c2 f64mix T=4" python bench.py --steps 4000 --warmup 32 $B --temporal-block 4
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
