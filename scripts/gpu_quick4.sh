#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
: > gpurun_out/quick.log
run() { echo "## $1" >> gpurun_out/quick.log; shift; timeout 300 "$@" >> gpurun_out/quick.log 2>&1; }
B="--no-e2e --no-cpu-baseline --no-also --temporal-block 4"
for R in 24 30 36 42 48 60; do run "c2 T=4 R=$R" env HEAT_ROWS_PER_TILE_TB=$R python bench.py --steps 4000 --warmup 20 $B; done
for R in 48 72 96 120; do run "c3 T=4 R=$R" env HEAT_ROWS_PER_TILE_TB=$R python bench.py --workload c3 --steps 400 --warmup 20 $B; done
for R in 96 144 192 240; do run "weak32k R=$R" env HEAT_ROWS_PER_TILE_TB=$R python bench.py --workload weak32k --steps 200 --warmup 8 $B; done
run "c2 T=4 auto" python bench.py --steps 4000 --warmup 20 $B
