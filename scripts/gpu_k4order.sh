#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/quick.log
run() { echo "## $1" >> gpurun_out/quick.log; shift; timeout 300 "$@" >> gpurun_out/quick.log 2>&1; }
B="--no-e2e --no-cpu-baseline --no-also --temporal-block 4"
run "c2 T=4" python bench.py --steps 4000 --warmup 64 $B
run "c2 T=4 R=24" env HEAT_ROWS_PER_TILE_TB=24 python bench.py --steps 4000 --warmup 64 $B
run "c2 T=4 R=72" env HEAT_ROWS_PER_TILE_TB=72 python bench.py --steps 4000 --warmup 64 $B
run "c2 T=4 R=96" env HEAT_ROWS_PER_TILE_TB=96 python bench.py --steps 4000 --warmup 64 $B
run "c3 T=4" python bench.py --workload c3 --steps 400 --warmup 16 $B
run "weak32k" python bench.py --workload weak32k --steps 400 --warmup 16 $B
run "c2 T=2" python bench.py --steps 4000 --warmup 64 --no-e2e --no-cpu-baseline --no-also --temporal-block 2
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
