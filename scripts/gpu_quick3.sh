#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/quick.log
run() { echo "## $1" >> gpurun_out/quick.log; shift; timeout 300 "$@" >> gpurun_out/quick.log 2>&1; }
B="--no-e2e --no-cpu-baseline --no-also"
run "c2 T=4 auto" python bench.py --steps 4000 --warmup 20 --temporal-block 4 $B
run "c2 T=4 R=48" env HEAT_ROWS_PER_TILE_TB=48 python bench.py --steps 4000 --warmup 20 --temporal-block 4 $B
run "c2 T=4 R=36" env HEAT_ROWS_PER_TILE_TB=36 python bench.py --steps 4000 --warmup 20 --temporal-block 4 $B
run "c2 T=4 R=72" env HEAT_ROWS_PER_TILE_TB=72 python bench.py --steps 4000 --warmup 20 --temporal-block 4 $B
run "c2 T=4 R=50" env HEAT_ROWS_PER_TILE_TB=50 python bench.py --steps 4000 --warmup 20 --temporal-block 4 $B
run "c2 T=2 auto" python bench.py --steps 4000 --warmup 20 --temporal-block 2 $B
run "c2 T=3 auto" python bench.py --steps 4000 --warmup 20 --temporal-block 3 $B
run "weak32k auto" python bench.py --workload weak32k --steps 400 --warmup 8 $B
run "c3 T=4 auto" python bench.py --workload c3 --temporal-block 4 --steps 400 --warmup 20 $B
run "strong N=1 T=4" python bench.py --workload strong --steps 100 --warmup 4 --temporal-block 4 $B
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
