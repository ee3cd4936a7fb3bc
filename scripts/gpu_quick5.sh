#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
: > gpurun_out/quick.log
run() { echo "## $1" >> gpurun_out/quick.log; shift; timeout 300 "$@" >> gpurun_out/quick.log 2>&1; }
B="--no-e2e --no-cpu-baseline --no-also"
run "c1 graphs" python bench.py --workload c1 --steps 8000 --warmup 32 $B
run "c1 nographs" env HEAT_GRAPHS=0 python bench.py --workload c1 --steps 8000 --warmup 32 $B
run "c1 T=4 graphs" python bench.py --workload c1 --steps 8000 --warmup 64 --temporal-block 4 $B
run "c1 T=4 nographs" env HEAT_GRAPHS=0 python bench.py --workload c1 --steps 8000 --warmup 64 --temporal-block 4 $B
run "c2 graphs" python bench.py --steps 4000 --warmup 32 $B
run "c2 nographs" env HEAT_GRAPHS=0 python bench.py --steps 4000 --warmup 32 $B
run "c2 T=4 graphs" python bench.py --steps 4000 --warmup 64 --temporal-block 4 $B
run "c2 T=4 nographs" env HEAT_GRAPHS=0 python bench.py --steps 4000 --warmup 64 --temporal-block 4 $B
run "c2 full" python bench.py --no-cpu-baseline
