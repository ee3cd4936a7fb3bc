#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
: > gpurun_out/r5.log
run() { echo "## $1" >> gpurun_out/r5.log; shift; timeout 300 "$@" >> gpurun_out/r5.log 2>&1; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run "c2 N=1 T=1 4000" python bench.py --steps 4000 --warmup 20 --no-e2e --no-cpu-baseline
run "c2 N=1 T=1 rpt16" env HEAT_ROWS_PER_TILE=16 python bench.py --steps 4000 --warmup 20 --no-e2e --no-cpu-baseline
run "c2 N=1 T=1 rpt4" env HEAT_ROWS_PER_TILE=4 python bench.py --steps 4000 --warmup 20 --no-e2e --no-cpu-baseline
run "c2 N=1 T=1 rpt12" env HEAT_ROWS_PER_TILE=12 python bench.py --steps 4000 --warmup 20 --no-e2e --no-cpu-baseline
run "c3 N=1" python bench.py --workload c3 --steps 400 --warmup 20 --no-e2e --no-cpu-baseline
run "c2 N=2 T=1" $TR --nproc-per-node 2 --master-port 29511 bench.py --gpus 2 --steps 4000 --warmup 20
run "c2 N=2 T=4" $TR --nproc-per-node 2 --master-port 29512 bench.py --gpus 2 --steps 4000 --warmup 20 --temporal-block 4
run "c2 N=1 T=4" python bench.py --steps 4000 --warmup 20 --temporal-block 4 --no-e2e --no-cpu-baseline
run "c2 N=1 f64mix" python bench.py --steps 2000 --warmup 20 --arith f64mix --no-e2e --no-cpu-baseline
grep -E "^##|value" gpurun_out/r5.log | cut -c1-150
