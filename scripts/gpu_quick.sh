#!/bin/bash
set -x
mkdir -p gpurun_out
: > gpurun_out/quick.log
echo "## c2 T=4 e2e timing" >> gpurun_out/quick.log; HEAT_TIMING=1 python bench.py --steps 4000 --warmup 20 --temporal-block 4 --no-cpu-baseline >> gpurun_out/quick.log 2>&1
echo "## c2 T=1 e2e timing" >> gpurun_out/quick.log; HEAT_TIMING=1 python bench.py --steps 4000 --warmup 20 --no-cpu-baseline >> gpurun_out/quick.log 2>&1
echo "## weak32k auto" >> gpurun_out/quick.log; python bench.py --workload weak32k --steps 400 --warmup 8 --no-e2e --no-cpu-baseline >> gpurun_out/quick.log 2>&1
echo "## c3 T=4 auto" >> gpurun_out/quick.log; python bench.py --workload c3 --temporal-block 4 --steps 400 --warmup 20 --no-e2e --no-cpu-baseline >> gpurun_out/quick.log 2>&1
echo "## c1 T=4" >> gpurun_out/quick.log; python bench.py --workload c1 --temporal-block 4 --steps 4000 --warmup 20 --no-e2e --no-cpu-baseline >> gpurun_out/quick.log 2>&1
grep -E "^##|value|heat_solve:" gpurun_out/quick.log | cut -c1-200
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
