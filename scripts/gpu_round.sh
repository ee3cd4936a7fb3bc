#!/bin/bash
# one gpurun call: smoke, GPU tests, short bench runs. Outputs under gpurun_out/.
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
: > gpurun_out/quick.log
run() { echo "## $1" >> gpurun_out/quick.log; shift; timeout 300 "$@" >> gpurun_out/quick.log 2>&1; }
run "c2" python bench.py --steps 4000 --warmup 32 --no-cpu-baseline
run "c3" python bench.py --workload c3 --steps 400 --warmup 20 --no-e2e --no-cpu-baseline
run "c3 T=4" python bench.py --workload c3 --steps 400 --warmup 20 --no-e2e --no-cpu-baseline --temporal-block 4
run "c1" python bench.py --workload c1 --steps 8000 --warmup 32 --no-e2e --no-cpu-baseline
tail -n 3 gpurun_out/smoke.log; tail -n 3 gpurun_out/pytest_gpu.log
