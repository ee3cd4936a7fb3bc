#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
B="python bench.py --steps 1000 --warmup 20 --no-e2e --no-cpu-baseline"
: > gpurun_out/sweep.log
for serp in 1 0; do for rpt in 8 16 32 64 128; do
  echo "## serp=$serp rpt=$rpt" >> gpurun_out/sweep.log
  HEAT_SERPENTINE=$serp HEAT_ROWS_PER_TILE=$rpt $B >> gpurun_out/sweep.log 2>&1
done; done
for a in nofma f64mix; do echo "## arith=$a" >> gpurun_out/sweep.log; $B --arith $a >> gpurun_out/sweep.log 2>&1; done
echo "## c3 16384 conv20" >> gpurun_out/sweep.log; python bench.py --workload c3 --steps 400 --warmup 20 --no-e2e --no-cpu-baseline >> gpurun_out/sweep.log 2>&1
echo "## c1 1024" >> gpurun_out/sweep.log; python bench.py --workload c1 --steps 2000 --warmup 20 --no-e2e --no-cpu-baseline >> gpurun_out/sweep.log 2>&1
tail -3 gpurun_out/pytest_noalloc_alone.log gpurun_out/pytest_gpu.log
cat gpurun_out/ref_cuda_conv.txt
