#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
B="python bench.py --steps 2000 --warmup 20 --no-e2e --no-cpu-baseline"
: > gpurun_out/tb.log
for T in 1 2 3 4; do for rpt in 32 64 128 256; do
  echo "## c2 T=$T rpt_tb=$rpt" >> gpurun_out/tb.log; HEAT_ROWS_PER_TILE_TB=$rpt $B --temporal-block $T >> gpurun_out/tb.log 2>&1
  [ $T -eq 1 ] && break
done; done
for T in 2 4; do echo "## c2 f64mix T=$T" >> gpurun_out/tb.log; $B --temporal-block $T --arith f64mix >> gpurun_out/tb.log 2>&1; done
echo "## weak32k T=4" >> gpurun_out/tb.log; python bench.py --workload weak32k --steps 400 --warmup 8 --no-e2e --no-cpu-baseline >> gpurun_out/tb.log 2>&1
echo "## c3 T=4" >> gpurun_out/tb.log; python bench.py --workload c3 --steps 400 --warmup 20 --temporal-block 4 --no-e2e --no-cpu-baseline >> gpurun_out/tb.log 2>&1
echo "## strong N=1" >> gpurun_out/tb.log; python bench.py --workload strong --steps 100 --warmup 4 --no-e2e --no-cpu-baseline >> gpurun_out/tb.log 2>&1
grep -E "^##|value" gpurun_out/tb.log | cut -c1-160
