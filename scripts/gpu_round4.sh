#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
B="python bench.py --steps 2000 --warmup 20 --no-e2e --no-cpu-baseline"
: > gpurun_out/tb.log
for T in 2 3 4; do for rpt in 32 48 64 96 128 192; do
  echo "## c2 T=$T rpt_tb=$rpt" >> gpurun_out/tb.log; HEAT_ROWS_PER_TILE_TB=$rpt $B --temporal-block $T >> gpurun_out/tb.log 2>&1
done; done
for rpt in 64 128 256; do echo "## weak32k T=4 rpt_tb=$rpt" >> gpurun_out/tb.log; HEAT_ROWS_PER_TILE_TB=$rpt python bench.py --workload weak32k --steps 400 --warmup 8 --no-e2e --no-cpu-baseline >> gpurun_out/tb.log 2>&1; done
echo "## c2 full default" >> gpurun_out/tb.log; python bench.py >> gpurun_out/tb.log 2>&1
echo "## c2 full T=4" >> gpurun_out/tb.log; python bench.py --temporal-block 4 >> gpurun_out/tb.log 2>&1
grep -E "^##|value" gpurun_out/tb.log | cut -c1-160
timeout 900 python scripts/ref_baselines.py > gpurun_out/ref_baselines.json 2> gpurun_out/ref_baselines.log
tail -15 gpurun_out/ref_baselines.log
