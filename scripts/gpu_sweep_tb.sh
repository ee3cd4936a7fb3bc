#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/sweep_tb.log
B="--no-e2e --no-cpu-baseline --no-also --temporal-block 4"
for R in 32 48 64 96 128 192 256 384 512; do
  echo "## c2 8192 R=$R" >> gpurun_out/sweep_tb.log; HEAT_ROWS_PER_TILE_TB=$R python bench.py --steps 2000 --warmup 20 $B >> gpurun_out/sweep_tb.log 2>&1
  echo "## c3 16384 R=$R" >> gpurun_out/sweep_tb.log; HEAT_ROWS_PER_TILE_TB=$R python bench.py --workload c3 --steps 400 --warmup 20 $B >> gpurun_out/sweep_tb.log 2>&1
  echo "## weak32k 32768 R=$R" >> gpurun_out/sweep_tb.log; HEAT_ROWS_PER_TILE_TB=$R python bench.py --workload weak32k --steps 200 --warmup 8 $B >> gpurun_out/sweep_tb.log 2>&1
done
for R in 128 256 512; do
  echo "## strong 65536 R=$R" >> gpurun_out/sweep_tb.log; HEAT_ROWS_PER_TILE_TB=$R python bench.py --workload strong --steps 60 --warmup 4 $B >> gpurun_out/sweep_tb.log 2>&1
done
grep -cE "value" gpurun_out/sweep_tb.log
