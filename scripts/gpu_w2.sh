#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/quick.log
run() { echo "## $1" >> gpurun_out/quick.log; shift; timeout 300 "$@" >> gpurun_out/quick.log 2>&1; }
B="--no-e2e --no-cpu-baseline --no-also --temporal-block 4"
for W in 4 2 1; do
  L=$PWD/parallel_heat_b200/libheat_b200_w$W.so; [ $W -eq 4 ] && L=$PWD/parallel_heat_b200/libheat_b200.so
  run "c2 T=4 WARPS=$W" env HEAT_B200_LIB=$L python bench.py --steps 4000 --warmup 64 $B
  run "c2 T=4 WARPS=$W R=24" env HEAT_B200_LIB=$L HEAT_ROWS_PER_TILE_TB=24 python bench.py --steps 4000 --warmup 64 $B
  run "c2 T=4 WARPS=$W R=96" env HEAT_B200_LIB=$L HEAT_ROWS_PER_TILE_TB=96 python bench.py --steps 4000 --warmup 64 $B
  run "weak32k WARPS=$W" env HEAT_B200_LIB=$L python bench.py --workload weak32k --steps 400 --warmup 16 $B
done
