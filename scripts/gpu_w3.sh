#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/quick.log
run() { echo "## $1" >> gpurun_out/quick.log; shift; timeout 300 "$@" >> gpurun_out/quick.log 2>&1; }
B="--no-e2e --no-cpu-baseline --no-also"
for W in 128 64 256; do
  L=$PWD/parallel_heat_b200/libheat_b200_t$W.so; [ $W -eq 128 ] && L=$PWD/parallel_heat_b200/libheat_b200.so
  run "c2 THREADS=$W" env HEAT_B200_LIB=$L python bench.py --steps 4000 --warmup 64 $B
  run "c3 THREADS=$W" env HEAT_B200_LIB=$L python bench.py --workload c3 --steps 400 --warmup 16 $B
done
run "c2 T=4 default" python bench.py --steps 4000 --warmup 64 $B --temporal-block 4
run "c3 T=4 default" python bench.py --workload c3 --steps 400 --warmup 16 $B --temporal-block 4
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
