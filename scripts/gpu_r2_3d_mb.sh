#!/bin/bash
# one GPU: 3-D kernel, resident CTAs per SM forced by launch bounds (registers 78 / 64 / 48)
mkdir -p gpurun_out
P=$PWD/parallel_heat_b200
: > gpurun_out/g3d_mb.log
run() { echo "## $1" >> gpurun_out/g3d_mb.log; shift; timeout 200 "$@" 2>>gpurun_out/g3d.err | grep '^{' >> gpurun_out/g3d_mb.log; }
for v in base mb4 mb5; do
  lib=$P/libheat_b200.so; [ $v != base ] && lib=$P/libheat_b200_$v.so
  HEAT_B200_LIB=$lib timeout 300 python -m pytest tests/test_gpu_3d.py -q -m gpu 2>&1 | tail -n 1
  for a in fma f64mix; do
    run "3-D 640 $v $a" env HEAT_B200_LIB=$lib python bench.py --steps 20 --warmup 5 --only-3d 640 --reps 15 --arith $a
  done
  run "3-D 1024 $v fma" env HEAT_B200_LIB=$lib python bench.py --steps 20 --warmup 5 --only-3d 1024 --reps 7
done
grep -o '^## .*\|"value": [0-9.]*, "unit": "GLUP/s"\|"bit_exact": [a-z]*' gpurun_out/g3d_mb.log | paste - - -
