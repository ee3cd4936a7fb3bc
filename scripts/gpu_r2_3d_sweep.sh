#!/bin/bash
# round 2, one GPU: 3-D kernel after the ring-row fix -- tests, then rows-per-thread x planes-per-tile sweep
mkdir -p gpurun_out
P=$PWD/parallel_heat_b200
timeout 900 python -m pytest tests/test_gpu_3d.py tests/test_gpu_f64.py -q -m gpu > gpurun_out/pytest_3d.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_3d.log
tail -n 6 gpurun_out/pytest_3d.log
: > gpurun_out/g3d_sweep.log
run() { echo "## $1" >> gpurun_out/g3d_sweep.log; shift; timeout 200 "$@" 2>>gpurun_out/g3d.err | grep '^{' >> gpurun_out/g3d_sweep.log; }
for v in base ty2 ty2w8 ty8; do
  lib=$P/libheat_b200.so; [ $v != base ] && lib=$P/libheat_b200_$v.so
  [ $v != base ] && { HEAT_B200_LIB=$lib timeout 300 python -m pytest tests/test_gpu_3d.py -q -m gpu 2>&1 | tail -n 1 >> gpurun_out/pytest_3d.log; }
  for ppt in 0 16 32 64; do
    run "3-D 640 $v planes=$ppt" env HEAT_B200_LIB=$lib HEAT_3D_PLANES=$ppt python bench.py --steps 20 --warmup 5 --only-3d 640 --reps 15
  done
  run "3-D 1024 $v planes=0" env HEAT_B200_LIB=$lib python bench.py --steps 20 --warmup 5 --only-3d 1024 --reps 7
  run "3-D 1024 $v planes=32" env HEAT_B200_LIB=$lib HEAT_3D_PLANES=32 python bench.py --steps 20 --warmup 5 --only-3d 1024 --reps 7
done
run "f64 grid leg parity" python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-also --workload g16k_f64
grep -o '^## .*\|"value": [0-9.]*, "unit": "GLUP/s"\|"bit_exact": [a-z]*' gpurun_out/g3d_sweep.log | paste - - -
tail -n 4 gpurun_out/pytest_3d.log
