#!/bin/bash
# round 2, one GPU: last 3-D shape sweep, then what the driver runs at round end on the final library
mkdir -p gpurun_out
P=$PWD/parallel_heat_b200
: > gpurun_out/g3d_sweep2.log
run() { echo "## $1" >> gpurun_out/g3d_sweep2.log; shift; timeout 200 "$@" 2>>gpurun_out/g3d.err | grep '^{' >> gpurun_out/g3d_sweep2.log; }
for v in base t1w8 t1w16 t2w16; do
  lib=$P/libheat_b200.so; [ $v != base ] && lib=$P/libheat_b200_$v.so
  [ $v != base ] && { HEAT_B200_LIB=$lib timeout 300 python -m pytest tests/test_gpu_3d.py -q -m gpu 2>&1 | tail -n 1; }
  run "3-D 640 $v" env HEAT_B200_LIB=$lib python bench.py --steps 20 --warmup 5 --only-3d 640 --reps 15
  run "3-D 1024 $v" env HEAT_B200_LIB=$lib python bench.py --steps 20 --warmup 5 --only-3d 1024 --reps 7
done
grep -o '^## .*\|"value": [0-9.]*, "unit": "GLUP/s"\|"bit_exact": [a-z]*' gpurun_out/g3d_sweep2.log | paste - - -
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log; tail -n 2 gpurun_out/smoke.log
( time timeout 900 python bench.py --impl reference --steps 20 --warmup 5 ) > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"
( time HEAT_TIMING=1 timeout 900 python bench.py --steps 20 --warmup 5 ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
grep "heat_solve" gpurun_out/bench_default.err | tail -n 4
tail -n 4 gpurun_out/bench_default.err
Q="--steps 20 --warmup 5 --only-3d 640 --reps 7"
python bench.py $Q > gpurun_out/plain_3d.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:heat_step_3d -s 30 -c 2 -f -o gpurun_out/r02_k3d_640 python bench.py $Q > gpurun_out/ncu_3d.log 2>&1
ncu -i gpurun_out/r02_k3d_640.ncu-rep --page raw --csv > gpurun_out/r02_k3d_640_raw.csv 2>/dev/null
ncu -i gpurun_out/r02_k3d_640.ncu-rep --page details --csv > gpurun_out/r02_k3d_640_details.csv 2>/dev/null
rm -f gpurun_out/r02_k3d_640.ncu-rep
