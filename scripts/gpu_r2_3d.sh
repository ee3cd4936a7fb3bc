#!/bin/bash
# round 2, one GPU: the 3-D variant (tests, bench, ncu), the whole suite and the driver's two commands on the final library
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_3d.py -q -m gpu > gpurun_out/pytest_3d.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_3d.log
tail -n 15 gpurun_out/pytest_3d.log
: > gpurun_out/g3d.log
run() { echo "## $1" >> gpurun_out/g3d.log; shift; timeout 300 "$@" 2>>gpurun_out/g3d.err | grep '^{' >> gpurun_out/g3d.log; }
run "3-D 640 fma" python bench.py --steps 20 --warmup 5 --only-3d 640
run "3-D 1024 fma" python bench.py --steps 20 --warmup 5 --only-3d 1024
run "3-D 640 nofma" python bench.py --steps 20 --warmup 5 --only-3d 640 --arith nofma
run "3-D 640 f64mix" python bench.py --steps 20 --warmup 5 --only-3d 640 --arith f64mix
grep -o '^## .*\|"value": [0-9.]*, "unit": "GLUP/s"\|"bit_exact": [a-z]*' gpurun_out/g3d.log | paste - - -
timeout 1500 python -m pytest tests/ -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log; tail -n 2 gpurun_out/smoke.log
( time timeout 900 python bench.py --impl reference --steps 20 --warmup 5 ) > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"
( time timeout 900 python bench.py --steps 20 --warmup 5 ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
tail -n 4 gpurun_out/bench_default.err
Q="--steps 20 --warmup 5 --only-3d 640 --reps 7"
python bench.py $Q > gpurun_out/plain_3d.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:heat_step_3d -s 30 -c 2 -f -o gpurun_out/r02_k3d_640 python bench.py $Q > gpurun_out/ncu_3d.log 2>&1
ncu -i gpurun_out/r02_k3d_640.ncu-rep --page raw --csv > gpurun_out/r02_k3d_640_raw.csv 2>/dev/null
ncu -i gpurun_out/r02_k3d_640.ncu-rep --page details --csv > gpurun_out/r02_k3d_640_details.csv 2>/dev/null
ls -la gpurun_out/*3d*
