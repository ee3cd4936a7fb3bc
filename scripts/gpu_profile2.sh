#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dropin_program.py -m gpu -x -q > gpurun_out/pytest_dropin.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_dropin.log
tail -5 gpurun_out/pytest_dropin.log
python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
tail -c 600 gpurun_out/bench_reference.json; tail -c 1500 gpurun_out/bench_default.json
CMD="python bench.py --steps 20 --warmup 4 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:heat_step -s 10 -c 24 -f -o gpurun_out/prof_final $CMD > gpurun_out/ncu_full.log 2>&1
CMD3="python bench.py --workload c3 --steps 40 --warmup 4 --no-e2e --no-cpu-baseline"
$CMD3 > gpurun_out/plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:heat_step_vec -s 18 -c 4 -f -o gpurun_out/prof_c3 $CMD3 > gpurun_out/ncu_c3.log 2>&1
ls -la gpurun_out/*.ncu-rep
