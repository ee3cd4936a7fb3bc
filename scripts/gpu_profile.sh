#!/bin/bash
mkdir -p gpurun_out
python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
tail -c 400 gpurun_out/bench_default.json
CMD="python bench.py --steps 40 --warmup 4 --no-e2e --no-cpu-baseline --no-also"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:heat_step_vec -s 20 -c 4 -f -o gpurun_out/prof_k1 $CMD > gpurun_out/ncu_k1.log 2>&1
CMD4="python bench.py --steps 64 --warmup 16 --temporal-block 4 --no-e2e --no-cpu-baseline --no-also"
$CMD4 > gpurun_out/plain4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:heat_step_tb -s 6 -c 4 -f -o gpurun_out/prof_k4 $CMD4 > gpurun_out/ncu_k4.log 2>&1
ls -la gpurun_out/*.ncu-rep
