#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --steps 20 --warmup 4 --arith f64mix --no-e2e --no-cpu-baseline --no-also"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:heat_step_vec -s 10 -c 2 -f -o gpurun_out/prof_f64 $CMD > gpurun_out/ncu_f64.log 2>&1
ls -la gpurun_out/prof_f64.ncu-rep
