#!/bin/bash
# round 2, one GPU, second A/B: product library with its new defaults (bulk-staged K1 for long rows, packed K4, integer
# widening in MPI_F64MIX, double grids) -- parity first, then the open choices, then ncu (CSV exported on the box)
mkdir -p gpurun_out
P=$PWD/parallel_heat_b200
B="--steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-also"
: > gpurun_out/ab2.log; : > gpurun_out/ab2.err; : > gpurun_out/ab2_parity.log
run() { echo "## $1" >> gpurun_out/ab2.log; shift; timeout 300 "$@" 2>>gpurun_out/ab2.err | grep '^{' >> gpurun_out/ab2.log; }
par() { echo "## parity $1" >> gpurun_out/ab2_parity.log; shift; timeout 900 "$@" >> gpurun_out/ab2_parity.log 2>&1; echo "rc=$?" >> gpurun_out/ab2_parity.log; }
par "whole suite, product library" python -m pytest tests -q -m gpu
par "bulk-staged K1 forced everywhere" env HEAT_K1_VARIANT=bulk python -m pytest tests/test_gpu_parity.py tests/test_bounds_debug.py -q -m gpu
par "register-streaming K1 forced everywhere" env HEAT_K1_VARIANT=vec python -m pytest tests/test_gpu_parity.py -q -m gpu -k "8192 or 16384 or 65536"
par "f64mix w1b (integer 2u)" env HEAT_B200_LIB=$P/libheat_b200_w1b.so python -m pytest tests/test_gpu_parity.py -q -m gpu
par "bulk rows 16" env HEAT_K1_VARIANT=bulk HEAT_B200_LIB=$P/libheat_b200_rows16.so python -m pytest tests/test_gpu_parity.py -q -m gpu -k "bit_exact or ring or tiling or 8192 or random or strip or conver"
par "bulk rows 4" env HEAT_K1_VARIANT=bulk HEAT_B200_LIB=$P/libheat_b200_rows4.so python -m pytest tests/test_gpu_parity.py -q -m gpu -k "bit_exact or ring or tiling or 8192 or random or strip or conver"
for v in w0 base w1b; do
  lib=$P/libheat_b200.so; [ $v != base ] && lib=$P/libheat_b200_$v.so
  for w in c2 g16k strong weak32k; do run "f64mix $w $v" env HEAT_B200_LIB=$lib HEAT_K1_VARIANT=vec python bench.py $B --workload $w --arith f64mix; done
done
run "f64mix g16k base bulk" env HEAT_K1_VARIANT=bulk python bench.py $B --workload g16k --arith f64mix
run "f64mix strong base bulk" env HEAT_K1_VARIANT=bulk python bench.py $B --workload strong --arith f64mix
for w in c2 g16k strong; do
  run "fma $w vec" env HEAT_K1_VARIANT=vec python bench.py $B --workload $w
  run "fma $w bulk rows 8" env HEAT_K1_VARIANT=bulk python bench.py $B --workload $w
  run "fma $w bulk rows 16" env HEAT_K1_VARIANT=bulk HEAT_B200_LIB=$P/libheat_b200_rows16.so python bench.py $B --workload $w
  run "fma $w bulk rows 4" env HEAT_K1_VARIANT=bulk HEAT_B200_LIB=$P/libheat_b200_rows4.so python bench.py $B --workload $w
done
run "fma c3 vec" env HEAT_K1_VARIANT=vec python bench.py $B --workload c3
run "fma c3 bulk" env HEAT_K1_VARIANT=bulk python bench.py $B --workload c3
run "fma weak32k (packed K4)" python bench.py $B --workload weak32k
run "f64 grid 16384" python bench.py $B --workload g16k_f64
run "f64 grid 16384 nofma" python bench.py $B --workload g16k_f64 --arith nofma
grep -o '^## .*\|"value": [0-9.]*, "unit": "GLUP/s", "n_gpus"' gpurun_out/ab2.log | paste - - | sed 's/"unit.*//'
grep "^##\|passed\|failed\|rc=" gpurun_out/ab2_parity.log
# --- ncu
Q="--steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-also --no-parity --reps 7"
cap() { # name, kernel regex, skip, count, then the command
  local name=$1 k=$2 s=$3 c=$4; shift 4
  "$@" > gpurun_out/plain_$name.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$k -s $s -c $c -f -o gpurun_out/r02_$name "$@" > gpurun_out/ncu_$name.log 2>&1
  ncu -i gpurun_out/r02_$name.ncu-rep --page raw --csv > gpurun_out/r02_${name}_raw.csv 2>/dev/null
  ncu -i gpurun_out/r02_$name.ncu-rep --page details --csv > gpurun_out/r02_${name}_details.csv 2>/dev/null
}
python bench.py $Q > gpurun_out/plain_launches.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_strong_launches.csv python bench.py $Q > gpurun_out/ncu_launches.log 2>&1
cap k1_bulk_strong heat_step_bulk 30 2 python bench.py $Q
cap k1_vec_strong heat_step_vec 30 2 env HEAT_K1_VARIANT=vec python bench.py $Q
cap k1k2_c3 "heat_step_(vec|bulk)" 24 3 python bench.py $Q --workload c3
cap k1_c2 heat_step_vec 30 2 python bench.py $Q --workload c2
cap k1_f64mix_c2 heat_step_vec 30 2 python bench.py $Q --workload c2 --arith f64mix
cap k4_weak32k heat_step_tb 8 2 python bench.py $Q --workload weak32k
cap f64grid_g16k heat_step_f64 30 2 python bench.py $Q --workload g16k_f64
# keep two reports for the source page, drop the rest (64 MiB limit on what travels back)
rm -f gpurun_out/r02_k1k2_c3.ncu-rep gpurun_out/r02_k1_c2.ncu-rep gpurun_out/r02_k4_weak32k.ncu-rep gpurun_out/r02_f64grid_g16k.ncu-rep gpurun_out/r02_k1_vec_strong.ncu-rep
ls -la gpurun_out | head -60; du -sh gpurun_out
