#!/bin/bash
# round 2, one GPU: A/B of the kernel variants (each after its own parity run), then the ncu captures
mkdir -p gpurun_out
P=$PWD/parallel_heat_b200
B="--steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-also"
: > gpurun_out/ab.log; : > gpurun_out/ab.err
run() { echo "## $1" >> gpurun_out/ab.log; shift; timeout 300 "$@" 2>>gpurun_out/ab.err | grep '^{' >> gpurun_out/ab.log; }
par() { echo "## parity $1" >> gpurun_out/ab_parity.log; shift; timeout 600 "$@" >> gpurun_out/ab_parity.log 2>&1; echo "rc=$?" >> gpurun_out/ab_parity.log; }
: > gpurun_out/ab_parity.log
par "new tests, product library" python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "strips_from_host or streamed or block"
# --- MPI_F64MIX: how many of the three widenings per point go through the integer pipes
for v in base w1 w2 w3; do
  lib=$P/libheat_b200.so; [ $v != base ] && lib=$P/libheat_b200_$v.so
  [ $v != base ] && par "f64mix $v" env HEAT_B200_LIB=$lib python -m pytest tests/test_gpu_parity.py -x -q -m gpu
  run "f64mix c2 $v" env HEAT_B200_LIB=$lib python bench.py $B --workload c2 --arith f64mix
  run "f64mix g16k $v" env HEAT_B200_LIB=$lib python bench.py $B --workload g16k --arith f64mix
  run "f64mix strong $v" env HEAT_B200_LIB=$lib python bench.py $B --workload strong --arith f64mix
  run "f64mix weak32k T4 $v" env HEAT_B200_LIB=$lib python bench.py $B --workload weak32k --arith f64mix
done
# --- packed FP32 (FADD2/FFMA2) in K1's plain path and K4's fast path
par "packed" env HEAT_B200_LIB=$P/libheat_b200_packed.so python -m pytest tests/test_gpu_parity.py -x -q -m gpu
for v in base packed; do
  lib=$P/libheat_b200.so; [ $v != base ] && lib=$P/libheat_b200_$v.so
  for w in c2 g16k strong weak32k; do run "fma $w $v" env HEAT_B200_LIB=$lib python bench.py $B --workload $w; done
  run "fma c2 T4 $v" env HEAT_B200_LIB=$lib python bench.py $B --workload c2 --temporal-block 4
  run "nofma weak32k $v" env HEAT_B200_LIB=$lib python bench.py $B --workload weak32k --arith nofma
done
# --- the shared-memory staged variant of K1 (cp.async.bulk rows into a tile with halo)
par "bulk" env HEAT_K1_VARIANT=bulk python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "bit_exact or ring or tiling or 8192 or random or streamed"
for w in c2 g16k strong; do run "fma $w bulk-staged" env HEAT_K1_VARIANT=bulk python bench.py $B --workload $w; done
grep -o '^## .*\|"value": [0-9.]*, "unit": "GLUP/s", "n_gpus"' gpurun_out/ab.log | paste - - | sed 's/"unit.*//' 
grep "^##\|passed\|failed\|rc=" gpurun_out/ab_parity.log
# --- ncu: launch list of the driver's command, then full captures of the top kernels
Q="--steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-also --no-parity --reps 7"
python bench.py $Q > gpurun_out/plain_strong.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_strong_launches.csv python bench.py $Q > gpurun_out/ncu_launches.log 2>&1
python bench.py $Q > gpurun_out/plain_strong.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:heat_step_vec -s 30 -c 2 -f -o gpurun_out/r02_k1_strong python bench.py $Q > gpurun_out/ncu_k1_strong.log 2>&1
python bench.py $Q --workload c3 > gpurun_out/plain_c3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:heat_step_vec -s 24 -c 3 -f -o gpurun_out/r02_k1k2_c3 python bench.py $Q --workload c3 > gpurun_out/ncu_c3.log 2>&1
python bench.py $Q --workload c2 --arith f64mix > gpurun_out/plain_f64.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:heat_step_vec -s 30 -c 2 -f -o gpurun_out/r02_k1_f64mix_c2 python bench.py $Q --workload c2 --arith f64mix > gpurun_out/ncu_f64.log 2>&1
HEAT_B200_LIB=$P/libheat_b200_w1.so python bench.py $Q --workload c2 --arith f64mix > gpurun_out/plain_f64w1.log 2>&1 &&
HEAT_B200_LIB=$P/libheat_b200_w1.so ncu --set full --clock-control none --import-source on -k regex:heat_step_vec -s 30 -c 2 -f -o gpurun_out/r02_k1_f64mix_w1_c2 python bench.py $Q --workload c2 --arith f64mix > gpurun_out/ncu_f64w1.log 2>&1
HEAT_K1_VARIANT=bulk python bench.py $Q --workload g16k > gpurun_out/plain_bulk.log 2>&1 &&
HEAT_K1_VARIANT=bulk ncu --set full --clock-control none --import-source on -k regex:heat_step_bulk -s 30 -c 2 -f -o gpurun_out/r02_k1_bulk_g16k python bench.py $Q --workload g16k > gpurun_out/ncu_bulk.log 2>&1
python bench.py $Q --workload g16k > gpurun_out/plain_g16k.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:heat_step_vec -s 30 -c 2 -f -o gpurun_out/r02_k1_g16k python bench.py $Q --workload g16k > gpurun_out/ncu_g16k.log 2>&1
python bench.py $Q --workload weak32k > gpurun_out/plain_tb.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:heat_step_tb -s 8 -c 2 -f -o gpurun_out/r02_k4_weak32k python bench.py $Q --workload weak32k > gpurun_out/ncu_tb.log 2>&1
ls -la gpurun_out/*.ncu-rep
