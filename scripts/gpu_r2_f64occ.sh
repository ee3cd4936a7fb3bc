#!/bin/bash
# one GPU: MPI_F64MIX K1 with other unroll depths / register budgets
mkdir -p gpurun_out
P=$PWD/parallel_heat_b200
B="--steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-also --arith f64mix"
: > gpurun_out/f64occ.log
run() { echo "## $1" >> gpurun_out/f64occ.log; shift; timeout 200 "$@" 2>>gpurun_out/f64occ.err | grep '^{' >> gpurun_out/f64occ.log; }
for v in base u2 u2t u4t u8; do
  lib=$P/libheat_b200.so; [ $v != base ] && lib=$P/libheat_b200_$v.so
  run "f64mix c2 $v" env HEAT_B200_LIB=$lib python bench.py $B --workload c2
  run "f64mix g16k $v" env HEAT_B200_LIB=$lib python bench.py $B --workload g16k
  run "f64mix strong $v" env HEAT_B200_LIB=$lib python bench.py $B --workload strong
done
[ -f $P/libheat_b200_u2t.so ] && HEAT_B200_LIB=$P/libheat_b200_u2t.so timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "bit_exact or denormal or random or golden or strip" 2>&1 | tail -n 2
grep -o '^## .*\|"value": [0-9.]*, "unit": "GLUP/s"\|"bit_exact": [a-z]*' gpurun_out/f64occ.log | paste - - -
