#!/bin/bash
# round 2, one GPU: cache-policy, CTA-width and tile-height variants of K1 / K1b at 65536^2 (and 16384^2)
mkdir -p gpurun_out
P=$PWD/parallel_heat_b200
B="--steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-also --reps 9"
: > gpurun_out/policy.log
run() { echo "## $1" >> gpurun_out/policy.log; shift; timeout 200 "$@" 2>>gpurun_out/policy.err | grep '^{' >> gpurun_out/policy.log; }
for v in base stcs ldcs both ldnc t256; do
  lib=$P/libheat_b200.so; [ $v != base ] && lib=$P/libheat_b200_$v.so
  run "strong K1 $v" env HEAT_B200_LIB=$lib HEAT_K1_VARIANT=vec python bench.py $B
  run "strong K1b $v" env HEAT_B200_LIB=$lib HEAT_K1_VARIANT=bulk python bench.py $B
  run "g16k K1 $v" env HEAT_B200_LIB=$lib HEAT_K1_VARIANT=vec python bench.py $B --workload g16k
done
for r in 4 16 32; do run "strong K1 base rows_per_tile=$r" env HEAT_K1_VARIANT=vec HEAT_ROWS_PER_TILE=$r python bench.py $B; done
run "strong K1 base no serpentine" env HEAT_K1_VARIANT=vec HEAT_SERPENTINE=0 python bench.py $B
run "weak32k T4 both" env HEAT_B200_LIB=$P/libheat_b200_both.so python bench.py $B --workload weak32k
run "weak32k T4 base" python bench.py $B --workload weak32k
grep -o '^## .*\|"value": [0-9.]*, "unit": "GLUP/s"\|"bit_exact": [a-z]*' gpurun_out/policy.log | paste - - -
