#!/bin/bash
# 1 GPU, no Python: does the SHAPE of a 1x2 block (16384 rows x 4098 columns) cost what 1x2 blocks lose to strips?
mkdir -p gpurun_out; : > gpurun_out/geometry.log
for ny in 8192 4096 4098 4097; do
  nx=$((8192 * 8192 / (ny / 2 * 2)))
  timeout 6 src/cuda_heat --nx $nx --ny $ny --steps 3000 --no-dump 2>> gpurun_out/geometry.log >/dev/null
done
cat gpurun_out/geometry.log
