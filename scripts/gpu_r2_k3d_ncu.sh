#!/bin/bash
# one GPU: ncu of the final 3-D kernel (64 registers, 2 rows x 8 warps)
mkdir -p gpurun_out
Q="--steps 20 --warmup 5 --only-3d 640 --reps 7"
python bench.py $Q > gpurun_out/plain_3d.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:heat_step_3d -s 30 -c 2 -f -o gpurun_out/r02_k3d_640 python bench.py $Q > gpurun_out/ncu_3d.log 2>&1
ncu -i gpurun_out/r02_k3d_640.ncu-rep --page raw --csv > gpurun_out/r02_k3d_640_raw.csv 2>/dev/null
ncu -i gpurun_out/r02_k3d_640.ncu-rep --page details --csv > gpurun_out/r02_k3d_640_details.csv 2>/dev/null
rm -f gpurun_out/r02_k3d_640.ncu-rep
tail -n 2 gpurun_out/plain_3d.log | cut -c1-300
