#!/bin/bash
# 2 GPUs: 2-D blocks across processes -- parity (incl. the padding-CTA case and HEAT_COLS_FIRST=M), then the
# weak 8192^2-per-GPU bench as 1x2 blocks with HEAT_COLS_FIRST=8 and 16 (M=0, M=1 and 2 strips: see the
# earlier run in profiles/r01_n2_cols_first.txt).
mkdir -p gpurun_out
timeout 66 python -m pytest tests/test_multi_gpu.py -m gpu -x -q -k blocks > gpurun_out/pytest_blocks2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_blocks2.log
tail -n 4 gpurun_out/pytest_blocks2.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
B="bench.py --gpus 2 --steps 3000 --warmup 50 --no-e2e --no-cpu-baseline --no-also"
: > gpurun_out/n2cols2.log
run() { echo "## $1" >> gpurun_out/n2cols2.log; shift; timeout 15 "$@" 2>/dev/null | grep '^{' >> gpurun_out/n2cols2.log; }
run "weak c2 N=2 1x2 blocks HEAT_COLS_FIRST=8" env HEAT_COLS_FIRST=8 $TR --master-port 29512 $B --cols 2
run "weak c2 N=2 1x2 blocks HEAT_COLS_FIRST=16" env HEAT_COLS_FIRST=16 $TR --master-port 29514 $B --cols 2
cut -c1-200 gpurun_out/n2cols2.log
