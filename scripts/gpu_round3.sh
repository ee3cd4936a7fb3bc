#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
B="python bench.py --steps 2000 --warmup 20 --no-e2e --no-cpu-baseline"
: > gpurun_out/tb.log
for T in 1 2 3 4; do for rpt in 32 64 128; do
  echo "## c2 T=$T rpt_tb=$rpt" >> gpurun_out/tb.log; HEAT_ROWS_PER_TILE_TB=$rpt $B --temporal-block $T >> gpurun_out/tb.log 2>&1
  [ $T -eq 1 ] && break
done; done
echo "## weak32k T=4" >> gpurun_out/tb.log; python bench.py --workload weak32k --steps 400 --warmup 8 --no-e2e --no-cpu-baseline >> gpurun_out/tb.log 2>&1
echo "## c3 T=1" >> gpurun_out/tb.log; python bench.py --workload c3 --steps 400 --warmup 20 --no-e2e --no-cpu-baseline >> gpurun_out/tb.log 2>&1
N=$(nvidia-smi -L | wc -l)
if [ "$N" -ge 2 ]; then
for T in 1 4; do
  echo "## p2p N=2 T=$T" >> gpurun_out/tb.log
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2000 --warmup 20 --temporal-block $T >> gpurun_out/tb.log 2>&1
done
fi
grep -E "^##|value" gpurun_out/tb.log | cut -c1-160
CMD="python bench.py --steps 40 --warmup 4 --temporal-block 4 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain_tb.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:heat_step_tb -s 4 -c 2 -f -o gpurun_out/prof_tb $CMD > gpurun_out/ncu_tb.log 2>&1
