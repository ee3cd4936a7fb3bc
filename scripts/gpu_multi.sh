#!/bin/bash
# 2-GPU call: multi-GPU tests, then bench at N=1,2 (+ tile sweep at N=1)
set -x
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi.log
tail -30 gpurun_out/pytest_multi.log
N=$(nvidia-smi -L | wc -l)
: > gpurun_out/multi_bench.log
for n in 2 $N; do
  [ "$n" -le "$N" ] || continue
  echo "## p2p N=$n" >> gpurun_out/multi_bench.log
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 2000 --warmup 20 >> gpurun_out/multi_bench.log 2>&1
  echo "## sendrecv N=$n" >> gpurun_out/multi_bench.log
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --steps 500 --warmup 20 --halo sendrecv >> gpurun_out/multi_bench.log 2>&1
  [ "$n" -eq "$N" ] && break
done
B="python bench.py --steps 1000 --warmup 20 --no-e2e --no-cpu-baseline"
: > gpurun_out/sweep2.log
for rpt in 2 4 6 8 12; do echo "## c2 rpt=$rpt" >> gpurun_out/sweep2.log; HEAT_ROWS_PER_TILE=$rpt $B >> gpurun_out/sweep2.log 2>&1; done
for rpt in 4 8 16 32; do echo "## c3 rpt=$rpt" >> gpurun_out/sweep2.log; HEAT_ROWS_PER_TILE=$rpt python bench.py --workload c3 --steps 400 --warmup 20 --no-e2e --no-cpu-baseline >> gpurun_out/sweep2.log 2>&1; done
grep -E "^##|value" gpurun_out/multi_bench.log | cut -c1-200
