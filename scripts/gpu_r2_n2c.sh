#!/bin/bash
# two GPUs: K1 against K1b in flag-ordered strips, alternating, 21 repetitions each
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
: > gpurun_out/n2c.log
run() { echo "## $1" >> gpurun_out/n2c.log; shift; timeout 300 "$@" >> gpurun_out/n2c.log 2>gpurun_out/n2c.err; echo "rc=$?" >> gpurun_out/n2c.log; }
port=29520
for k in 1 2 3; do
  for v in vec bulk; do
    port=$((port+1))
    run "strong N=2 $v #$k" env HEAT_K1_VARIANT=$v $TR --master-port $port bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --no-also --reps 21
  done
done
python - <<'PY'
import json
for line in open('gpurun_out/n2c.log'):
    line=line.strip()
    if line.startswith('##'): print(line, end='  ')
    elif line.startswith('{'):
        d=json.loads(line); print(round(d['value'],1), {k:round(v,2) for k,v in d['rep_ms'].items()}, d['parity']['bit_exact'])
PY
