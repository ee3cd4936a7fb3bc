#!/bin/bash
# two GPUs: NVLink data counters around a strips run (nvidia-smi, since NVML's field values are unsupported here)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
{ echo "## before"; nvidia-smi nvlink -gt d 2>&1 | head -60; } > gpurun_out/nvlink_counters.txt
timeout 300 $TR --master-port 29531 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --no-also --k3-profile > gpurun_out/nvlink_run.json 2> gpurun_out/nvlink_run.err; echo "rc=$?"
{ echo "## after"; nvidia-smi nvlink -gt d 2>&1 | head -60; } >> gpurun_out/nvlink_counters.txt
{ echo "## status"; nvidia-smi nvlink -s 2>&1 | head -24; } >> gpurun_out/nvlink_counters.txt
head -c 1500 gpurun_out/nvlink_counters.txt; grep -c . gpurun_out/nvlink_counters.txt
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/nvlink_run.json') if l.startswith('{')][-1])
print(d['value'], d['steps'], d['reps'], d['k3'])
PY
