#!/bin/bash
# one GPU: 3-D kernel with the peer stores templated out of the single-strip instantiation -- tests and two timings
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_3d.py -q -m gpu 2>&1 | tail -n 2
for n in 640 1024; do timeout 200 python bench.py --steps 20 --warmup 5 --only-3d $n --reps 15 2>/dev/null | grep '^{' | python -c "import json,sys; d=json.loads(sys.stdin.read()); print($n, round(d['value'],1), d['rep_ms'], d['parity']['bit_exact'], d['clocks']['sm_mhz'])"; done
