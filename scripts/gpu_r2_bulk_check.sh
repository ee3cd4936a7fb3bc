#!/bin/bash
mkdir -p gpurun_out
HEAT_K1_VARIANT=bulk timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "bit_exact or strip or conver or 16384 or 65536 or random" 2>&1 | tail -n 2
for w in strong g16k; do timeout 200 python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-also --reps 9 --workload $w 2>/dev/null | grep '^{' | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$w', d['roofline']['kernel'], round(d['value'],1), {k:round(v,2) for k,v in d['rep_ms'].items()}, d['parity']['bit_exact'])"; done
