#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_f64.py -q -m gpu 2>&1 | tail -n 2
timeout 200 python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-also --workload g16k_f64 2>/dev/null | grep '^{' | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['roofline']['frac'],3), d['parity']['bit_exact'])"
