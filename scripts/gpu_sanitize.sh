#!/bin/bash
# compute-sanitizer memcheck on a small, representative subset (one tool per gpurun call)
mkdir -p gpurun_out
timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest tests/test_gpu_parity.py -m gpu -x -q \
  -k "(updates_bit_exact and (33-47 or 4-64 or 129-257 or 3-3)) or (temporal_blocking_is_bit_identical and 0-4) or (strip_decomposition and 3-0) or temporal_blocking_in_strips or ring_and_both or stateless" \
  > gpurun_out/sanitizer_memcheck.log 2>&1
echo "rc=$?" >> gpurun_out/sanitizer_memcheck.log
tail -15 gpurun_out/sanitizer_memcheck.log
