#!/bin/bash
# round 2, first call on one GPU: the suite, smoke, then the driver's own bench commands
mkdir -p gpurun_out
{ free -g; nproc; nvidia-smi -L; nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.limit --format=csv; } > gpurun_out/box.txt 2>&1
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -n 2 gpurun_out/smoke.log
( time timeout 900 python bench.py --steps 20 --warmup 5 ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
tail -n 5 gpurun_out/bench_default.err
( time timeout 600 python bench.py --impl reference --steps 20 --warmup 5 ) > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"
tail -n 4 gpurun_out/bench_reference.err
HEAT_TIMING=1 timeout 300 python - > gpurun_out/solve_phases.log 2>&1 <<'PY'
import ctypes as C, time, torch, os
from parallel_heat_b200 import api
lib = api.load()
for n in (8192, 16384):
    host = torch.empty((n, n), dtype=torch.float32, pin_memory=True)
    hn = host.numpy()
    with api.HeatSolver(n, n) as t:
        t.init_inidat(); t.download(hn)
    for streamed in ("1", "0", "1", "0"):
        os.environ["HEAT_SOLVE_STREAMED"] = streamed
        p = api.default_params(nx=n, ny=n, steps=20)
        d, c = C.c_int(0), C.c_int(0)
        t0 = time.perf_counter()
        rc = lib.heat_solve(C.byref(p), hn.ctypes.data_as(C.POINTER(C.c_float)), C.byref(d), C.byref(c))
        print(n, "streamed" if streamed == "1" else "plain", rc, f"{(time.perf_counter()-t0)*1e3:.3f} ms", flush=True)
PY
cat gpurun_out/solve_phases.log | tail -n 20
