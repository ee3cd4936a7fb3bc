#!/bin/bash
# one GPU: heat_solve with its buffers kept between calls -- test, phases at 8192^2 / 16384^2, whole suite, default bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "keeps_and_trims or streamed or strips_from_host" 2>&1 | tail -n 3
HEAT_TIMING=1 timeout 300 python - > gpurun_out/solve_phases_keep.log 2>&1 <<'PY'
import ctypes as C, time, torch, os
from parallel_heat_b200 import api
lib = api.load()
for n in (8192, 16384):
    host = torch.empty((n, n), dtype=torch.float32, pin_memory=True)
    hn = host.numpy()
    with api.HeatSolver(n, n) as t:
        t.init_inidat(); t.download(hn)
    for keep in ("1", "1", "1", "0", "0"):
        os.environ["HEAT_SOLVE_KEEP_BUFFERS"] = keep
        p = api.default_params(nx=n, ny=n, steps=20)
        d, c = C.c_int(0), C.c_int(0)
        t0 = time.perf_counter()
        rc = lib.heat_solve(C.byref(p), hn.ctypes.data_as(C.POINTER(C.c_float)), C.byref(d), C.byref(c))
        print(n, "keep" if keep == "1" else "no keep", rc, f"{(time.perf_counter()-t0)*1e3:.3f} ms", flush=True)
    api.trim()
PY
cat gpurun_out/solve_phases_keep.log
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -n 3 gpurun_out/pytest_gpu.log
( time timeout 900 python bench.py --steps 20 --warmup 5 ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_default.json').read().strip().splitlines()[-1])
print("main", d['value'], "e2e", d['e2e']['value'], d['e2e']['seconds_all'], d['e2e']['parity']['bit_exact'])
PY
