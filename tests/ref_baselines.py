#!/usr/bin/env python3
"""Reported baselines: the reference's own programs on this box.
Run on the GPU box:  python tests/ref_baselines.py > gpurun_out/ref_baselines.json
Loop-only time: MPI program = the stand-in's loop stamps; CUDA program = wall(STEPS=S) - wall(STEPS=0),
because its own timer stops without a device sync (cuda_heat.cu:239)."""
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from oracle import pyoracle as O  # noqa: E402

out = {"cores": os.cpu_count(), "rows": []}


def add(name, **kw):
    kw["name"] = name
    out["rows"].append(kw)
    print(json.dumps(kw), file=sys.stderr, flush=True)


def mpi(nx, ny, steps, threads=None, **kw):
    r = O.run_reference("mpi", nx, ny, steps, capture=False, threads=threads, **kw)
    lups = (nx - 2) * (ny - 2) * (steps + 1)
    return dict(nx=nx, ny=ny, updates=steps + 1, loop_secs=r["loop_secs"], elapsed_print=r["elapsed_print"],
                glups=lups / r["loop_secs"] / 1e9, converged_print=r["converged_print"])


def cuda(nx, ny, steps, **kw):
    r0 = O.run_reference("cuda", nx, ny, 0, capture=False, **kw)
    r = O.run_reference("cuda", nx, ny, steps, capture=False, **kw)
    secs = r["wall_secs"] - r0["wall_secs"]
    return dict(nx=nx, ny=ny, updates=steps, loop_secs=secs, wall=r["wall_secs"], wall_steps0=r0["wall_secs"],
                elapsed_print=r["elapsed_print"], glups=(nx - 2) * (ny - 2) * steps / secs / 1e9,
                converged_print=r["converged_print"])


try:
    add("B1 mpi 1024^2 STEPS=1000 as shipped (-O0), 1 rank (stand-in)", **mpi(1024, 1024, 1000, opt="O0"))
    add("B1 mpi 1024^2 STEPS=1000 -O3, 1 rank (stand-in)", **mpi(1024, 1024, 1000, opt="O3"))
    add("B2 mpi 1024^2 STEPS=1000 -O3 -DOMPCH, 4 threads", **mpi(1024, 1024, 1000, omp=True, threads=4))
    add("B2 mpi 1024^2 STEPS=1000 -O3, 4 ranks 2x2 (BASELINE config[0]; shared-memory stand-in, not real MPI)",
        **mpi(1024, 1024, 1000, ranks=4))
    add("B2 mpi 1024^2 STEPS=1000 as shipped (-O0), 4 ranks 2x2", **mpi(1024, 1024, 1000, ranks=4, opt="O0"))
    add("B2 mpi 1024^2 STEPS=1000 -O3 -DOMPCH, all threads", **mpi(1024, 1024, 1000, omp=True))
    add("B2 mpi 1024^2 STEPS=1000 -O3, 16 ranks 4x4", **mpi(1024, 1024, 1000, ranks=16))
    add("B2 mpi 1024^2 -DCONVERGE -DSTEP=20 -O3 -DOMPCH, all threads", **mpi(1024, 1024, 1000, omp=True, conv=True))
    add("B3 mpi 8192^2 STEPS=100 -O3 -DOMPCH, all threads", **mpi(8192, 8192, 100, omp=True))
    add("B3 mpi 8192^2 STEPS=100 -O3 -DOMPCH, 16 ranks 4x4 x 1 thread", **mpi(8192, 8192, 100, omp=True, ranks=16, threads=1))
    add("B3 mpi 8192^2 STEPS=100 -O3 -DOMPCH, 4 ranks 2x2 x 4 threads", **mpi(8192, 8192, 100, omp=True, ranks=4, threads=4))
    add("B3 mpi 8192^2 STEPS=10 as shipped (-O0) -DOMPCH, all threads", **mpi(8192, 8192, 10, omp=True, opt="O0"))
    add("B3 mpi 16384^2 -DCONVERGE STEPS=40 -O3 -DOMPCH, all threads", **mpi(16384, 16384, 40, omp=True, conv=True))
    add("B4 cuda_heat 8192^2 STEPS=2000", **cuda(8192, 8192, 2000))
    add("B4 cuda_heat 1024^2 STEPS=10000", **cuda(1024, 1024, 10000))
    add("B5 cuda_heat -DCONVERGE 16384^2 STEPS=200", **cuda(16384, 16384, 200, conv=True))
except Exception as e:  # keep what we have
    out["error"] = repr(e)
print(json.dumps(out, indent=1))
