#!/usr/bin/env python3
"""Under torchrun: row strips on every GPU of the box against the undecomposed CPU oracle (bit-exact).
    python -m torch.distributed.run --nproc-per-node N tests/check_strips_torchrun.py"""
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import parallel_heat_b200 as ph  # noqa: E402
from parallel_heat_b200.strips import StripSolver  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ok = True
for shape, steps, params in [((1024, 520), 50, dict(arith=0)), ((1031, 2050), 37, dict(arith=1, temporal_block=4)),
                             ((256, 64), 100000, dict(arith=2, conv_every=20, conv_phase=0, step_convention=1, temporal_block=2))]:
    u = ph.inidat(*shape)
    s = StripSolver(shape[0], shape[1], device=local, **params)
    s.upload(u)
    done, conv = s.run(steps)
    full = s.gather(0)
    s.close()
    if rank == 0:
        from oracle import pyoracle as O
        want, wdone, wconv = O.solve(u, steps, arith=params["arith"], conv_every=params.get("conv_every", 0),
                                     conv_phase=params.get("conv_phase", 1), step_convention=params.get("step_convention", 0))
        same = np.array_equal(full.view(np.uint32), want.view(np.uint32)) and (done, conv) == (wdone, wconv)
        print(f"strips x{world} {shape} {params}: updates {done} converged {conv} bit-exact {same}", flush=True)
        ok = ok and same
dist.destroy_process_group()
sys.exit(0 if ok else 1)
