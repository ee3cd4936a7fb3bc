"""Row strips over several GPUs (needs >= 2 GPUs; skipped otherwise).

* one process per GPU, CUDA-IPC peer stores + device-side step flags (the bench path),
  and the NCCL send/recv fallback -- both bit-identical to the undecomposed oracle run;
* one process driving two GPUs (heat_connect_local across devices)."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]


def ngpus():
    import torch
    return torch.cuda.device_count()


needs2 = pytest.mark.skipif(ngpus() < 2, reason="needs 2 GPUs")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, shape, steps, halo, params, out_path, cols=1):
    sys.path.insert(0, str(ROOT))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    os.environ["HEAT_COLS_FIRST"] = str(int(params.pop("cols_first", 0)))  # read by heat_create (0 = grid order)
    rdzv = "file://" + out_path + ".rdzv"  # a file store: no TCP port to collide on
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=rdzv, rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import parallel_heat_b200 as ph
        from parallel_heat_b200.strips import StripSolver
        u = (ph.inidat3d(*shape) if len(shape) == 3 else ph.inidat(*shape)) if params.pop("inidat", True) else None
        if len(shape) == 3:
            params = dict(params, nz=shape[2])
        s = StripSolver(shape[0], shape[1], device=rank, halo=halo, cols=cols, **params)
        s.upload(u)
        done, conv = s.run(steps)
        full = s.gather(0)
        if rank == 0:
            np.savez(out_path, full=full, done=done, conv=conv)
        s.close()
    finally:
        dist.destroy_process_group()


@needs2
@pytest.mark.timeout(600)
@pytest.mark.parametrize("halo", ["p2p", "sendrecv"])
@pytest.mark.parametrize("shape,steps,params", [
    ((64, 64), 50, dict(arith=0)),
    ((257, 1030), 33, dict(arith=1)),
    ((40, 24), 100000, dict(arith=2, step_convention=1, conv_every=20, conv_phase=0)),
    ((2048, 4096), 40, dict(arith=0)),
])
def test_strips_one_process_per_gpu(tmp_path, oracle, ph, halo, shape, steps, params):
    import torch.multiprocessing as mp
    world = min(ngpus(), 4) if shape[0] >= 64 else 2
    out = tmp_path / "out.npz"
    mp.spawn(_worker, args=(world, _free_port(), shape, steps, halo, dict(params), str(out)), nprocs=world, join=True)
    r = np.load(out)
    want, wdone, wconv = oracle.solve(oracle.inidat(*shape), steps, arith=params.get("arith", 0),
                                      step_convention=params.get("step_convention", 0),
                                      conv_every=params.get("conv_every", 0), conv_phase=params.get("conv_phase", 1))
    assert (int(r["done"]), bool(r["conv"])) == (wdone, wconv)
    assert np.array_equal(r["full"].view(np.uint32), want.view(np.uint32))


@needs2
def test_one_process_two_gpus(ph, oracle):
    g = min(ngpus(), 4)
    u = ph.inidat(300, 520)
    ss = [ph.HeatSolver(300, 520, rank=r, nranks=g, device=r, arith=ph.F32_NOFMA) for r in range(g)]
    try:
        for r, s in enumerate(ss):
            s.connect_local(ss[r - 1] if r > 0 else None, ss[r + 1] if r < g - 1 else None)
        for s in ss:
            s.upload(u)
        done, _ = ph.group_run(ss, 60)
        out = np.zeros_like(u)
        for s in ss:
            s.download(out)
    finally:
        for s in ss:
            s.close()
    want, _, _ = oracle.solve(u, 60, arith=1)
    assert done == 60 and np.array_equal(out.view(np.uint32), want.view(np.uint32))


@needs2
def test_solve_gpus_one_shot(ph, oracle):
    u = ph.inidat(400, 300)
    want, wdone, wconv = oracle.solve(u, 61, arith=2, conv_every=20, conv_phase=0, step_convention=1)
    got, done, conv = ph.solve_gpus(u, 61, min(ngpus(), 4), arith=ph.MPI_F64MIX, conv_every=20, conv_phase=ph.PHASE_MPI,
                                    step_convention=ph.STEPS_MPI, temporal_block=2)
    assert (done, conv) == (wdone, wconv)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@needs2
@pytest.mark.timeout(600)
@pytest.mark.parametrize("shape,steps,params", [
    ((64, 64), 50, dict(arith=0)),
    ((257, 1030), 33, dict(arith=1)),
    ((40, 48), 100000, dict(arith=2, step_convention=1, conv_every=20, conv_phase=0)),
    # 1024 columns in two blocks: the west block's last owned column ends a CTA and a CTA of padding follows it,
    # so "the east-edge CTA" is not the last one of the tile row
    ((130, 1024), 40, dict(arith=0)),
    # HEAT_COLS_FIRST=M: column-edge CTAs first, M tile rows each (three CTA columns per block, with and
    # without the padding CTA; M dividing the tile rows or not; both sweep directions)
    ((130, 2048), 40, dict(arith=0, cols_first=1)),
    ((300, 2052), 31, dict(arith=1, cols_first=3)),
    ((517, 2048), 24, dict(arith=0, cols_first=8)),
])
def test_blocks_one_process_per_gpu(tmp_path, oracle, ph, shape, steps, params):
    # column blocks across processes: ghost columns by peer stores, four flags per block
    import torch.multiprocessing as mp
    world = 4 if ngpus() >= 4 else 2
    cols = 2
    out = tmp_path / "out.npz"
    mp.spawn(_worker, args=(world, _free_port(), shape, steps, "p2p", dict(params), str(out), cols), nprocs=world, join=True)
    r = np.load(out)
    want, wdone, wconv = oracle.solve(oracle.inidat(*shape), steps, arith=params.get("arith", 0),
                                      step_convention=params.get("step_convention", 0),
                                      conv_every=params.get("conv_every", 0), conv_phase=params.get("conv_phase", 1))
    assert (int(r["done"]), bool(r["conv"])) == (wdone, wconv)
    assert np.array_equal(r["full"].view(np.uint32), want.view(np.uint32))


@needs2
@pytest.mark.timeout(900)
@pytest.mark.parametrize("cols,shape,steps,params", [
    # many passes across processes (hundreds of flag epochs in both sweep directions): row strips, then 2-D blocks
    # with both kinds of neighbours (cols = world on 2 GPUs: column neighbours only; on 4 and 8 GPUs: 2x2, 2x4)
    (1, (1030, 2052), 600, dict(arith=0)),
    (1, (1030, 2052), 600, dict(arith=2, temporal_block=4)),
    (0, (1100, 4100), 500, dict(arith=0)),
    (0, (1100, 4100), 500, dict(arith=1, cols_first=4)),
])
def test_long_runs_across_processes(tmp_path, oracle, ph, cols, shape, steps, params):
    import torch.multiprocessing as mp
    world = 8 if ngpus() >= 8 else 4 if ngpus() >= 4 else 2
    if cols == 0:
        cols = world // 2 if world >= 4 else 2
    out = tmp_path / "out.npz"
    mp.spawn(_worker, args=(world, _free_port(), shape, steps, "p2p", dict(params), str(out), cols), nprocs=world, join=True)
    r = np.load(out)
    want, wdone, _ = oracle.solve(oracle.inidat(*shape), steps, arith=params.get("arith", 0))
    assert int(r["done"]) == wdone
    assert np.array_equal(r["full"].view(np.uint32), want.view(np.uint32))


@needs2
@pytest.mark.timeout(600)
def test_3d_strips_of_planes_across_processes(tmp_path, oracle, ph):
    # 3-D grid, one process per GPU: ghost planes by peer stores, device-side step flags
    import torch.multiprocessing as mp
    world = min(ngpus(), 4)
    shape, steps = (96, 40, 132), 60
    out = tmp_path / "out.npz"
    mp.spawn(_worker, args=(world, _free_port(), shape, steps, "p2p", dict(arith=1), str(out)), nprocs=world, join=True)
    r = np.load(out)
    want, wdone, _ = oracle.solve3d(oracle.inidat3d(*shape), steps, arith=1)
    assert int(r["done"]) == wdone
    assert np.array_equal(r["full"].view(np.uint32), want.view(np.uint32))
