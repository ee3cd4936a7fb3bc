"""world_size-2 (and 3) gloo runs of the row-strip driver on CPU.

parallel_heat_b200.strips.StripSolver is the real host logic (partition, halo transport over
torch.distributed P2P, max-allreduce for convergence, gather); only the per-strip compute is
swapped for an oracle-backed stand-in, because there is no GPU here.  The result must be bit
identical to the undecomposed oracle run."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]


class OracleStrip:
    """Stand-in for api.HeatSolver with the same surface, computing with the CPU oracle."""

    class _P:
        pass

    def __init__(self, nx, ny, rank=0, nranks=1, device=0, **params):
        from oracle import pyoracle as O
        from parallel_heat_b200 import api
        self.O = O
        p = api.default_params(nx=nx, ny=ny, rank=rank, nranks=nranks, **params)
        self.params = p
        r0, n = api.strip_rows(nx, nranks, rank)
        self.row0, self.rows = r0, n
        self.lo = r0 - 1 if rank > 0 else 0
        self.hi = r0 + n + 1 if rank < nranks - 1 else nx
        self.buf = [None, None]
        self.cur = 0

    def set_halo_mode(self, mode):
        self.external = mode

    def set_allreduce_max(self, fn):
        self.allreduce = fn

    def upload(self, full):
        loc = np.ascontiguousarray(full[self.lo:self.hi], np.float32)
        self.buf = [loc.copy(), loc.copy()]
        self.cur = 0

    def enqueue(self, updates):
        for _ in range(updates):
            self.O.step(self.buf[self.cur], self.params.arith, self.params.cx, self.params.cy, out=self.buf[1 - self.cur])
            self.cur = 1 - self.cur

    def residual(self):
        a, b = self.buf
        return float(np.abs(a[1:-1] - b[1:-1]).max()) if a.shape[0] > 2 else 0.0

    def sync(self):
        pass

    def download(self, out):
        first = self.row0 - self.lo
        out[self.row0:self.row0 + self.rows] = self.buf[self.cur][first:first + self.rows]
        return out

    def download_local(self, out):
        first = self.row0 - self.lo
        out[:] = self.buf[self.cur][first:first + self.rows]
        return out

    def close(self):
        pass


class NumpyRows:
    def __init__(self, solver):
        self.s = solver

    def row(self, which, local_row):
        b = self.s.buf[self.s.cur if which == 0 else 1 - self.s.cur]
        return torch.from_numpy(b[local_row])  # shares memory: irecv writes the ghost row in place


def _worker(rank, world, port, shape, steps, params, out_path):
    sys.path.insert(0, str(ROOT))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    rdzv = "file://" + out_path + ".rdzv"  # a file store: no TCP port to collide on
    dist.init_process_group("gloo", init_method=rdzv, rank=rank, world_size=world)
    try:
        from oracle import pyoracle as O
        from parallel_heat_b200.strips import StripSolver
        u = O.inidat(*shape)

        class CpuStrips(StripSolver):  # same host logic, strips computed by the oracle (no GPU here)
            make_solver = staticmethod(OracleStrip)
            make_rows = staticmethod(NumpyRows)
        s = CpuStrips(shape[0], shape[1], halo="sendrecv", **params)
        s.upload(u)
        done, conv = s.run(steps)
        full = s.gather(0)
        if rank == 0:
            np.savez(out_path, full=full, done=done, conv=conv)
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,shape,steps,params", [
    (2, (20, 20), 100, dict(arith=2, step_convention=1)),
    (2, (33, 47), 40, dict(arith=1)),
    (3, (31, 18), 25, dict(arith=0)),
    (2, (20, 20), 10000, dict(arith=2, step_convention=1, conv_every=20, conv_phase=0)),
])
def test_strips_over_gloo_match_the_undecomposed_oracle(tmp_path, oracle, ph, world, shape, steps, params):
    out = tmp_path / "out.npz"
    mp.spawn(_worker, args=(world, _free_port(), shape, steps, params, str(out)), nprocs=world, join=True)
    r = np.load(out)
    want, wdone, wconv = oracle.solve(oracle.inidat(*shape), steps, arith=params.get("arith", 0),
                                      step_convention=params.get("step_convention", 0),
                                      conv_every=params.get("conv_every", 0), conv_phase=params.get("conv_phase", 1))
    assert int(r["done"]) == wdone and bool(r["conv"]) == wconv
    assert np.array_equal(r["full"].view(np.uint32), want.view(np.uint32))
    if "conv_every" in params:
        assert wconv and wdone - 1 == 1979  # what the reference prints for this case


def test_partition_and_windows(ph):
    from parallel_heat_b200 import strips
    assert strips.partition(10, 3) == [(0, 4), (4, 3), (7, 3)]
    assert strips.local_window(10, 3, 0) == (0, 5)
    assert strips.local_window(10, 3, 1) == (3, 8)
    assert strips.local_window(10, 3, 2) == (6, 10)
    assert strips.local_window(10, 1, 0) == (0, 10)


def test_overlap_windows(ph):
    from parallel_heat_b200 import strips
    # own rows + depth on either side, clipped to the grid
    assert strips.overlap_window(100, 4, 0, 10) == (0, 35)
    assert strips.overlap_window(100, 4, 1, 10) == (15, 60)
    assert strips.overlap_window(100, 4, 3, 10) == (65, 100)
    assert strips.overlap_window(100, 4, 1, 1000) == (0, 100)
    assert strips.overlap_window(100, 1, 0, 7) == (0, 100)
