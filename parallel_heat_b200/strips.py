"""Row strips across GPUs, one process per GPU (torchrun), torch.distributed as plumbing.

Replaces the reference's 2-D cartesian block decomposition and persistent halo
requests (mpi/mpi_heat_improved_persistent_stat.c:51-84,100-127,130-155,160-161,
177,263,270-297) with a G x 1 row-strip split inside one NVSwitch box:

  halo="p2p"      (default, the product path) every rank exports its buffers with CUDA
                  IPC; the update kernel stores a strip's first/last owned row straight
                  into the neighbour's ghost row over NVLink and step flags order the
                  strips on the device -- no host work per update, no collective.
  halo="sendrecv" the ghost rows travel as torch.distributed P2P ops (NCCL on GPUs, gloo
                  on CPUs) between updates; the portable fallback and the cross-check.

The convergence check is one 4-byte max-allreduce every conv_every updates
(replaces MPI_Allreduce(LAND), M:255).  torch.distributed is only used for the
blob exchange, barriers, that allreduce and the final gather.
"""
from __future__ import annotations

import numpy as np

from . import api


def partition(nx, nranks):
    """[(row0, rows)] per rank -- heat_strip_rows for every rank."""
    return [api.strip_rows(nx, nranks, r) for r in range(nranks)]


def local_window(nx, nranks, rank):
    """Global rows [lo, hi) a strip keeps: its own rows plus one ghost row per neighbour."""
    r0, n = api.strip_rows(nx, nranks, rank)
    lo = r0 - 1 if rank > 0 else 0
    hi = r0 + n + 1 if rank < nranks - 1 else nx
    return lo, hi


def overlap_window(nx, nranks, rank, depth):
    """Global rows [lo, hi) from which strip `rank` can compute its own rows for `depth` updates on its own:
    its rows plus `depth` more on either side (clipped to the grid).  After k updates only the k outermost
    rows of such a window are wrong (its first and last row are treated as a fixed ring), so after `depth`
    updates every owned row is exact -- redundant work of 2*depth rows instead of any exchange."""
    r0, n = api.strip_rows(nx, nranks, rank)
    return max(0, r0 - depth), min(nx, r0 + n + depth)


def solve_strip_on_host_rows(window_rows, nx, nranks, rank, steps, **params):
    """One-shot solve of ONE strip from host memory, without communication: heat_solve() on the overlap window
    (window_rows: float32 array holding global rows overlap_window(nx, nranks, rank, updates), all ny columns;
    overwritten).  For few updates on a field that lives in host memory the PCIe copies are the cost, and this
    keeps every GPU's H2D, updates and D2H overlapped (heat_solve streams its window in bands) with no step that
    waits for a neighbour; the strips of the reference exchange a halo per update instead (M:160-177).
    Returns (view of the rows this strip owns, updates_done).  Convergence checks need the exchanging path."""
    import ctypes as C
    if params.get("conv_every", 0):
        raise ValueError("solve_strip_on_host_rows: no convergence check (use StripSolver.run)")
    updates = steps + (1 if params.get("step_convention", api.STEPS_CUDA) == api.STEPS_MPI else 0)
    lo, hi = overlap_window(nx, nranks, rank, updates)
    r0, n = api.strip_rows(nx, nranks, rank)
    if window_rows.shape[0] != hi - lo or window_rows.dtype != np.float32 or not window_rows.flags.c_contiguous:
        raise ValueError(f"window must be C-contiguous float32 with rows [{lo}, {hi})")
    p = api.default_params(nx=hi - lo, ny=window_rows.shape[1], steps=steps, **params)
    done, conv = C.c_int(0), C.c_int(0)
    api._check(api.load().heat_solve(C.byref(p), window_rows.ctypes.data_as(C.POINTER(C.c_float)), C.byref(done), C.byref(conv)))
    return window_rows[r0 - lo:r0 - lo + n], done.value


class _CudaRows:
    """torch views of single rows of a strip's device buffer (for halo='sendrecv')."""

    def __init__(self, solver):
        self.solver = solver

    def row(self, which, local_row):
        import torch
        ptr, pitch, rows, _ = self.solver.device_view(which)
        ny = self.solver.params.ny

        class _Arr:  # minimal __cuda_array_interface__ holder
            pass
        a = _Arr()
        a.__cuda_array_interface__ = {"shape": (ny,), "typestr": "<f4", "version": 3,
                                      "data": (ptr + 4 * pitch * local_row, False)}
        return torch.as_tensor(a, device=f"cuda:{self.solver.params.device}")


class StripSolver:
    """The solver loop of one strip; every rank of the process group makes one."""

    # what computes a strip and what exposes single rows of it to torch.distributed: the CUDA library
    make_solver = staticmethod(api.HeatSolver)
    make_rows = staticmethod(_CudaRows)

    def __init__(self, nx, ny, *, group=None, device=None, halo="p2p", cols=1, **params):
        """cols > 1: 2-D blocks, world = (world // cols) row blocks x cols column blocks (the reference's
        cartesian topology, M:51-75); rank r is block (r // cols, r % cols).  halo='p2p' only."""
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        if halo not in ("p2p", "sendrecv"):
            raise ValueError("halo must be 'p2p' or 'sendrecv'")
        self.halo = halo if self.world > 1 else "none"
        self.nx, self.ny = nx, ny
        self.nz = int(params.get("nz", 1) or 1)
        self.row_len = ny * self.nz if self.nz > 1 else ny  # a 3-D grid is cut into strips of ix-planes of ny*nz floats
        if self.nz > 1 and (cols > 1 or halo != "p2p"):
            raise ValueError("3-D grids: strips of planes with halo='p2p' only")
        if device is None:
            device = 0
        self.device = device
        if cols < 1 or self.world % cols:
            raise ValueError(f"{self.world} ranks cannot form blocks with {cols} columns")
        if cols > 1 and halo != "p2p":
            raise ValueError("2-D blocks move their ghost columns with peer stores only (halo='p2p')")
        self.cols, self.gx = cols, self.world // cols
        self.bx, self.by = divmod(self.rank, cols)
        if cols > 1:
            params = dict(params, rank_y=self.by, nranks_y=cols)
        self.solver = self.make_solver(nx, ny, rank=self.bx, nranks=self.gx, device=device, **params)
        self.row0, self.rows = api.strip_rows(nx, self.gx, self.bx)
        self.col0, self.ncols = api.strip_rows(ny, cols, self.by)
        self.lo, self.hi = local_window(nx, self.gx, self.bx)
        self.nloc = self.hi - self.lo
        self.conv_every = int(self.solver.params.conv_every)
        self._rows = self.make_rows(self.solver)
        if self.halo == "p2p":
            self._connect_p2p()
        elif self.halo == "sendrecv":
            if int(self.solver.params.temporal_block) > 1:
                raise ValueError("halo='sendrecv' moves one ghost row per update: use temporal_block=1 (or halo='p2p')")
            self.solver.set_halo_mode(api.HALO_EXTERNAL)
            if self.dist.get_backend(self.group) == "nccl":
                # NCCL P2P ops are ordered against torch's current stream: run the updates there too
                import torch
                self.solver.set_stream(torch.cuda.current_stream(device).cuda_stream)
        if self.world > 1:
            self.solver.set_allreduce_max(self._allreduce_max)
        self.halo = self.halo if self.world > 1 else "none"

    # ------------------------------------------------------------------ wiring
    def _connect_p2p(self):
        blobs = [None] * self.world
        self.dist.all_gather_object(blobs, self.solver.ipc_export(), group=self.group)
        def at(bx, by):
            return blobs[bx * self.cols + by] if 0 <= bx < self.gx and 0 <= by < self.cols else None
        if self.cols > 1:
            self.solver.ipc_connect(at(self.bx - 1, self.by), at(self.bx + 1, self.by), at(self.bx, self.by - 1),
                                    at(self.bx, self.by + 1))
        else:
            self.solver.ipc_connect(at(self.bx - 1, 0), at(self.bx + 1, 0))
        self.dist.barrier(group=self.group)

    def _allreduce_max(self, value):
        import torch
        dev = "cpu" if self.dist.get_backend(self.group) == "gloo" else f"cuda:{self.device}"
        t = torch.tensor([value], dtype=torch.float32, device=dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
        return float(t.item())

    # ------------------------------------------------------------------ data
    def upload(self, full_grid):
        self.solver.upload(full_grid)
        if self.world > 1:
            self.dist.barrier(group=self.group)

    def init_inidat(self):
        self.solver.init_inidat()
        if self.world > 1:
            self.dist.barrier(group=self.group)

    def _exchange(self):
        """halo='sendrecv': newest buffer, first/last owned row -> neighbours' ghost rows."""
        dist = self.dist
        ops = []
        if self.rank > 0:
            ops.append(dist.P2POp(dist.isend, self._rows.row(0, 1), self.rank - 1, self.group))
            ops.append(dist.P2POp(dist.irecv, self._rows.row(0, 0), self.rank - 1, self.group))
        if self.rank < self.world - 1:
            ops.append(dist.P2POp(dist.isend, self._rows.row(0, self.nloc - 2), self.rank + 1, self.group))
            ops.append(dist.P2POp(dist.irecv, self._rows.row(0, self.nloc - 1), self.rank + 1, self.group))
        for w in dist.batch_isend_irecv(ops):
            w.wait()

    # ------------------------------------------------------------------ time loop
    def enqueue(self, updates):
        """`updates` plain updates, asynchronous where the transport allows it."""
        if self.halo == "sendrecv":
            for _ in range(updates):
                self.solver.enqueue(1)
                self._exchange()
        else:
            self.solver.enqueue(updates)

    def sync(self):
        self.solver.sync()

    def run(self, steps):
        """The whole loop (M:159-265): returns (updates_done, converged), same on every rank."""
        if self.halo != "sendrecv":
            return self.solver.run(steps)
        # host-driven exchange: same loop structure as heat_group_run, one update at a time
        p = self.solver.params
        total = steps + (1 if p.step_convention == api.STEPS_MPI else 0)
        done, conv, it = 0, False, getattr(self, "_it", 0)
        K = self.conv_every
        while done < total and not conv:
            check = K > 0 and ((it % K == 0) if p.conv_phase == api.PHASE_CUDA else ((it + 1) % K == 0))
            self.solver.enqueue(1)
            if check:
                r = self._allreduce_max(self.solver.residual())
                conv = r < p.conv_eps
            self._exchange()
            done += 1
            it += 1
        self._it = it
        self.solver.sync()
        return done, conv

    # ------------------------------------------------------------------ result
    def gather(self, dst=0):
        """Newest field assembled on rank `dst` (replaces the END messages, M:270-297)."""
        if self.world == 1:
            return self.solver.download(np.zeros((self.nx, self.ny), np.float32))
        if self.cols == 1:
            # a row strip hands over its own rows only: no rank but `dst` ever holds the whole field
            part = self.solver.download_local(np.empty((self.rows, self.row_len), np.float32))
        else:
            mine = np.zeros((self.nx, self.ny), np.float32)
            self.solver.download(mine)
            part = np.ascontiguousarray(mine[self.row0:self.row0 + self.rows, self.col0:self.col0 + self.ncols])
            del mine
        parts = [None] * self.world if self.rank == dst else None
        self.dist.gather_object(part, parts, dst=dst, group=self.group)
        if self.rank != dst:
            return None
        full = np.concatenate([np.concatenate(parts[bx * self.cols:(bx + 1) * self.cols], axis=1) for bx in range(self.gx)], axis=0)
        return full.reshape(self.nx, self.ny, self.nz) if self.nz > 1 else full

    def close(self):
        self.solver.close()
