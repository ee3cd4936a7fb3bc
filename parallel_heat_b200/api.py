"""ctypes mirror of include/heat.h -- the host-side face of libheat_b200.so.

The reference has no library interface (its solver loop is inline in main(),
mpi/mpi_heat_improved_persistent_stat.c:129-269 and cuda/cuda_heat.cu:176-242);
the names and meanings here are the reference's own parameters: NXPROB/NYPROB ->
nx/ny, STEPS -> steps, STEP/CHECK_INTERVAL -> conv_every, parms.cx/cy -> cx/cy,
inidat(), prtdat().  There is no CPU path: if libheat_b200.so is missing or no
CUDA device is usable, calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

PKG = Path(__file__).resolve().parent
# HEAT_B200_LIB selects another build of the same library (the bounds-checked debug twin)
LIB_PATH = Path(os.environ.get("HEAT_B200_LIB") or PKG / "libheat_b200.so")

OK, E_INVALID, E_CUDA, E_NOMEM, E_STATE, E_PEER, E_TIMEOUT = 0, -1, -2, -3, -4, -5, -6
F32_FMA, F32_NOFMA, MPI_F64MIX = 0, 1, 2
ARITH_NAMES = {F32_FMA: "F32_FMA", F32_NOFMA: "F32_NOFMA", MPI_F64MIX: "MPI_F64MIX"}
STEPS_CUDA, STEPS_MPI = 0, 1
PHASE_MPI, PHASE_CUDA = 0, 1
IPC_BLOB_BYTES = 256
HALO_P2P, HALO_EXTERNAL = 0, 1
GRID_F32, GRID_F64 = 0, 1


class HeatError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libheat_b200 error {code}: {msg}")
        self.code = code


class HeatParams(C.Structure):
    """struct heat_params (include/heat.h)."""
    _fields_ = [("nx", C.c_int), ("ny", C.c_int), ("steps", C.c_int), ("cx", C.c_float), ("cy", C.c_float),
                ("conv_every", C.c_int), ("conv_eps", C.c_float), ("conv_phase", C.c_int),
                ("step_convention", C.c_int), ("arith", C.c_int), ("temporal_block", C.c_int),
                ("device", C.c_int), ("rank", C.c_int), ("nranks", C.c_int), ("rank_y", C.c_int), ("nranks_y", C.c_int),
                ("grid_type", C.c_int), ("nz", C.c_int), ("cz", C.c_float)]


ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.POINTER(C.c_float), C.c_void_p)

_fp = C.POINTER(C.c_float)
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_H = C.c_void_p

# every symbol include/heat.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "heat_params_default": (C.c_int, [C.POINTER(HeatParams)]),
    "heat_last_error": (C.c_char_p, []),
    "heat_device_count": (C.c_int, []),
    "heat_inidat": (C.c_int, [C.c_int, C.c_int, _fp]),
    "heat_prtdat": (C.c_int, [C.c_int, C.c_int, _fp, C.c_char_p]),
    "heat_prtdat_bin": (C.c_int, [C.c_int, C.c_int, _fp, C.c_char_p]),
    "heat_readdat_bin": (C.c_int, [C.c_char_p, _ip, _ip, _fp, C.c_size_t]),
    "heat_solve": (C.c_int, [C.POINTER(HeatParams), _fp, _ip, _ip]),
    "heat_solve_f64": (C.c_int, [C.POINTER(HeatParams), _dp, _ip, _ip]),
    "heat_trim": (C.c_int, []),
    "heat_inidat_f64": (C.c_int, [C.c_int, C.c_int, _dp]),
    "heat_inidat3d": (C.c_int, [C.c_int, C.c_int, C.c_int, _fp]),
    "heat_upload_f64": (C.c_int, [_H, _dp]),
    "heat_download_f64": (C.c_int, [_H, _dp]),
    "heat_solve_gpus": (C.c_int, [C.POINTER(HeatParams), C.c_int, _fp, _ip, _ip]),
    "heat_create": (C.c_int, [C.POINTER(HeatParams), C.POINTER(_H)]),
    "heat_destroy": (C.c_int, [_H]),
    "heat_set_stream": (C.c_int, [_H, C.c_void_p]),
    "heat_upload": (C.c_int, [_H, _fp]),
    "heat_upload_device": (C.c_int, [_H, C.c_void_p]),
    "heat_init_inidat": (C.c_int, [_H]),
    "heat_run": (C.c_int, [_H, C.c_int, _ip, _ip]),
    "heat_enqueue": (C.c_int, [_H, C.c_int]),
    "heat_sync": (C.c_int, [_H]),
    "heat_elapsed_ms": (C.c_int, [_H, _fp]),
    "heat_residual": (C.c_int, [_H, _fp]),
    "heat_download": (C.c_int, [_H, _fp]),
    "heat_upload_local": (C.c_int, [_H, _fp]),
    "heat_download_local": (C.c_int, [_H, _fp]),
    "heat_device_view": (C.c_int, [_H, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), _ip, _ip]),
    "heat_launch_count": (C.c_int, [_H, C.POINTER(C.c_int64)]),
    "heat_step_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_void_p]),
    "heat_step_residual_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int,
                                            C.c_void_p, C.c_void_p]),
    "heat_step_residual_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int,
                                             C.c_void_p, C.c_void_p]),
    "heat_strip_rows": (C.c_int, [C.c_int, C.c_int, C.c_int, _ip, _ip]),
    "heat_set_halo_mode": (C.c_int, [_H, C.c_int]),
    "heat_ipc_export": (C.c_int, [_H, C.c_void_p, C.c_size_t]),
    "heat_ipc_connect": (C.c_int, [_H, C.c_void_p, C.c_void_p]),
    "heat_connect_local": (C.c_int, [_H, _H, _H]),
    "heat_connect_local2d": (C.c_int, [_H, _H, _H, _H, _H]),
    "heat_ipc_connect2d": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "heat_group_run": (C.c_int, [C.POINTER(_H), C.c_int, C.c_int, _ip, _ip]),
    "heat_group_enqueue": (C.c_int, [C.POINTER(_H), C.c_int, C.c_int]),
    "heat_set_allreduce_max": (C.c_int, [_H, ALLREDUCE_FN, C.c_void_p]),
    "heat_debug_wait_stats": (C.c_int, [_H, C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)]),
}

_lib = None


def load():
    """dlopen libheat_b200.so (built by parallel_heat_b200.build.build()).  Raises if absent."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise HeatError(E_STATE, f"{LIB_PATH} is not built (python -m parallel_heat_b200.build); "
                                     "there is no CPU fallback")
        lib = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def _check(rc):
    if rc != OK:
        raise HeatError(rc, load().heat_last_error().decode(errors="replace"))


def _f32(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(_fp)


def default_params(**kw):
    p = HeatParams()
    _check(load().heat_params_default(C.byref(p)))
    for k, v in kw.items():
        if not hasattr(p, k):
            raise TypeError(f"heat_params has no field {k!r}")
        setattr(p, k, v)
    return p


def device_count():
    return load().heat_device_count()


def inidat(nx, ny):
    """inidat(nx, ny, u) of the reference (M:315-321), wrapped int32 product included."""
    u = np.empty((nx, ny), np.float32)
    _check(load().heat_inidat(nx, ny, u.ctypes.data_as(_fp)))
    return u


def inidat_f64(nx, ny):
    u = np.empty((nx, ny), np.float64)
    _check(load().heat_inidat_f64(nx, ny, u.ctypes.data_as(_dp)))
    return u


def inidat3d(nx, ny, nz):
    u = np.empty((nx, ny, nz), np.float32)
    _check(load().heat_inidat3d(nx, ny, nz, u.ctypes.data_as(_fp)))
    return u


def prtdat(u, fnam):
    u, p = _f32(u)
    _check(load().heat_prtdat(u.shape[0], u.shape[1], p, str(fnam).encode()))


def prtdat_bin(u, fnam):
    u, p = _f32(u)
    _check(load().heat_prtdat_bin(u.shape[0], u.shape[1], p, str(fnam).encode()))


def readdat_bin(fnam):
    nx, ny = C.c_int(0), C.c_int(0)
    _check(load().heat_readdat_bin(str(fnam).encode(), C.byref(nx), C.byref(ny), None, 0))
    u = np.empty((nx.value, ny.value), np.float32)
    _check(load().heat_readdat_bin(str(fnam).encode(), C.byref(nx), C.byref(ny), u.ctypes.data_as(_fp), u.size))
    return u


def strip_rows(nx, nranks, rank):
    r0, n = C.c_int(0), C.c_int(0)
    _check(load().heat_strip_rows(nx, nranks, rank, C.byref(r0), C.byref(n)))
    return r0.value, n.value


def solve(grid, steps, **params):
    """One-shot heat_solve(): host grid in, newest field out.  Returns (field, updates_done, converged).
    grid_type=GRID_F64 takes / returns float64 through heat_solve_f64()."""
    f64 = params.get("grid_type", GRID_F32) == GRID_F64
    g = np.array(grid, dtype=np.float64 if f64 else np.float32, order="C", copy=True)
    if g.ndim == 3:  # 3-D grid u[ix][iy][iz]
        params = dict(params, nz=g.shape[2])
    p = default_params(nx=g.shape[0], ny=g.shape[1], steps=steps, **params)
    done, conv = C.c_int(0), C.c_int(0)
    if f64:
        _check(load().heat_solve_f64(C.byref(p), g.ctypes.data_as(_dp), C.byref(done), C.byref(conv)))
    else:
        _check(load().heat_solve(C.byref(p), g.ctypes.data_as(_fp), C.byref(done), C.byref(conv)))
    return g, done.value, bool(conv.value)


def trim():
    """heat_trim(): release the device buffers heat_solve keeps between calls."""
    _check(load().heat_trim())


def solve_gpus(grid, steps, ngpus, **params):
    """heat_solve_gpus(): one row strip per GPU of this process.  Returns (field, updates_done, converged)."""
    g = np.array(grid, dtype=np.float32, order="C", copy=True)
    p = default_params(nx=g.shape[0], ny=g.shape[1], steps=steps, **params)
    done, conv = C.c_int(0), C.c_int(0)
    _check(load().heat_solve_gpus(C.byref(p), ngpus, g.ctypes.data_as(_fp), C.byref(done), C.byref(conv)))
    return g, done.value, bool(conv.value)


class HeatSolver:
    """Device-resident solver handle (heat_create ... heat_destroy)."""

    def __init__(self, nx, ny, **params):
        self.params = default_params(nx=nx, ny=ny, **params)
        self._h = _H()
        self._cb = None
        _check(load().heat_create(C.byref(self.params), C.byref(self._h)))

    @property
    def shape(self):
        """Shape of the full grid: (nx, ny), or (nx, ny, nz) for a 3-D grid."""
        p = self.params
        return (p.nx, p.ny, p.nz) if p.nz > 1 else (p.nx, p.ny)

    # -- lifetime
    def close(self):
        if self._h:
            load().heat_destroy(self._h)
            self._h = _H()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- data
    def set_stream(self, cuda_stream_ptr):
        _check(load().heat_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def upload(self, full_grid):
        if self.params.grid_type == GRID_F64:
            g = np.ascontiguousarray(full_grid, dtype=np.float64)
        else:
            g, p = _f32(full_grid)
        if g.shape != self.shape:
            raise ValueError(f"grid shape {g.shape} != {self.shape}")
        if self.params.grid_type == GRID_F64:
            _check(load().heat_upload_f64(self._h, g.ctypes.data_as(_dp)))
        else:
            _check(load().heat_upload(self._h, p))

    def upload_device(self, device_ptr):
        _check(load().heat_upload_device(self._h, C.c_void_p(device_ptr)))

    def init_inidat(self):
        _check(load().heat_init_inidat(self._h))

    def download(self, out=None):
        f64 = self.params.grid_type == GRID_F64
        if out is None:
            out = np.zeros(self.shape, np.float64 if f64 else np.float32)
        assert out.dtype == (np.float64 if f64 else np.float32) and out.flags.c_contiguous
        if f64:
            _check(load().heat_download_f64(self._h, out.ctypes.data_as(_dp)))
        else:
            _check(load().heat_download(self._h, out.ctypes.data_as(_fp)))
        return out

    def upload_local(self, local_rows):
        """local_rows: this strip's window (device_view()[2] rows from global row device_view()[3]), dense."""
        g, p = _f32(local_rows)
        _check(load().heat_upload_local(self._h, p))

    def download_local(self, out):
        """out: (rows this strip owns, ny) float32, C-contiguous (3-D grids: (planes owned, ny*nz))."""
        assert out.dtype == np.float32 and out.flags.c_contiguous
        _check(load().heat_download_local(self._h, out.ctypes.data_as(_fp)))
        return out

    def device_view(self, which=0):
        ptr, pitch, rows, first = C.c_void_p(0), C.c_int64(0), C.c_int(0), C.c_int(0)
        _check(load().heat_device_view(self._h, which, C.byref(ptr), C.byref(pitch), C.byref(rows), C.byref(first)))
        return ptr.value, pitch.value, rows.value, first.value

    # -- time loop
    def run(self, steps):
        done, conv = C.c_int(0), C.c_int(0)
        _check(load().heat_run(self._h, steps, C.byref(done), C.byref(conv)))
        return done.value, bool(conv.value)

    def enqueue(self, updates):
        _check(load().heat_enqueue(self._h, updates))

    def sync(self):
        _check(load().heat_sync(self._h))

    def elapsed_ms(self):
        ms = C.c_float(0)
        _check(load().heat_elapsed_ms(self._h, C.byref(ms)))
        return ms.value

    def residual(self):
        r = C.c_float(0)
        _check(load().heat_residual(self._h, C.byref(r)))
        return r.value

    def launch_count(self):
        n = C.c_int64(0)
        _check(load().heat_launch_count(self._h, C.byref(n)))
        return n.value

    # -- strips
    def set_halo_mode(self, mode):
        _check(load().heat_set_halo_mode(self._h, mode))

    def ipc_export(self):
        blob = C.create_string_buffer(IPC_BLOB_BYTES)
        _check(load().heat_ipc_export(self._h, blob, IPC_BLOB_BYTES))
        return blob.raw

    def ipc_connect(self, north_blob, south_blob, west_blob=None, east_blob=None):
        self._blobs = tuple(C.create_string_buffer(b, IPC_BLOB_BYTES) if b else None
                            for b in (north_blob, south_blob, west_blob, east_blob))
        _check(load().heat_ipc_connect2d(self._h, *self._blobs))

    def connect_local(self, north, south, west=None, east=None):
        _check(load().heat_connect_local2d(self._h, *[o._h if o else None for o in (north, south, west, east)]))

    def wait_stats(self):
        """(ns waited per side, waits per side) of the edge CTAs since the last call; needs HEAT_PROFILE_WAIT=1."""
        ns, cnt = (C.c_ulonglong * 4)(), (C.c_ulonglong * 4)()
        _check(load().heat_debug_wait_stats(self._h, ns, cnt))
        return list(ns), list(cnt)

    def set_allreduce_max(self, fn):
        """fn(value: float) -> float, the maximum over every strip's value."""
        def _cb(pv, _ctx):
            try:
                pv[0] = float(fn(float(pv[0])))
                return 0
            except Exception:  # never unwind through C
                return 1
        self._cb = ALLREDUCE_FN(_cb)
        _check(load().heat_set_allreduce_max(self._h, self._cb, None))


def group_run(solvers, steps):
    """heat_group_run over strips living in this process."""
    arr = (_H * len(solvers))(*[s._h for s in solvers])
    done, conv = C.c_int(0), C.c_int(0)
    _check(load().heat_group_run(arr, len(solvers), steps, C.byref(done), C.byref(conv)))
    return done.value, bool(conv.value)


def group_enqueue(solvers, updates):
    arr = (_H * len(solvers))(*[s._h for s in solvers])
    _check(load().heat_group_enqueue(arr, len(solvers), updates))


def step_device(d_in, d_out, nx, ny, cx=0.1, cy=0.1, arith=F32_FMA, stream=0):
    _check(load().heat_step_device(C.c_void_p(d_in), C.c_void_p(d_out), nx, ny, cx, cy, arith, C.c_void_p(stream)))


def step_residual_device(d_in, d_out, nx, ny, d_residual_bits, cx=0.1, cy=0.1, arith=F32_FMA, stream=0):
    _check(load().heat_step_residual_device(C.c_void_p(d_in), C.c_void_p(d_out), nx, ny, cx, cy, arith,
                                            C.c_void_p(d_residual_bits), C.c_void_p(stream)))


NOALLOC_MAX_NY = 16384 * 256 + 2


def step_residual_noalloc(d_in, d_out, nx, ny, d_residual, cx=0.1, cy=0.1, arith=F32_FMA, stream=0):
    """heat_step_residual_noalloc(): the update + residual for callers that cannot allocate (their
    own pool owns the GPU's memory, or cudaMalloc's device-wide sync is unaffordable).  d_residual
    is one device float; the reduction works in a fixed array inside the library, shared by all
    calls on a GPU (do not overlap them); ny is limited to NOALLOC_MAX_NY."""
    _check(load().heat_step_residual_noalloc(C.c_void_p(d_in), C.c_void_p(d_out), nx, ny, cx, cy, arith,
                                             C.c_void_p(d_residual), C.c_void_p(stream)))
