"""ctypes front-end of the CPU oracle and runner for the prebuilt reference binaries.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package
(parallel_heat_b200) must never import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import re
import resource
import subprocess
import tempfile
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB = HERE / "_build" / "libheat_oracle.so"
REF = HERE / "_ref"

F32_FMA, F32_NOFMA, MPI_F64MIX = 0, 1, 2
ARITH_NAMES = {F32_FMA: "F32_FMA", F32_NOFMA: "F32_NOFMA", MPI_F64MIX: "MPI_F64MIX"}

_lib = None
_fp = C.POINTER(C.c_float)


def build():
    subprocess.run(["make", "-s", "-C", str(HERE), "all"], check=True)


def lib():
    global _lib
    if _lib is None:
        src_m = max((HERE / n).stat().st_mtime for n in ("heat_oracle.c", "heat_oracle.h"))
        if not LIB.exists() or LIB.stat().st_mtime < src_m:
            build()
        L = C.CDLL(str(LIB))
        L.oracle_inidat.argtypes = [C.c_int, C.c_int, _fp]
        L.oracle_inidat.restype = None
        L.oracle_prtdat.argtypes = [C.c_int, C.c_int, _fp, C.c_char_p]
        L.oracle_prtdat.restype = C.c_int
        L.oracle_step.argtypes = [C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, _fp, _fp]
        L.oracle_step.restype = None
        L.oracle_max_abs_diff.argtypes = [C.c_int64, _fp, _fp]
        L.oracle_max_abs_diff.restype = C.c_float
        L.oracle_solve.argtypes = [C.c_int] * 4 + [C.c_float, C.c_float, C.c_int, C.c_float, C.c_int, C.c_int,
                                                   _fp, _fp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.oracle_solve.restype = C.c_int
        L.oracle_threads.restype = C.c_int
        L.oracle_inidat3d.argtypes = [C.c_int, C.c_int, C.c_int, _fp]
        L.oracle_inidat3d.restype = None
        L.oracle_solve3d.argtypes = [C.c_int] * 5 + [C.c_float] * 3 + [C.c_int, C.c_float, C.c_int, C.c_int, _fp, _fp,
                                                                     C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.oracle_solve3d.restype = C.c_int
        _dp = C.POINTER(C.c_double)
        L.oracle_inidat_f64.argtypes = [C.c_int, C.c_int, _dp]
        L.oracle_inidat_f64.restype = None
        L.oracle_solve_f64.argtypes = [C.c_int] * 4 + [C.c_float, C.c_float, C.c_int, C.c_float, C.c_int, C.c_int,
                                                       _dp, _dp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.oracle_solve_f64.restype = C.c_int
        _lib = L
    return _lib


def _p(a):
    assert a.dtype == np.float32 and a.flags.c_contiguous
    return a.ctypes.data_as(_fp)


def inidat(nx, ny):
    u = np.empty((nx, ny), np.float32)
    lib().oracle_inidat(nx, ny, _p(u))
    return u


def prtdat(u, fnam):
    nx, ny = u.shape
    if lib().oracle_prtdat(nx, ny, _p(np.ascontiguousarray(u, np.float32)), str(fnam).encode()):
        raise OSError(f"oracle_prtdat({fnam})")


def step(u, arith=F32_NOFMA, cx=0.1, cy=0.1, out=None):
    """One Jacobi update; `out` (default: a copy of u) keeps its ring."""
    u = np.ascontiguousarray(u, np.float32)
    nx, ny = u.shape
    if out is None:
        out = u.copy()
    lib().oracle_step(arith, nx, ny, cx, cy, _p(u), _p(out))
    return out


def max_abs_diff(a, b):
    return float(lib().oracle_max_abs_diff(a.size, _p(a), _p(b)))


def solve(grid, steps, arith=F32_NOFMA, cx=0.1, cy=0.1, conv_every=0, conv_eps=1e-3, conv_phase=0,
          step_convention=0):
    """Returns (final grid, updates_done, converged)."""
    g = np.array(grid, np.float32, order="C", copy=True)
    nx, ny = g.shape
    work = np.empty_like(g)
    done, conv = C.c_int(0), C.c_int(0)
    rc = lib().oracle_solve(arith, nx, ny, steps, cx, cy, conv_every, np.float32(conv_eps), conv_phase,
                            step_convention, _p(g), _p(work), C.byref(done), C.byref(conv))
    if rc:
        raise ValueError("oracle_solve: bad arguments")
    return g, done.value, bool(conv.value)


def threads():
    return lib().oracle_threads()


# 3-D 7-point grids: PARITY UNPINNED (the reference is 2-D only, see heat_oracle.c)
def inidat3d(nx, ny, nz):
    u = np.empty((nx, ny, nz), np.float32)
    lib().oracle_inidat3d(nx, ny, nz, _p(u))
    return u


def solve3d(grid, steps, arith=F32_NOFMA, cx=0.1, cy=0.1, cz=0.1, conv_every=0, conv_eps=1e-3, conv_phase=0, step_convention=0):
    """Returns (final grid, updates_done, converged) for a float32 grid of shape (nx, ny, nz)."""
    g = np.array(grid, np.float32, order="C", copy=True)
    nx, ny, nz = g.shape
    work = np.empty_like(g)
    done, conv = C.c_int(0), C.c_int(0)
    rc = lib().oracle_solve3d(arith, nx, ny, nz, steps, cx, cy, cz, conv_every, np.float32(conv_eps), conv_phase,
                              step_convention, _p(g), _p(work), C.byref(done), C.byref(conv))
    if rc:
        raise ValueError("oracle_solve3d: bad arguments")
    return g, done.value, bool(conv.value)


# double-precision grids: PARITY UNPINNED (no reference build exists for them, see heat_oracle.c)
def inidat_f64(nx, ny):
    u = np.empty((nx, ny), np.float64)
    lib().oracle_inidat_f64(nx, ny, u.ctypes.data_as(C.POINTER(C.c_double)))
    return u


def solve_f64(grid, steps, fused=False, cx=0.1, cy=0.1, conv_every=0, conv_eps=1e-3, conv_phase=0, step_convention=0):
    """Returns (final grid, updates_done, converged) for a float64 grid."""
    g = np.array(grid, np.float64, order="C", copy=True)
    nx, ny = g.shape
    work = np.empty_like(g)
    done, conv = C.c_int(0), C.c_int(0)
    dp = C.POINTER(C.c_double)
    rc = lib().oracle_solve_f64(int(bool(fused)), nx, ny, steps, cx, cy, conv_every, np.float32(conv_eps), conv_phase,
                                step_convention, g.ctypes.data_as(dp), work.ctypes.data_as(dp), C.byref(done), C.byref(conv))
    if rc:
        raise ValueError("oracle_solve_f64: bad arguments")
    return g, done.value, bool(conv.value)


# ----------------------------------------------------------------------------
# numpy restatement of one update (second, independent statement of M:168-174 /
# C:155-161 for the two unfused modes; numpy never contracts to FMA)
# ----------------------------------------------------------------------------
def step_numpy(u, arith=F32_NOFMA, cx=0.1, cy=0.1):
    u = np.asarray(u, np.float32)
    out = u.copy()
    c = u[1:-1, 1:-1]
    ns = u[2:, 1:-1] + u[:-2, 1:-1]
    ew = u[1:-1, 2:] + u[1:-1, :-2]
    cx32, cy32 = np.float32(cx), np.float32(cy)
    if arith == F32_NOFMA:
        two = np.float32(2.0)
        out[1:-1, 1:-1] = (c + cx32 * (ns - two * c)) + cy32 * (ew - two * c)
    elif arith == MPI_F64MIX:
        cd = c.astype(np.float64)
        r = cd + np.float64(cx32) * (ns.astype(np.float64) - 2.0 * cd) + np.float64(cy32) * (ew.astype(np.float64) - 2.0 * cd)
        out[1:-1, 1:-1] = r.astype(np.float32)
    else:
        raise ValueError("numpy has no fused multiply-add")
    return out


# ----------------------------------------------------------------------------
# the prebuilt, unmodified reference programs (oracle/_ref, see build_ref.py)
# ----------------------------------------------------------------------------
def _unlimit_stack():
    # the MPI program keeps whole grids on main()'s stack (M:87,280)
    try:
        resource.setrlimit(resource.RLIMIT_STACK, (resource.RLIM_INFINITY, resource.RLIM_INFINITY))
    except (ValueError, OSError):
        soft, hard = resource.getrlimit(resource.RLIMIT_STACK)
        resource.setrlimit(resource.RLIMIT_STACK, (hard, hard))


def _from_prtdat_order(raw, nx, ny):
    # prtdat writes iy = ny-1..0 (outer), ix = 0..nx-1 (inner): M:332-333
    return np.ascontiguousarray(raw.reshape(ny, nx)[::-1, :].T)


def ref_binary(kind, nx, ny, conv=False, omp=False, opt="O3", nofma=False):
    if kind == "mpi":
        name = f"mpi_heat_{nx}x{ny}" + ("_con" if conv else "") + ("_omp" if omp else "") + f"_{opt}"
    else:
        name = f"cuda_heat_{nx}x{ny}" + ("_con" if conv else "") + ("_nofma" if nofma else "")
    return REF / name


def run_reference(kind, nx, ny, steps, conv=False, omp=False, opt="O3", nofma=False, step=20,
                  capture=True, text=False, threads=None, ranks=1, timeout=3600):
    """Run a prebuilt reference binary in a scratch dir.  Returns a dict with
    stdout, wall seconds, loop_secs (MPI stub's loop stamp), converged_print (the
    number the program printed, or None) and -- if capture -- 'initial'/'final'
    arrays with the exact float bits prtdat was handed.  ranks > 1 (MPI program only): the stand-in
    forks that many processes, i.e. the reference's own 2-D block decomposition with its halo messages
    (nx and ny must be divisible by the rank grid MPI_Dims_create picks, as the reference assumes, M:72-73)."""
    import time
    exe = ref_binary(kind, nx, ny, conv, omp, opt, nofma)
    if not exe.exists():
        raise FileNotFoundError(f"{exe} (run oracle/build_ref.py where /root/reference exists)")
    env = dict(os.environ, HEAT_REF_STEPS=str(steps), HEAT_REF_STEP=str(step),
               HEAT_REF_BIN="1" if capture else "0", HEAT_REF_TEXT="1" if text else "0",
               HEAT_REF_RANKS=str(int(ranks)))
    if threads:
        env["OMP_NUM_THREADS"] = str(threads)
    tmp_root = "/dev/shm" if os.path.isdir("/dev/shm") else None
    with tempfile.TemporaryDirectory(dir=tmp_root, prefix="heat_ref_") as d:
        t0 = time.perf_counter()
        pr = subprocess.run([str(exe)], cwd=d, env=env, capture_output=True, text=True, timeout=timeout,
                            preexec_fn=_unlimit_stack)
        wall = time.perf_counter() - t0
        if pr.returncode != 0:
            raise RuntimeError(f"{exe.name} exited {pr.returncode}: {pr.stderr[-400:]}")
        res = {"stdout": pr.stdout, "wall_secs": wall, "loop_secs": None, "converged_print": None,
               "elapsed_print": None}
        m = re.search(r"HEAT_REF_LOOP_SECS ([0-9.]+)", pr.stdout)
        if m:
            res["loop_secs"] = float(m.group(1))
        m = re.search(r"Converged (?:after|at) (\d+) steps", pr.stdout)
        if m:
            res["converged_print"] = int(m.group(1))
        m = re.search(r"Elapsed time:? ([0-9.]+) (m?)secs", pr.stdout)
        if m:
            res["elapsed_print"] = float(m.group(1)) * (1e-3 if m.group(2) else 1.0)
        if capture or text:
            for f in Path(d).iterdir():
                key = "initial" if f.name.startswith("initial") else "final"
                if f.suffix == ".f32":
                    res[key] = _from_prtdat_order(np.fromfile(f, np.float32), nx, ny)
                elif text and f.suffix == ".dat":
                    import hashlib
                    res[key + "_text_md5"] = hashlib.md5(f.read_bytes()).hexdigest()
        return res
