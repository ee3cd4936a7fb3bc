#!/usr/bin/env python3
"""bench.py -- the solver-loop benchmark (BASELINE.json metric: GLUP/s and % of HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A bench "step" is ONE Jacobi update of the whole grid (one pass of the hot path).
Workloads (BASELINE.json configs):
    strong  (default)  65536 x 65536 float32 cut into N row strips (config[3]; at N=1 the whole grid on one GPU)
    c3                 16384 x 16384 with the convergence check every 20 updates (config[2], 1 GPU)
    c2                 8192 x 8192, no convergence check (config[1]); at N GPUs (8192*N) x 8192 in N strips
    weak32k            (32768*N) x 32768, temporal_block 4 (config[4])
    c1                 1024 x 1024 (config[0]'s grid; a parity case, here only for curiosity)
At N=1 the default run also measures c3, c2, weak32k and the MPI arithmetic (`also`), each the same way.

How a number is taken (every workload, every N):
    W warm-up updates, then one untimed repetition of the timed call (K updates), device sync, barrier;
    2 untimed updates so that flag-ordered neighbour strips are in lockstep; then R >= 7 repetitions of
    K updates back to back, each between its own pair of CUDA events on the launching stream; per
    repetition the MAX over ranks; `ms_per_step` = MEDIAN repetition / K.  Nothing but event records and
    the K-update call sits between the first and the last event (the clock sampler and every lazily
    built thing were started before the barrier).
`value` = LUPs of all ranks / that time, inputs resident in HBM.  `e2e` = the same through heat_solve()
with HOST buffers (H2D + K updates + D2H inside the timed region).  --impl reference times the
reference's own MPI program (oracle/_ref, unmodified source, OpenMP build, all host cores).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "GLUP/s (5-point Jacobi lattice-point updates per second, solver loop)"
BYTES_PER_LUP = 8.0  # float32 single-step: 4 B read + 4 B written per point-update
MIN_REPS, MAX_REPS, TARGET_REGION_MS = 7, 200, 250.0
LOCKSTEP_UPDATES = 2
PARITY_UPDATES = 24


def workload(name, n):
    if name == "strong":
        return dict(name=name, nx=65536, ny=65536, conv_every=0, temporal_block=1, scaling="strong",
                    text=f"65536x65536 float32 inidat field over {n} row strip(s), no convergence check")
    if name == "c2":
        return dict(name=name, nx=8192 * n, ny=8192, conv_every=0, temporal_block=1, scaling="weak",
                    text=f"{8192 * n}x8192 float32 inidat field, {n} row strip(s) of 8192 rows, no convergence check")
    if name == "c3":
        return dict(name=name, nx=16384, ny=16384, conv_every=20, temporal_block=1, scaling="weak",
                    text="16384x16384 float32 inidat field, convergence check every 20 updates")
    if name == "weak32k":
        return dict(name=name, nx=32768 * n, ny=32768, conv_every=0, temporal_block=4, scaling="weak",
                    text=f"{32768 * n}x32768 float32 inidat field, {n} strip(s), 4 updates per HBM pass")
    if name == "g16k":
        return dict(name=name, nx=16384, ny=16384, conv_every=0, temporal_block=1, scaling="weak",
                    text="16384x16384 float32 inidat field, no convergence check (A/B runs of kernel variants)")
    if name == "g16k_f64":
        return dict(name=name, nx=16384, ny=16384, conv_every=0, temporal_block=1, scaling="weak", f64=True,
                    text="16384x16384 DOUBLE-precision inidat field (HEAT_GRID_F64; 16 B per point-update), no convergence check")
    if name == "c1":
        return dict(name=name, nx=1024, ny=1024, conv_every=0, temporal_block=1, scaling="weak", text="1024x1024 float32 inidat field")
    raise SystemExit(f"unknown workload {name}")


def config_of(wl, arith_name, world, halo, cols):
    """The `config` object of a JSON line; both arms print the same keys."""
    nx, ny = wl["nx"], wl["ny"]
    return {"workload": wl["text"], "nx": nx, "ny": ny, "arith": arith_name,
            "conv_every": wl["conv_every"], "temporal_block": wl["temporal_block"],
            "halo": (halo if world > 1 else "none"),
            "decomposition": (f"{world // cols}x{cols} blocks" if world > 1 else "single"),
            "l2": f"working set {2 * nx * ny * 4 >> 20} MiB vs 126 MB L2: inputs larger than L2, no flush "
                  "(consecutive updates ARE the workload)"}


def measured_peak():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            return float(json.loads(f.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(name):
    """Per-launch DRAM bytes of a kernel from the committed ncu capture, if any (profiles/traffic.json)."""
    f = ROOT / "profiles" / "traffic.json"
    if f.exists():
        try:
            return json.loads(f.read_text()).get(name)
        except Exception:
            return None
    return None


def host_cores():
    """Cores this process may use (what `nproc` prints)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


class ClockSampler:
    """nvidia-smi clocks/throttle reasons (through NVML) sampled every few ms by a thread that is started
    BEFORE the warm-up; window() then keeps the samples that fall inside a timed region."""

    NAMES = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
             "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10, "applications_clocks_setting": 0x2}

    def __init__(self, index=0, period=0.004):
        self.index, self.period = index, period
        self.samples = []  # (host time, sm MHz, reasons bitmask, power W)
        self.max_mhz = None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _once(self):
        nv = self.nv
        try:
            mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            try:
                w = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
            except Exception:
                w = None
            self.samples.append((time.perf_counter(), mhz, r, w))
        except Exception:
            pass

    def _loop(self):
        while not self._stop.is_set():
            self._once()
            self._stop.wait(self.period)

    def start(self):
        if self.nv:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()
        return self

    def stop(self):
        if self._t:
            self._stop.set()
            self._t.join()
            self._t = None

    def nvlink_kib(self):
        """(tx, rx) KiB moved over all NVLinks of this GPU so far (NVML field values), or None."""
        nv = self.nv
        if nv is None:
            return None
        ids = (nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX)
        try:  # scope UINT_MAX = the sum over the links
            v = nv.nvmlDeviceGetFieldValues(self.h, [(ids[0], 0xFFFFFFFF), (ids[1], 0xFFFFFFFF)])
            if v[0].nvmlReturn == 0 and v[1].nvmlReturn == 0:
                return int(v[0].value.ullVal), int(v[1].value.ullVal)
        except Exception:
            pass
        try:  # else link by link
            tot = [0, 0]
            seen = False
            for link in range(18):
                v = nv.nvmlDeviceGetFieldValues(self.h, [(ids[0], link), (ids[1], link)])
                for k in range(2):
                    if v[k].nvmlReturn == 0:
                        tot[k] += int(v[k].value.ullVal)
                        seen = True
            return tuple(tot) if seen else None
        except Exception:
            return None

    def window(self, t0, t1):
        """Summary of the samples taken in [t0, t1] (host clock); the nearest one if the region was too short."""
        inside = [s for s in self.samples if t0 <= s[0] <= t1]
        if not inside and self.samples:
            mid = 0.5 * (t0 + t1)
            inside = [min(self.samples, key=lambda s: abs(s[0] - mid))]
        mhz = sorted(s[1] for s in inside)
        bits = 0
        for s in inside:
            bits |= s[2]
        watts = [s[3] for s in inside if s[3] is not None]
        return {"sm_mhz": (mhz[len(mhz) // 2] if mhz else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(k for k, b in self.NAMES.items() if bits & b), "samples": len(inside),
                "power_w_max": (max(watts) if watts else None)}


# ------------------------------------------------------------------------------------ reference arm
def rank_grid(n):
    """MPI_Dims_create(n, 2): the balanced factorisation, larger factor first (M:52)."""
    b = max(f for f in range(1, int(n ** 0.5) + 1) if n % f == 0)
    return n // b, b


def reference_layouts(nx, ny, cores):
    """Ways to spend `cores` host threads on the reference MPI program: (ranks, OpenMP threads per rank).
    One rank with an OpenMP team (-DOMPCH, M:163), or as many ranks as divide the grid the way the reference
    needs (M:72-73) with one thread each -- its own block decomposition, carried by the stand-in MPI."""
    outs = [(1, cores)]
    for r in range(cores, 1, -1):
        a, b = rank_grid(r)
        if nx % a == 0 and ny % b == 0:
            outs.append((r, 1))
            break
    return outs


def reference_grid(wl):
    """The grid the unmodified reference program runs for a workload, and why if it is not the workload's own.
    The program keeps 5 copies of the field in static/stack arrays (M:46,87,280), initialises it on one thread
    and walks it column-wise twice in prtdat (M:326-341, one fprintf per value): at 65536^2 that is 80 GiB and
    about ten minutes outside the loop, at 32768^2 20 GiB and two to three minutes.  Its loop speed does not
    depend on the grid size once the field is far larger than the caches, so those workloads are timed on
    16384 x 16384 (the largest grid whose whole program ends inside the time-box) and the line says so."""
    nx, ny = wl["nx"], wl["ny"]
    if nx * ny > 16384 * 16384 * 2 and wl["conv_every"] == 0:
        return 16384, 16384, (f"reference timed on 16384x16384 instead of {nx}x{ny}: its setup (single-threaded inidat, two "
                              "column-wise prtdat walks, 5 field copies in static/stack arrays) does not fit the time-box "
                              "at that size; its loop GLUP/s is size-independent beyond the caches")
    return nx, ny, None


def time_reference(O, nx, ny, conv, cores, loop_budget_secs, max_updates):
    """Times the unmodified reference MPI program (oracle/_ref) on a bounded sample of the workload: every
    layout is probed with 3 updates, the fastest one runs the sample.  Returns (glups, updates, secs, text)."""
    probed = []
    for ranks, threads in reference_layouts(nx, ny, cores):
        try:
            probe = O.run_reference("mpi", nx, ny, 2, conv=conv, omp=True, opt="O3", capture=False, threads=threads, ranks=ranks,
                                    timeout=900)
        except Exception:
            if ranks == 1:
                raise
            continue  # a rank layout that cannot run here (fork or shared memory refused) is skipped
        probed.append((max(probe["loop_secs"] / 3, 1e-6), ranks, threads))
    r = None
    for per_update, ranks, threads in sorted(probed):  # fastest first; the next one if the sample run fails
        updates = int(max(3, min(max_updates, loop_budget_secs / per_update)))
        try:
            r = O.run_reference("mpi", nx, ny, updates - 1, conv=conv, omp=True, opt="O3", capture=False, threads=threads, ranks=ranks,
                                timeout=1800)
            break
        except Exception:
            if (per_update, ranks, threads) == sorted(probed)[-1]:
                raise
    a, b = rank_grid(ranks)
    how = (f"{ranks} ranks ({a}x{b} blocks; forked processes over a shared-memory MPI stand-in -- a shim, not real MPI) x "
           f"{threads} OpenMP thread(s)" if ranks > 1 else f"1 rank x {threads} OpenMP threads (single-rank MPI stand-in)")
    text = (f"{updates} updates of {nx}x{ny}: unmodified mpi_heat_improved_persistent_stat.c, -O3 -fopenmp -DOMPCH, {how}, "
            f"the faster of {reference_layouts(nx, ny, cores)} (ranks, threads) layouts, loop-only time (M:159-265)")
    return (nx - 2) * (ny - 2) * updates / r["loop_secs"] / 1e9, updates, r["loop_secs"], text


def reference_sample(wl, cores, loop_budget_secs, max_updates):
    """(glups, updates, secs, kind, sample text, reference_workload note) for one workload on the host cores."""
    os.environ["OMP_NUM_THREADS"] = str(cores)  # torchrun exports OMP_NUM_THREADS=1 to its workers
    from oracle import pyoracle as O
    nx, ny, note = reference_grid(wl)
    conv = wl["conv_every"] > 0
    exe = O.ref_binary("mpi", nx, ny, conv=conv, omp=True, opt="O3")
    if exe.exists():
        glups, updates, secs, sample = time_reference(O, nx, ny, conv, cores, loop_budget_secs, max_updates)
        return glups, updates, secs, "reference", sample, note, (nx, ny), cores
    # no prebuilt reference binary for this grid: the C restatement (oracle/heat_oracle.c), same arithmetic
    if nx * ny > 16384 * 16384:
        nx = ny = 16384
        note = f"oracle port timed on 16384x16384 instead of {wl['nx']}x{wl['ny']} (bounded sample)"
    ini = O.inidat(nx, ny)
    t0 = time.perf_counter()
    O.solve(ini, 2, arith=O.MPI_F64MIX)
    per_update = max((time.perf_counter() - t0) / 2, 1e-6)
    updates = int(max(3, min(max_updates, loop_budget_secs / per_update)))
    t0 = time.perf_counter()
    O.solve(ini, updates, arith=O.MPI_F64MIX)
    secs = time.perf_counter() - t0
    threads = O.threads()
    glups = (nx - 2) * (ny - 2) * updates / secs / 1e9
    sample = f"{updates} updates of {nx}x{ny}: oracle/heat_oracle.c (C restatement, MPI_F64MIX), {threads} OpenMP threads"
    return glups, updates, secs, "port", sample, note, (nx, ny), threads


def run_reference_arm(args, wl, rank, world):
    if rank != 0:
        return
    cores = host_cores()
    glups, updates, secs, kind, sample, note, (rnx, rny), used = reference_sample(wl, cores, 20.0, args.steps + max(args.warmup, 3))
    cfg = config_of(dict(wl, nx=rnx, ny=rny), "MPI_F64MIX", 1, "none", 1)  # the grid that was really timed
    out = {"metric": METRIC, "value": glups, "unit": "GLUP/s", "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": secs / updates * 1e3, "higher_is_better": True, "scaling": wl["scaling"],
           "vs_baseline": None, "dtype": "f32 grid, f64-mixed arithmetic (M:168-174)", "data": "synthetic", "config": cfg,
           "reference_workload": note,
           "cpu_baseline": {"value": glups, "unit": "GLUP/s", "cores": used, "kind": kind, "sample": sample},
           "e2e": {"value": glups, "unit": "GLUP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


# ------------------------------------------------------------------------------------ cpu baseline (rank 0, N=1)
def cpu_baseline(wl):
    cores = host_cores()
    try:
        glups, updates, secs, kind, sample, note, _, used = reference_sample(wl, cores, 12.0, 2000)
        if note:
            sample += " [" + note + "]"
        return {"value": glups, "unit": "GLUP/s", "cores": used, "kind": kind, "sample": sample}
    except Exception as e:  # a baseline that cannot run must not take the GPU number down with it
        return {"value": None, "unit": "GLUP/s", "cores": cores, "kind": "unavailable", "sample": repr(e)[:200]}


def reference_cuda_leg(peak):
    """The reference's own GPU program on this GPU (north_star: "... and the reference's own cuda_heat.cu"):
    unmodified cuda/cuda_heat.cu built for sm_100a at its largest grid that the 1-GPU configs name (8192^2).
    Its own timer stops without a device sync (C:239), so the loop time is wall(STEPS=S) - wall(STEPS=0)."""
    try:
        from oracle import pyoracle as O
        nx = ny = 8192
        if not O.ref_binary("cuda", nx, ny).exists():
            return {"value": None, "note": "oracle/_ref/cuda_heat_8192x8192 not built"}
        S = 2000  # 2000 updates = 2-3 s of loop against a process wall of several seconds; the smaller of two walls each
        w0 = min(O.run_reference("cuda", nx, ny, 0, capture=False, timeout=600)["wall_secs"] for _ in range(2))
        w1 = min(O.run_reference("cuda", nx, ny, S, capture=False, timeout=600)["wall_secs"] for _ in range(2))
        secs = max(w1 - w0, 1e-6)
        v = (nx - 2) * (ny - 2) * S / secs / 1e9
        return {"value": v, "unit": "GLUP/s", "ms_per_step": secs / S * 1e3, "steps": S, "frac": v * BYTES_PER_LUP / peak,
                "note": "unmodified cuda/cuda_heat.cu (C:140-163 kernel, C:200-211 launch loop), nvcc -O2 sm_100a, 8192x8192, "
                        f"min-of-2 wall(STEPS={S}) - min-of-2 wall(STEPS=0); a reported baseline, not a target"}
    except Exception as e:
        return {"value": None, "note": "reference cuda_heat leg failed: " + repr(e)[:160]}


# ------------------------------------------------------------------------------------ our arm
class Ctx:
    """What every measurement needs: torch, the process group, the stream, the clock sampler."""

    def __init__(self, torch, dist, rank, world, local_rank, stream, clocks):
        self.torch, self.dist, self.rank, self.world, self.local_rank = torch, dist, rank, world, local_rank
        self.stream, self.clocks = stream, clocks

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize()


def measure(ctx, enqueue, sync, K, W, reps=None, launch_count=lambda: 0, before_barrier=None, after_timed=None):
    """The timing scheme of the module docstring.  Returns a dict with the per-repetition times (max over ranks)."""
    torch = ctx.torch
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    enqueue(W)
    w0, w1 = ev(), ev()
    w0.record(ctx.stream)
    enqueue(K)  # the timed call itself, once, untimed: every lazily built thing exists afterwards
    w1.record(ctx.stream)
    sync()
    warm_ms = w0.elapsed_time(w1)
    if ctx.world > 1:
        t = torch.tensor([warm_ms], device="cuda")
        ctx.dist.all_reduce(t, op=ctx.dist.ReduceOp.MAX)
        warm_ms = float(t.item())
    R = reps or int(min(MAX_REPS, max(MIN_REPS, math.ceil(TARGET_REGION_MS / max(warm_ms, 1e-3)))))
    events = [ev() for _ in range(R + 1)]
    if before_barrier:
        before_barrier()
    ctx.barrier()
    enqueue(LOCKSTEP_UPDATES)  # neighbours wait for each other's flags here, not inside the first repetition
    n0 = launch_count()
    t0 = time.perf_counter()
    events[0].record(ctx.stream)
    for r in range(R):
        enqueue(K)
        events[r + 1].record(ctx.stream)
    sync()
    t1 = time.perf_counter()
    n1 = launch_count()
    extra = after_timed() if after_timed else None
    ctx.barrier()
    mine = [events[r].elapsed_time(events[r + 1]) for r in range(R)]
    per_rank_median = [sorted(mine)[R // 2]]
    times = mine
    if ctx.world > 1:
        t = torch.tensor(mine, device="cuda", dtype=torch.float64)
        ctx.dist.all_reduce(t, op=ctx.dist.ReduceOp.MAX)
        times = [float(x) for x in t.tolist()]
        g = torch.zeros(ctx.world, device="cuda", dtype=torch.float64)
        g[ctx.rank] = per_rank_median[0]
        ctx.dist.all_reduce(g)
        per_rank_median = [float(x) for x in g.tolist()]
    s = sorted(times)
    return {"reps": R, "K": K, "median_ms": s[R // 2], "best_ms": s[0], "worst_ms": s[-1], "mean_ms": sum(times) / R,
            "per_rank_median_ms": per_rank_median, "warm_rep_ms": warm_ms, "clocks": ctx.clocks.window(t0, t1),
            "host_secs": t1 - t0, "launches": n1 - n0, "extra": extra}


def device_window(torch, ptr, pitch, r0, r1, c0, c1, device, es=4):
    """Rows [r0, r1) x columns [c0, c1) of a strip's local device array (pitch in elements of es bytes), as numpy."""
    class _Arr:
        pass
    a = _Arr()
    a.__cuda_array_interface__ = {"shape": (r1 - r0, c1 - c0), "strides": (es * pitch, es), "typestr": f"<f{es}", "version": 3,
                                  "data": (ptr + es * (pitch * r0 + c0), False)}
    return torch.as_tensor(a, device=f"cuda:{device}").cpu().numpy()


def inidat_window(nx, ny, r0, r1, c0, c1, dtype=None):
    """inidat (M:320, int32 wrap included) on a window of the global grid; dtype float64 for double grids (the int
    product converts exactly there, it is rounded only for float)."""
    import numpy as np
    dtype = dtype or np.float32
    ix = np.arange(r0, r1, dtype=np.uint32)[:, None]
    iy = np.arange(c0, c1, dtype=np.uint32)[None, :]
    prod = ix * (np.uint32(nx) - ix - np.uint32(1)) * iy * (np.uint32(ny) - iy - np.uint32(1))
    return prod.view(np.int32).astype(dtype)


def parity_windows(ctx, wl, arith, view, updates, get):
    """Bit-exact check of what the timed code path computes, outside the timed region.  After `updates` updates of
    the inidat field a point depends only on initial values at most `updates` rows/columns away, so a small
    window can be checked against the oracle run on the window plus that margin (cut edges of the oracle's
    window are wrong by then only within the margin, which is discarded; where the window touches the real
    ring the oracle's fixed ring IS the ring).  Windows: this rank's first rows (strip boundary, ghost rows
    included), middle, last rows  x  first columns, a CTA-column boundary in the middle, last columns.
    view = (lo, nloc, lo_c, nyloc): the global rows / columns this rank's local array holds;
    get(r0, r1, c0, c1) -> numpy window of the newest local array in GLOBAL coordinates."""
    import numpy as np
    from oracle import pyoracle as O
    nx, ny, U = wl["nx"], wl["ny"], updates
    lo, nloc, lo_c, nyloc = view
    hi, hi_c = lo + nloc, lo_c + nyloc
    H, Wd = 40, 64
    rows = {(lo, min(lo + H, hi)), (max(lo, (lo + hi) // 2 - H // 2), min(hi, (lo + hi) // 2 + H // 2)), (max(lo, hi - H), hi)}
    mid_c = max(lo_c, min(hi_c - Wd, (lo_c + hi_c) // 2 // 512 * 512 - Wd // 2))
    cols = {(lo_c, min(lo_c + Wd, hi_c)), (mid_c, min(mid_c + Wd, hi_c)), (max(lo_c, hi_c - Wd), hi_c)}
    n_win = n_pts = n_bad = 0
    for (r0, r1) in sorted(rows):
        for (c0, c1) in sorted(cols):
            R0, R1, C0, C1 = max(0, r0 - U), min(nx, r1 + U), max(0, c0 - U), min(ny, c1 + U)
            ini = inidat_window(nx, ny, R0, R1, C0, C1, np.float64 if wl.get("f64") else np.float32)
            if wl.get("f64"):  # double grids: the oracle's double mode (parity unpinned, see oracle/heat_oracle.c)
                want, _, _ = O.solve_f64(ini, U, fused=(arith == 0))
            else:
                want, _, _ = O.solve(ini, U, arith=arith)
            want = want[r0 - R0:r1 - R0, c0 - C0:c1 - C0]
            got = get(r0, r1, c0, c1)
            bits = np.uint64 if wl.get("f64") else np.uint32
            n_bad += int((got.view(bits) != want.view(bits)).sum())
            n_win += 1
            n_pts += got.size
    return n_win, n_pts, n_bad


def run_one(ctx, api, wl, arith, args):
    """One workload on this rank's GPU: build the solver, measure, check parity.  Returns a dict."""
    torch, dist, world = ctx.torch, ctx.dist, ctx.world
    nx, ny = wl["nx"], wl["ny"]
    T = wl["temporal_block"]
    params = dict(arith=arith, conv_every=wl["conv_every"], temporal_block=T)
    es = 4
    if wl.get("f64"):
        params["grid_type"] = api.GRID_F64
        es = 8
    bytes_per_lup = BYTES_PER_LUP * es / 4
    s = None
    if args.k3_profile:
        os.environ["HEAT_PROFILE_WAIT"] = "1"  # read by heat_create: the edge CTAs time their flag waits
    if world > 1:
        from parallel_heat_b200.strips import StripSolver
        s = StripSolver(nx, ny, device=ctx.local_rank, halo=args.halo, cols=args.cols, **params)
        solver = s.solver
        if args.halo == "p2p":
            solver.set_stream(ctx.stream.cuda_stream)
        init, enqueue, sync = s.init_inidat, s.enqueue, s.sync
        rows_own, cols_own = s.rows, s.ncols
        lo_c = s.col0 - 1 if (args.cols > 1 and s.by > 0) else 0
        hi_c = s.col0 + s.ncols + 1 if (args.cols > 1 and s.by < args.cols - 1) else ny
    else:
        solver = api.HeatSolver(nx, ny, device=ctx.local_rank, **params)
        solver.set_stream(ctx.stream.cuda_stream)
        init, enqueue, sync = solver.init_inidat, solver.enqueue, solver.sync
        rows_own, cols_own, lo_c, hi_c = nx, ny, 0, ny
    try:
        init()
        K, W = args.steps, max(args.warmup, 3)
        if T > 1:
            K = max(K // T, 1) * T
        if wl["conv_every"] > 0:
            # the convergence-check loop is host-driven every conv_every updates: time heat_run itself
            def enqueue(n):  # noqa: F811
                solver.run(n)
        k3 = {}

        def k3_begin():
            if args.k3_profile and world > 1:
                solver.wait_stats()  # clears the counters (synchronises the stream: before the barrier, not after)
                k3["nvl0"] = ctx.clocks.nvlink_kib()

        def k3_end():
            if args.k3_profile and world > 1:
                k3["nvl1"] = ctx.clocks.nvlink_kib()
                k3["wait"] = solver.wait_stats()
        m = measure(ctx, enqueue, sync, K, W, reps=args.reps or None, launch_count=solver.launch_count,
                    before_barrier=k3_begin, after_timed=k3_end)
        launches = m["launches"]  # this rank's kernel launches between the first and the last timed event
        ms = m["median_ms"]
        lups_per_update = (nx - 2) * (ny - 2)
        value = lups_per_update * K / (ms * 1e-3) / 1e9

        # roofline of the dominant kernel (the update kernel): this rank's algorithmic bytes per launch
        peak, peak_src = measured_peak()
        step_launches = K // T
        own_lups = (min(rows_own, nx - 2) if world > 1 else nx - 2) * min(cols_own, ny - 2) * T
        kernel_ms = ms / max(step_launches, 1)
        achieved = bytes_per_lup * own_lups / (kernel_ms * 1e-3) / 1e9
        staged = ny >= 16384 and arith != api.MPI_F64MIX and world == 1 and os.environ.get("HEAT_K1_VARIANT", "auto") in ("", "auto")
        staged = staged or os.environ.get("HEAT_K1_VARIANT") == "bulk"
        kname = "heat_step_tb" if T > 1 else "heat_step_f64" if es == 8 else "heat_step_bulk" if staged else "heat_step_vec"
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": ncu_traffic(f"{kname}@{nx}x{ny}"), "kernel": kname, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": bytes_per_lup * own_lups,
                    "note": f"{bytes_per_lup:.0f} B per point-update x interior points of this GPU's strip / (median repetition / launches per repetition)"
                            + (" (effective: temporal blocking re-uses tiles on chip)" if T > 1 else "")}

        # parity of this very code path (same solver object, same kernels, same neighbours), outside the timed region
        parity = None
        if not args.no_parity:
            init()
            U = PARITY_UPDATES
            enqueue(U)  # the timed call (heat_run when there is a convergence check, else heat_enqueue)
            sync()
            ctx.barrier()
            ptr, pitch, nloc, lo = solver.device_view(0)

            def get(r0, r1, c0, c1):
                return device_window(torch, ptr, pitch, r0 - lo, r1 - lo, c0 - lo_c, c1 - lo_c, ctx.local_rank, es)
            # 2-D blocks keep their ghost COLUMNS in dense side arrays, not in the field: owned columns only there
            vc0, vc1 = (s.col0, s.col0 + s.ncols) if (world > 1 and args.cols > 1) else (lo_c, hi_c)
            nw, npts, nbad = parity_windows(ctx, wl, arith, (lo, nloc, vc0, vc1 - vc0), U, get)
            if world > 1:
                t = torch.tensor([nw, npts, nbad], device="cuda", dtype=torch.float64)
                dist.all_reduce(t)
                nw, npts, nbad = (int(x) for x in t.tolist())
            parity = {"windows": nw, "points": npts, "mismatches": nbad, "bit_exact": nbad == 0, "updates": U,
                      "against": "oracle.solve on each window + margin (40x64 windows: strip boundaries incl. ghost rows, "
                                 "middle, field edges), after re-initialising and running the same enqueue path"}
        k3_out = None
        if args.k3_profile and world > 1:
            # K3 evidence, per rank: how long the edge CTAs waited for the neighbours' flags and what went over NVLink
            passes = (m["reps"] * K + LOCKSTEP_UPDATES) // T
            ns, cnt = k3["wait"]
            mine = [float(ctx.rank), ms / max(step_launches, 1) * 1e3] + [float(x) / max(passes, 1) / 1e3 for x in ns[:2]] + \
                   [float(x) / max(passes, 1) for x in cnt[:2]]
            if k3.get("nvl0") and k3.get("nvl1"):
                mine += [(k3["nvl1"][0] - k3["nvl0"][0]) / max(passes, 1), (k3["nvl1"][1] - k3["nvl0"][1]) / max(passes, 1)]
            else:
                mine += [-1.0, -1.0]
            g = torch.zeros((world, len(mine)), device="cuda", dtype=torch.float64)
            g[ctx.rank] = torch.tensor(mine, dtype=torch.float64)
            dist.all_reduce(g)
            k3_out = {"per_rank": [{"rank": int(r[0]), "pass_us": r[1], "edge_wait_us_per_pass_north": r[2],
                                    "edge_wait_us_per_pass_south": r[3], "edge_waits_per_pass_north": r[4],
                                    "edge_waits_per_pass_south": r[5], "nvlink_tx_kib_per_pass": r[6], "nvlink_rx_kib_per_pass": r[7]}
                                   for r in g.tolist()],
                      "note": "edge_wait = SUM over the edge CTAs of one pass of the time their thread 0 spun on the neighbour's step "
                              "flag (globaltimer; a pass has edge_waits CTAs that wait, the other CTAs never do); nvlink = NVML "
                              "NVLINK_THROUGHPUT_DATA counters over the timed region / passes (ghost rows + flags; expected "
                              f"{T * ny * 4 / 1024:.0f} KiB per neighbour and pass)"}
        out = {"value": value, "unit": "GLUP/s", "ms_per_step": ms / K, "steps": K, "reps": m["reps"], "k3": k3_out,
               "rep_ms": {"best": m["best_ms"], "median": m["median_ms"], "worst": m["worst_ms"], "mean": m["mean_ms"]},
               "per_rank_median_ms": m["per_rank_median_ms"], "roofline": roofline, "clocks": m["clocks"],
               "gpu_launches": int(launches), "parity": parity}
        return out
    finally:
        if s is not None:
            s.close()
        else:
            solver.close()


def leg_3d(ctx, api, n, arith, args, peak):
    """the 7-point update on an n^3 float grid, measured like every other leg; parity = three
    slabs of planes against the oracle's 3-D mode (unpinned: the reference is 2-D only)."""
    import numpy as np
    from oracle import pyoracle as O
    torch = ctx.torch
    with api.HeatSolver(n, n, nz=n, device=ctx.local_rank, arith=arith) as s:
        s.set_stream(ctx.stream.cuda_stream)
        s.init_inidat()
        K = args.steps
        m = measure(ctx, s.enqueue, s.sync, K, max(args.warmup, 3), reps=args.reps or None, launch_count=s.launch_count)
        ms = m["median_ms"]
        lups = (n - 2) ** 3
        value = lups * K / (ms * 1e-3) / 1e9
        s.init_inidat()
        U = 8
        s.enqueue(U)
        s.sync()
        ptr, pitch, nloc, lo = s.device_view(0)
        nbad = npts = 0
        for x0 in (0, n // 2, n - 16):
            got = device_window(torch, ptr, pitch, x0, x0 + 16, 0, n * n, ctx.local_rank).reshape(16, n, n)
            a, b = max(0, x0 - U), min(n, x0 + 16 + U)
            ix = np.arange(a, b, dtype=np.uint32)[:, None, None]
            iy = np.arange(n, dtype=np.uint32)[None, :, None]
            iz = np.arange(n, dtype=np.uint32)[None, None, :]
            one, nn = np.uint32(1), np.uint32(n)
            ini = (ix * (nn - ix - one) * iy * (nn - iy - one) * iz * (nn - iz - one)).view(np.int32).astype(np.float32)
            want = O.solve3d(ini, U, arith=arith)[0][x0 - a:x0 - a + 16]
            nbad += int((got.view(np.uint32) != want.view(np.uint32)).sum())
            npts += got.size
    achieved = BYTES_PER_LUP * lups / (ms / K * 1e-3) / 1e9
    return {"value": value, "unit": "GLUP/s", "ms_per_step": ms / K, "steps": K, "reps": m["reps"],
            "rep_ms": {"best": m["best_ms"], "median": m["median_ms"], "worst": m["worst_ms"]},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "kernel": "heat_step_3d", "traffic": ncu_traffic(f"heat_step_3d@{n}x{n}x{n}"),
                         "algorithmic_bytes_per_launch": BYTES_PER_LUP * lups},
            "clocks": m["clocks"], "gpu_launches": m["launches"],
            "parity": {"windows": 3, "points": npts, "mismatches": nbad, "bit_exact": nbad == 0, "updates": U,
                       "against": "oracle.solve3d on three 16-plane slabs + margin (unpinned: the reference is 2-D only)"},
            "workload": f"{n}x{n}x{n} float32 3-D inidat field, 7-point update, no convergence check", "arith": api.ARITH_NAMES[arith],
            "note": "SURVEY 8(f) rank 4: 3-D 7-point variant, 8 B per point-update"}


def e2e_single(ctx, api, wl, arith, K):
    """heat_solve() on pinned HOST memory: H2D + K updates + D2H inside the timed region, wall clock, median of 3."""
    import ctypes as C
    torch = ctx.torch
    nx, ny, T = wl["nx"], wl["ny"], wl["temporal_block"]
    params = dict(arith=arith, conv_every=wl["conv_every"], temporal_block=T)
    host = torch.empty((nx, ny), dtype=torch.float32, pin_memory=True)
    start = torch.empty((nx, ny), dtype=torch.float32, pin_memory=True) if nx * ny <= 16384 * 16384 else None
    hn = host.numpy()
    with api.HeatSolver(nx, ny, device=ctx.local_rank, arith=arith) as tmp:  # the inidat field, made on the device
        tmp.init_inidat()
        tmp.download(hn)
    if start is not None:
        start.copy_(host)
    p = api.default_params(nx=nx, ny=ny, steps=K, device=ctx.local_rank, **params)
    lib = api.load()
    secs, done = [], C.c_int(0)
    conv = C.c_int(0)
    runs = 3
    for k in range(runs + 1):  # the first call is a warm-up (context pools, pinned-page mappings)
        if k > 0:
            if start is not None:
                host.copy_(start)
            else:  # 65536^2: re-create the field in place instead of keeping a second 16 GiB pinned copy
                with api.HeatSolver(nx, ny, device=ctx.local_rank, arith=arith) as tmp:
                    tmp.init_inidat()
                    tmp.download(hn)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rc = lib.heat_solve(C.byref(p), hn.ctypes.data_as(C.POINTER(C.c_float)), C.byref(done), C.byref(conv))
        dt = time.perf_counter() - t0
        if rc != 0:
            raise RuntimeError(lib.heat_last_error().decode())
        if k > 0:
            secs.append(dt)
    dt = sorted(secs)[len(secs) // 2]
    api.trim()  # give the kept buffers back before the other legs allocate theirs
    # the host result of the last call, checked like the device result
    import numpy as np
    from oracle import pyoracle as O  # noqa: F401
    nw, npts, nbad = parity_windows(ctx, wl, arith, (0, nx, 0, ny), done.value, lambda r0, r1, c0, c1: np.ascontiguousarray(hn[r0:r1, c0:c1]))
    lups = (nx - 2) * (ny - 2)
    return {"value": lups * done.value / dt / 1e9, "unit": "GLUP/s",
            "h2d_bytes_per_step": nx * ny * 4 / max(done.value, 1), "d2h_bytes_per_step": nx * ny * 4 / max(done.value, 1),
            "call": f"heat_solve(steps={K}) on pinned host memory: H2D of the {nx * ny * 4 >> 20} MiB field + {done.value} updates + D2H "
                    f"(the three overlap band by band), wall clock, median of {runs} calls after one warm-up call; the library keeps "
                    "the two device buffers of the previous heat_solve of the same size (heat_trim() frees them), so the timed calls "
                    "do not pay cudaMalloc/cudaFree -- the warm-up call did",
            "seconds": dt, "seconds_all": secs, "pcie_gbs_each_way": nx * ny * 4 / dt / 1e9,
            "parity": {"windows": nw, "points": npts, "mismatches": nbad, "bit_exact": nbad == 0, "updates": done.value}}


def e2e_strips(ctx, api, wl, arith, args):
    """N GPUs, field in HOST memory: every rank holds the rows of its strip plus K more on either side (pinned),
    and calls heat_solve() on that window -- H2D, K updates and D2H overlapped band by band, no exchange with the
    neighbours (the K outer rows are recomputed on both sides of a cut instead; strips.solve_strip_on_host_rows).
    Wall clock between barriers, so the slowest rank counts."""
    import numpy as np
    torch, dist = ctx.torch, ctx.dist
    from parallel_heat_b200 import strips
    nx, ny, T = wl["nx"], wl["ny"], wl["temporal_block"]
    K = max(args.steps // T, 1) * T
    lo, hi = strips.overlap_window(nx, ctx.world, ctx.rank, K)
    row0, rows = api.strip_rows(nx, ctx.world, ctx.rank)
    host = torch.empty((hi - lo, ny), dtype=torch.float32, pin_memory=True)
    hn = host.numpy()
    secs, own = [], None
    for k in range(3):  # first round = warm-up
        hn[...] = inidat_window(nx, ny, lo, hi, 0, ny)
        ctx.barrier()
        t0 = time.perf_counter()
        own, done = strips.solve_strip_on_host_rows(hn, nx, ctx.world, ctx.rank, K, device=ctx.local_rank, arith=arith,
                                                    temporal_block=T)
        ctx.barrier()
        if k > 0:
            secs.append(time.perf_counter() - t0)
    nw, npts, nbad = parity_windows(ctx, wl, arith, (row0, rows, 0, ny), K,
                                    lambda r0, r1, c0, c1: np.ascontiguousarray(own[r0 - row0:r1 - row0, c0:c1]))
    dt = sorted(secs)[len(secs) // 2]
    tot = torch.tensor([float((hi - lo) * ny * 4), float((hi - lo) * ny * 4), float(nw), float(npts), float(nbad)], device="cuda",
                       dtype=torch.float64)
    mx = torch.tensor([dt], device="cuda", dtype=torch.float64)
    dist.all_reduce(tot)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    dt = float(mx.item())
    tot = tot.tolist()
    return {"value": (nx - 2) * (ny - 2) * K / dt / 1e9, "unit": "GLUP/s",
            "h2d_bytes_per_step": tot[0] / K, "d2h_bytes_per_step": tot[1] / K,
            "call": f"per rank: heat_solve(steps={K}) on its pinned host window = own rows + {K} rows either side (no halo exchange: "
                    "the overlap is recomputed), H2D + updates + D2H overlapped band by band; wall clock between barriers, "
                    "median of 2 rounds after one warm-up round, max over ranks", "seconds": dt,
            "parity": {"windows": int(tot[2]), "points": int(tot[3]), "mismatches": int(tot[4]), "bit_exact": tot[4] == 0, "updates": K}}


def run_ours(args, wl, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    from parallel_heat_b200 import api, build
    if rank == 0 and build.stale():
        build.build()
    if world > 1:
        dist.barrier()
    api.load()
    torch.cuda.set_device(local_rank)
    arith = {"fma": api.F32_FMA, "nofma": api.F32_NOFMA, "f64mix": api.MPI_F64MIX}[args.arith]
    stream = torch.cuda.current_stream()
    clocks = ClockSampler(local_rank).start()  # NVML initialised here, long before any timed region
    ctx = Ctx(torch, dist, rank, world, local_rank, stream, clocks)
    peak, _ = measured_peak()
    if args.only_3d:
        r = leg_3d(ctx, api, args.only_3d, arith, args, peak)
        clocks.stop()
        print(json.dumps(r), flush=True)
        return 0 if r["parity"]["bit_exact"] else 3
    try:
        main = run_one(ctx, api, wl, arith, args)

        e2e = None
        if not args.no_e2e:
            if world == 1:
                try:
                    e2e = e2e_single(ctx, api, wl, arith, max(args.steps // wl["temporal_block"], 1) * wl["temporal_block"])
                except (MemoryError, RuntimeError) as e:  # e.g. the host cannot pin a second copy of the field
                    e2e = {"value": None, "unit": "GLUP/s", "h2d_bytes_per_step": None, "d2h_bytes_per_step": None,
                           "error": repr(e)[:300]}
            elif args.cols == 1:
                e2e = e2e_strips(ctx, api, wl, arith, args)

        # the other 1-GPU configurations of BASELINE.json, measured the same way, reported next to the headline
        also = None
        if world == 1 and not args.no_also:
            also = {}

            def leg(key, name, arith_, note, tb=None):
                w = workload(name, 1)
                if tb:
                    w["temporal_block"] = tb
                try:  # a side leg that cannot run (memory, a missing binary) must not take the headline down with it
                    r = run_one(ctx, api, w, arith_, args)
                except Exception as e:
                    also[key] = {"value": None, "error": repr(e)[:300], "workload": w["text"]}
                    return
                r["workload"] = w["text"] + (f", {tb} updates per HBM pass" if tb else "")
                r["arith"] = api.ARITH_NAMES[arith_]
                r["note"] = note
                if w["temporal_block"] > 1:
                    r["effective_gbs"] = r["value"] * BYTES_PER_LUP
                    r["effective_frac"] = r["value"] * BYTES_PER_LUP / peak
                also[key] = r

            if wl["name"] != "c3":
                leg("c3_16384", "c3", arith, "BASELINE configs[2]: convergence check every 20 updates fused into the update "
                    "kernel (K2), one 4-byte read-back per check; north_star's >= 80 % target is quoted on this grid")
            if wl["name"] != "c2":
                leg("c2_8192", "c2", arith, "BASELINE configs[1]: 8192x8192, no convergence check")
            if wl["name"] != "weak32k":
                leg("weak32k_T4", "weak32k", arith, "BASELINE configs[4] at 1 GPU: heat_step_tb<T=4>, 4 updates per pass, same bits "
                    "as 4 single updates; 8 B/LUP counted as for single-step, so frac is EFFECTIVE bandwidth")
            if wl["temporal_block"] == 1 and wl["conv_every"] == 0:
                leg(wl["name"] + "_T4", wl["name"], arith, "the headline workload with 4 updates fused per HBM pass (bit-identical)", tb=4)
            if arith != api.MPI_F64MIX:
                leg("mpi_f64mix", wl["name"], api.MPI_F64MIX, "the headline workload in the MPI program's mixed float/double "
                    "arithmetic (M:168-174): bit-identical to what the --impl reference arm computes")
                if wl["name"] != "c2":
                    leg("mpi_f64mix_c2_8192", "c2", api.MPI_F64MIX, "8192x8192 in the MPI program's arithmetic")
            leg("f64_grid_16384", "g16k_f64", arith, "SURVEY 8(f) rank 4: the same update on a DOUBLE-precision grid (16 B per point-update, "
                "so the HBM roofline is half as many LUP/s); parity unpinned -- the reference has no double build")
            try:
                also["grid3d_640"] = leg_3d(ctx, api, 640, arith, args, peak)
            except Exception as e:
                also["grid3d_640"] = {"value": None, "error": repr(e)[:300]}
            also["reference_cuda"] = reference_cuda_leg(peak)

        if world > 1 and not args.no_also and args.cols == 1 and wl["name"] != "weak32k":
            # BASELINE configs[4] on the same N GPUs: weak scaling, 32768^2 per GPU, 4 updates per HBM pass
            w = workload("weak32k", world)
            r = run_one(ctx, api, w, arith, args)
            r.update(workload=w["text"], arith=api.ARITH_NAMES[arith], scaling="weak", effective_gbs=r["value"] * BYTES_PER_LUP,
                     effective_frac_per_gpu=r["value"] * BYTES_PER_LUP / peak / world,
                     note="weak scaling of configs[4]: compare with also.weak32k_T4 of the 1-GPU line")
            also = {"weak32k_T4": r}

        base = None
        if rank == 0 and world == 1 and not args.no_cpu_baseline:
            base = cpu_baseline(wl)
    finally:
        clocks.stop()

    if rank == 0:
        cfg = config_of(wl, api.ARITH_NAMES[arith], world, args.halo, args.cols)
        cfg["arith_note"] = ("F32_FMA = bit-identical to the reference cuda_heat.cu (nvcc default); the MPI reference computes in "
                             "MPI_F64MIX (see also.mpi_f64mix); the two differ by < 1e-5 relative")
        cfg["timing"] = (f"{main['reps']} repetitions of {main['steps']} updates back to back, each between its own CUDA events; per "
                         "repetition max over ranks; ms_per_step = median repetition / steps; warm-up = W updates + one untimed "
                         f"repetition + {LOCKSTEP_UPDATES} lockstep updates after the barrier")
        out = {"metric": METRIC, "value": main["value"], "unit": "GLUP/s", "n_gpus": world, "steps": main["steps"],
               "warmup": max(args.warmup, 3), "ms_per_step": main["ms_per_step"], "higher_is_better": True,
               "scaling": wl["scaling"], "vs_baseline": None,
               "dtype": "f32" if arith != api.MPI_F64MIX else "f32 grid, f64-mixed arithmetic", "data": "synthetic",
               "config": cfg, "reps": main["reps"], "rep_ms": main["rep_ms"], "per_rank_median_ms": main["per_rank_median_ms"],
               "roofline": main["roofline"], "cpu_baseline": base, "e2e": e2e, "gpu_launches": main["gpu_launches"],
               "clocks": main["clocks"], "parity": main["parity"], "k3": main.get("k3"), "also": also}
        print(json.dumps(out), flush=True)
    bad = main["parity"] is not None and not main["parity"]["bit_exact"]
    if e2e and e2e.get("parity") and not e2e["parity"]["bit_exact"]:
        bad = True
    if also:
        bad = bad or any(isinstance(v, dict) and v.get("parity") and not v["parity"]["bit_exact"] for v in also.values())
    return 3 if bad else 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--reps", type=int, default=0, help="timed repetitions (default: >= 7, enough for a 250 ms region)")
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", default="strong")
    ap.add_argument("--arith", choices=["fma", "nofma", "f64mix"], default="fma")
    ap.add_argument("--halo", choices=["p2p", "sendrecv"], default="p2p")
    ap.add_argument("--cols", type=int, default=1, help="N>1: 2-D blocks with this many column blocks (default: row strips)")
    ap.add_argument("--temporal-block", type=int, default=0, help="override the workload's updates per HBM pass (1, 2, 4)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-also", action="store_true", help="skip the other 1-GPU configurations")
    ap.add_argument("--k3-profile", action="store_true", help="N>1: per-rank edge-CTA flag-wait time and NVLink bytes per pass")
    ap.add_argument("--only-3d", type=int, default=0, help="measure only the 3-D leg on an N^3 grid (kernel tuning)")
    ap.add_argument("--no-parity", action="store_true", help="skip the bit-exact window check after the timed region")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    wl = workload(args.workload, max(world, 1) if args.impl == "ours" else max(args.gpus, 1))
    if args.temporal_block:
        wl["temporal_block"] = args.temporal_block
        wl["text"] += f", {args.temporal_block} updates per HBM pass"

    if args.impl == "reference":
        run_reference_arm(args, wl, rank, world)
        return 0

    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        return run_ours(args, wl, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    sys.exit(main())
