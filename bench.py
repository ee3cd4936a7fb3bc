#!/usr/bin/env python3
"""bench.py -- the solver-loop benchmark (BASELINE.json metric: GLUP/s and % of HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A bench "step" is ONE Jacobi update of the whole grid (one pass of the hot path).
Workloads (BASELINE.json configs):
    c2      (default)  8192 x 8192 float32, no convergence check; at N GPUs the grid is
                       (8192*N) x 8192 cut into N row strips (weak scaling of config[1])
    c3                 16384 x 16384 with the convergence check every 20 updates (1 GPU)
    strong             65536 x 65536 row strips over N GPUs (config[3])
    weak32k            (32768*N) x 32768, temporal_block 4 (config[4])
    c1                 1024 x 1024 (config[0]'s grid; a parity case, here only for curiosity)
`value` = LUPs of all ranks / device time (CUDA events on the launching stream, max over
ranks), inputs resident in HBM.  `e2e` = the same through heat_solve() with HOST buffers
(H2D + K updates + D2H inside the timed region).  --impl reference times the reference's own
MPI program (oracle/_ref, unmodified source, OpenMP build, all host cores) on the same grid.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "GLUP/s (5-point Jacobi lattice-point updates per second, solver loop)"
BYTES_PER_LUP = 8.0  # float32 single-step: 4 B read + 4 B written per point-update


def workload(name, n):
    if name == "c2":
        return dict(nx=8192 * n, ny=8192, conv_every=0, temporal_block=1, scaling="weak",
                    text=f"{8192 * n}x8192 float32 inidat field, {n} row strip(s) of 8192 rows, no convergence check")
    if name == "c3":
        return dict(nx=16384, ny=16384, conv_every=20, temporal_block=1, scaling="weak",
                    text="16384x16384 float32 inidat field, convergence check every 20 updates")
    if name == "strong":
        return dict(nx=65536, ny=65536, conv_every=0, temporal_block=1, scaling="strong",
                    text=f"65536x65536 float32 inidat field over {n} row strip(s)")
    if name == "weak32k":
        return dict(nx=32768 * n, ny=32768, conv_every=0, temporal_block=4, scaling="weak",
                    text=f"{32768 * n}x32768 float32 inidat field, {n} strip(s), 4 updates per HBM pass")
    if name == "c1":
        return dict(nx=1024, ny=1024, conv_every=0, temporal_block=1, scaling="weak", text="1024x1024 float32 inidat field")
    raise SystemExit(f"unknown workload {name}")


def measured_peak():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            return float(json.loads(f.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(name):
    """Per-launch DRAM bytes of the step kernel from the committed ncu capture, if any."""
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
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    def __init__(self, index=0, period=0.1):
        self.index, self.period = index, period
        self.samples, self.reasons = [], set()
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
            self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                     "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10, "applications_clocks_setting": 0x2}
            for k, bit in names.items():
                if r & bit:
                    self.reasons.add(k)
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

    def stop(self):
        if self._t:
            self._once()
            self._stop.set()
            self._t.join()
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


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


def time_reference(O, wl, cores, loop_budget_secs, max_updates):
    """Times the unmodified reference MPI program (oracle/_ref) on a bounded sample of the workload: every
    layout is probed with 3 updates, the fastest one runs the sample.  Returns (glups, updates, secs, text)."""
    nx, ny, conv = wl["nx"], wl["ny"], wl["conv_every"] > 0
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


def run_reference_arm(args, wl, rank, world):
    if rank != 0:
        return
    cores = host_cores()
    os.environ["OMP_NUM_THREADS"] = str(cores)  # torchrun exports OMP_NUM_THREADS=1 to its workers
    from oracle import pyoracle as O
    nx, ny = wl["nx"], wl["ny"]
    out = {"metric": METRIC, "unit": "GLUP/s", "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None,
           "dtype": "f32 grid, f64-mixed arithmetic (M:168-174)", "data": "synthetic",
           "config": {"workload": wl["text"], "nx": nx, "ny": ny}}
    exe = O.ref_binary("mpi", nx, ny, conv=wl["conv_every"] > 0, omp=True, opt="O3")
    if exe.exists():
        # bounded sample: about 20 s of loop time
        glups, updates, secs, sample = time_reference(O, wl, cores, 20.0, args.steps + args.warmup)
        kind = "reference"
    else:
        ini = O.inidat(nx, ny)
        t0 = time.perf_counter()
        O.solve(ini, 2, arith=O.MPI_F64MIX)
        per_update = max((time.perf_counter() - t0) / 2, 1e-6)
        updates = int(max(3, min(args.steps + args.warmup, 20.0 / per_update)))
        t0 = time.perf_counter()
        O.solve(ini, updates, arith=O.MPI_F64MIX)
        secs, kind = time.perf_counter() - t0, "port"
        cores = O.threads()
        glups = (nx - 2) * (ny - 2) * updates / secs / 1e9
        sample = f"{updates} updates of {nx}x{ny}: oracle/heat_oracle.c (C restatement, MPI_F64MIX), {cores} OpenMP threads"
    out.update(value=glups, ms_per_step=secs / updates * 1e3,
               cpu_baseline={"value": glups, "unit": "GLUP/s", "cores": cores, "kind": kind, "sample": sample},
               e2e={"value": glups, "unit": "GLUP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
    print(json.dumps(out), flush=True)


# ------------------------------------------------------------------------------------ cpu baseline (rank 0, N=1)
def cpu_baseline(wl):
    cores = host_cores()
    os.environ["OMP_NUM_THREADS"] = str(cores)
    from oracle import pyoracle as O
    nx, ny = wl["nx"], wl["ny"]
    lups = (nx - 2) * (ny - 2)
    exe = O.ref_binary("mpi", nx, ny, conv=wl["conv_every"] > 0, omp=True, opt="O3")
    try:
        if exe.exists():
            glups, updates, secs, sample = time_reference(O, wl, cores, 15.0, 2000)
            return {"value": glups, "unit": "GLUP/s", "cores": cores, "kind": "reference", "sample": sample}
        ini = O.inidat(nx, ny)
        t0 = time.perf_counter()
        O.solve(ini, 2, arith=O.MPI_F64MIX)
        updates = int(max(3, min(2000, 15.0 / max((time.perf_counter() - t0) / 2, 1e-6))))
        t0 = time.perf_counter()
        O.solve(ini, updates, arith=O.MPI_F64MIX)
        return {"value": lups * updates / (time.perf_counter() - t0) / 1e9, "unit": "GLUP/s", "cores": O.threads(),
                "kind": "port", "sample": f"{updates} updates of {nx}x{ny}, oracle/heat_oracle.c MPI_F64MIX"}
    except Exception as e:  # a baseline that cannot run must not take the GPU number down with it
        return {"value": None, "unit": "GLUP/s", "cores": cores, "kind": "unavailable", "sample": repr(e)[:200]}


# ------------------------------------------------------------------------------------ our arm
def run_ours(args, wl, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist

    from parallel_heat_b200 import api, build
    if rank == 0 and build.stale():
        build.build()
    if world > 1:
        dist.barrier()
    api.load()
    torch.cuda.set_device(local_rank)
    nx, ny = wl["nx"], wl["ny"]
    arith = {"fma": api.F32_FMA, "nofma": api.F32_NOFMA, "f64mix": api.MPI_F64MIX}[args.arith]
    params = dict(arith=arith, conv_every=wl["conv_every"], temporal_block=wl["temporal_block"])
    stream = torch.cuda.current_stream()

    if world > 1:
        from parallel_heat_b200.strips import StripSolver
        s = StripSolver(nx, ny, device=local_rank, halo=args.halo, cols=args.cols, **params)
        solver = s.solver
        if args.halo == "p2p":
            solver.set_stream(stream.cuda_stream)
        s.init_inidat()
        enqueue, sync = s.enqueue, s.sync
        rows_own = s.rows
        cols_own = s.ncols
    else:
        solver = api.HeatSolver(nx, ny, device=local_rank, **params)
        solver.set_stream(stream.cuda_stream)
        solver.init_inidat()
        enqueue, sync = solver.enqueue, solver.sync
        rows_own = nx
        cols_own = ny

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    K, W = args.steps, max(args.warmup, 3)
    if wl["conv_every"] > 0:
        # the convergence-check loop is host-driven every conv_every updates: time heat_run itself
        def enqueue(n):  # noqa: F811
            solver.run(n)
    enqueue(W)
    sync()
    barrier()
    launches0 = solver.launch_count()
    clocks = ClockSampler(local_rank)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clocks.start()
    e0.record(stream)
    enqueue(K)
    e1.record(stream)
    sync()
    barrier()
    clk = clocks.stop()
    ms = e0.elapsed_time(e1)
    launches = solver.launch_count() - launches0
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    lups_per_update = (nx - 2) * (ny - 2)
    value = lups_per_update * K / (ms * 1e-3) / 1e9

    # roofline of the dominant kernel (the update kernel): this rank's algorithmic bytes per launch
    peak, peak_src = measured_peak()
    step_launches = K // wl["temporal_block"] if wl["temporal_block"] > 1 else K
    own_lups = (min(rows_own, nx - 2) if world > 1 else nx - 2) * min(cols_own, ny - 2) * wl["temporal_block"]
    kernel_ms = ms / max(step_launches, 1)
    achieved = BYTES_PER_LUP * own_lups / (kernel_ms * 1e-3) / 1e9
    kname = "heat_step_tb" if wl["temporal_block"] > 1 else "heat_step_vec"
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": ncu_traffic(kname), "kernel": kname, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": BYTES_PER_LUP * own_lups,
                "note": "8 B per point-update x interior points of this GPU's strip / mean launch time over the timed region"
                        + (" (effective: temporal blocking re-uses tiles on chip)" if wl["temporal_block"] > 1 else "")}

    # end to end: heat_solve() on HOST memory (pinned), H2D + K updates + D2H inside the timed region
    e2e = None
    if world == 1 and not args.no_e2e:
        host = torch.empty((nx, ny), dtype=torch.float32, pin_memory=True)
        hn = host.numpy()
        hn[...] = api.inidat(nx, ny)
        p = api.default_params(nx=nx, ny=ny, steps=K, device=local_rank, **params)
        import ctypes as C
        done, conv = C.c_int(0), C.c_int(0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rc = api.load().heat_solve(C.byref(p), hn.ctypes.data_as(C.POINTER(C.c_float)), C.byref(done), C.byref(conv))
        dt = time.perf_counter() - t0
        if rc != 0:
            raise RuntimeError(api.load().heat_last_error().decode())
        e2e = {"value": lups_per_update * done.value / dt / 1e9, "unit": "GLUP/s",
               "h2d_bytes_per_step": nx * ny * 4 / max(done.value, 1), "d2h_bytes_per_step": nx * ny * 4 / max(done.value, 1),
               "call": f"heat_solve(steps={K}) on pinned host memory: cudaMalloc + H2D of the {nx * ny * 4 >> 20} MiB field "
                       f"+ {done.value} updates + D2H, wall clock", "seconds": dt}
        launches_e2e = done.value
    elif world > 1 and not args.no_e2e and args.cols == 1:
        # strips: every rank brings its own window of the field in pinned host memory, uploads it,
        # runs K updates and reads its rows back; wall clock between barriers (max over ranks)
        _, _, nloc, lo = solver.device_view(0)
        host = torch.empty((nloc, ny), dtype=torch.float32, pin_memory=True)
        ix = np.arange(lo, lo + nloc, dtype=np.uint32)[:, None]
        iy = np.arange(ny, dtype=np.uint32)[None, :]
        prod = ix * (np.uint32(nx) - ix - np.uint32(1)) * iy * (np.uint32(ny) - iy - np.uint32(1))  # M:320, wraps like int
        host.numpy()[...] = prod.view(np.int32).astype(np.float32)
        back = torch.empty((rows_own, ny), dtype=torch.float32, pin_memory=True)
        barrier()
        t0 = time.perf_counter()
        solver.upload_local(host.numpy())
        dist.barrier()  # nobody stores into a neighbour whose upload has not landed (flags order it; this is belt and braces)
        enqueue(K)
        sync()
        solver.download_local(back.numpy())
        barrier()
        dt = time.perf_counter() - t0
        tot_in = torch.tensor([float(nloc * ny * 4)], device="cuda")
        tot_out = torch.tensor([float(rows_own * ny * 4)], device="cuda")
        dist.all_reduce(tot_in)
        dist.all_reduce(tot_out)
        e2e = {"value": lups_per_update * K / dt / 1e9, "unit": "GLUP/s",
               "h2d_bytes_per_step": float(tot_in.item()) / K, "d2h_bytes_per_step": float(tot_out.item()) / K,
               "call": f"per rank: heat_upload_local (pinned host window) + {K} updates + heat_download_local, "
                       "wall clock between barriers", "seconds": dt}

    # the same job with 4 updates fused per pass over HBM (bit-identical result): reported next to the
    # single-step number, never instead of it
    also = None
    if world == 1 and wl["temporal_block"] == 1 and wl["conv_every"] == 0 and not args.no_also:
        with api.HeatSolver(nx, ny, device=local_rank, arith=arith, temporal_block=4) as s4:
            s4.set_stream(stream.cuda_stream)
            s4.init_inidat()
            s4.enqueue(4 * max(W // 4, 1))
            s4.sync()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            k4 = max(K // 4, 1) * 4
            a0.record(stream)
            s4.enqueue(k4)
            a1.record(stream)
            s4.sync()
            ms4 = a0.elapsed_time(a1)
        v4 = lups_per_update * k4 / (ms4 * 1e-3) / 1e9
        also = {"temporal_block_4": {"value": v4, "unit": "GLUP/s", "ms_per_step": ms4 / k4, "steps": k4,
                                     "effective_gbs": v4 * BYTES_PER_LUP, "effective_frac": v4 * BYTES_PER_LUP / peak,
                                     "traffic": ncu_traffic("heat_step_tb"),
                                     "note": "heat_step_tb<T=4>: 4 updates per pass, same bits as 4 single updates; "
                                             "8 B/LUP counted as for single-step, so this is EFFECTIVE bandwidth"}}

        # and in the arithmetic of the MPI reference (bit-identical to the program --impl reference times)
        if arith != api.MPI_F64MIX:
            with api.HeatSolver(nx, ny, device=local_rank, arith=api.MPI_F64MIX) as s6:
                s6.set_stream(stream.cuda_stream)
                s6.init_inidat()
                s6.enqueue(W)
                s6.sync()
                k6 = max(K // 4, 16)
                a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a0.record(stream)
                s6.enqueue(k6)
                a1.record(stream)
                s6.sync()
                ms6 = a0.elapsed_time(a1)
            v6 = lups_per_update * k6 / (ms6 * 1e-3) / 1e9
            also["mpi_f64mix"] = {"value": v6, "unit": "GLUP/s", "ms_per_step": ms6 / k6, "steps": k6,
                                  "frac": v6 * BYTES_PER_LUP / peak,
                                  "note": "same kernel in the MPI program's mixed float/double arithmetic (M:168-174), "
                                          "bit-identical to the reference arm's output; FP64-pipe-bound"}

    base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        base = cpu_baseline(wl)

    if rank == 0:
        out = {"metric": METRIC, "value": value, "unit": "GLUP/s", "n_gpus": world, "steps": K, "warmup": W,
               "ms_per_step": ms / K, "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None,
               "dtype": "f32" if arith != api.MPI_F64MIX else "f32 grid, f64-mixed arithmetic",
               "data": "synthetic",
               "config": {"workload": wl["text"], "nx": nx, "ny": ny, "arith": api.ARITH_NAMES[arith],
                          "arith_note": "F32_FMA = bit-identical to the reference cuda_heat.cu (nvcc default); the MPI reference "
                                        "computes in MPI_F64MIX (see also.mpi_f64mix); the two differ by < 1e-5 relative",
                          "conv_every": wl["conv_every"], "temporal_block": wl["temporal_block"],
                          "halo": (args.halo if world > 1 else "none"),
                          "decomposition": (f"{world // args.cols}x{args.cols} blocks" if world > 1 else "single"),
                          "l2": f"working set {2 * nx * ny * 4 >> 20} MiB vs 126 MB L2: inputs larger than L2, no flush "
                                "(consecutive updates ARE the workload)"},
               "roofline": roofline, "cpu_baseline": base, "e2e": e2e, "gpu_launches": int(launches), "clocks": clk,
               "also": also}
        print(json.dumps(out), flush=True)
    if world > 1:
        s.close()
    else:
        solver.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--arith", choices=["fma", "nofma", "f64mix"], default="fma")
    ap.add_argument("--halo", choices=["p2p", "sendrecv"], default="p2p")
    ap.add_argument("--cols", type=int, default=1, help="N>1: 2-D blocks with this many column blocks (default: row strips)")
    ap.add_argument("--temporal-block", type=int, default=0, help="override the workload's updates per HBM pass (1, 2, 4)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-also", action="store_true", help="skip the extra temporal-block-4 measurement")
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
        run_ours(args, wl, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
