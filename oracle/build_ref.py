#!/usr/bin/env python3
"""Compile the UNMODIFIED reference programs from /root/reference into oracle/_ref/.

TEST INFRASTRUCTURE.  Nothing is copied: gcc / nvcc read the sources where they
lie.  The recipe is a plain compiler command per binary (the reference's own
Makefiles need mpicc and /opt/cuda, which this image does not have):

  MPI program (mpi/mpi_heat_improved_persistent_stat.c)
      gcc <opt> -I oracle/refbuild/mpi_stub -include oracle/refbuild/ref_shim.h
          -DNXPROB=.. -DNYPROB=.. -DSTEPS=heat_ref_steps_v -DSTEP=heat_ref_step_v
          [-DCONVERGE] [-fopenmp -DOMPCH -DTHREAD_COUNT=omp_get_max_threads()]
          <ref>.c oracle/refbuild/ref_shim.c -lm
      mpi_stub/mpi.h is an MPI stand-in (1 rank, or $HEAT_REF_RANKS forked ranks) written for this repo (no MPI in the
      image); ref_shim.h makes STEPS/STEP run-time and captures prtdat's values.
      <opt> is "-O0" (as shipped: mpi/Makefile:2,13 has no -O) or
      "-O3 -fwrapv -ffp-contract=off" (the stronger CPU baseline).

  CUDA program (cuda/cuda_heat.cu)
      nvcc -O2 -gencode arch=compute_100a,code=sm_100a -I<ref>/cuda -DUNIX ...
      (cuda/Makefile:5 flags, plus the arch this box needs) [-fmad=false]

`/root/reference` exists only in the build container; the GPU box uses the
binaries this script left in oracle/_ref/ (git-ignored, shipped by gpurun).
"""
from __future__ import annotations

import json
import os
import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF = Path(os.environ.get("HEAT_REFERENCE_DIR", "/root/reference"))
OUT = HERE / "_ref"
SHIM = HERE / "refbuild"

MPI_SRC = REF / "mpi" / "mpi_heat_improved_persistent_stat.c"
CUDA_SRC = REF / "cuda" / "cuda_heat.cu"

OPT = {"O0": ["-O0"], "O3": ["-O3", "-fwrapv", "-ffp-contract=off"]}


def mpi_name(nx, ny, conv=False, omp=False, opt="O3"):
    return f"mpi_heat_{nx}x{ny}" + ("_con" if conv else "") + ("_omp" if omp else "") + f"_{opt}"


def cuda_name(nx, ny, conv=False, nofma=False):
    return f"cuda_heat_{nx}x{ny}" + ("_con" if conv else "") + ("_nofma" if nofma else "")


def build_mpi(nx, ny, conv=False, omp=False, opt="O3"):
    out = OUT / mpi_name(nx, ny, conv, omp, opt)
    cmd = ["gcc", *OPT[opt], "-w", "-I", str(SHIM / "mpi_stub"), "-include", str(SHIM / "ref_shim.h"),
           f"-DNXPROB={nx}", f"-DNYPROB={ny}", "-DSTEPS=heat_ref_steps_v", "-DSTEP=heat_ref_step_v"]
    if conv:
        cmd.append("-DCONVERGE")
    if omp:
        cmd += ["-fopenmp", "-DOMPCH", "-DTHREAD_COUNT=omp_get_max_threads()"]
    # grids above 2 GiB of bss need the medium code model
    if 2 * (nx + 2) * (ny + 2) * 4 > (1 << 31) - (1 << 24):
        cmd.append("-mcmodel=medium")
    cmd += [str(MPI_SRC), str(SHIM / "ref_shim.c"), "-lm", "-o", str(out)]
    subprocess.run(cmd, check=True)
    return out, cmd


def build_cuda(nx, ny, conv=False, nofma=False):
    out = OUT / cuda_name(nx, ny, conv, nofma)
    cmd = ["nvcc", "-O2", "-w", "-gencode", "arch=compute_100a,code=sm_100a", "-I", str(REF / "cuda"),
           "--compiler-options", "-fno-strict-aliasing", "-DUNIX", "-DHEAT_REF_NO_MPI",
           "-include", str(SHIM / "ref_shim.h"),
           f"-DNXPROB={nx}", f"-DNYPROB={ny}", "-DSTEPS=heat_ref_steps_v"]
    if conv:
        cmd.append("-DCONVERGE")
    if nofma:
        cmd.append("-fmad=false")
    cmd += [str(CUDA_SRC), str(SHIM / "ref_shim.c"), "-o", str(out)]
    subprocess.run(cmd, check=True)
    return out, cmd


# what tests/, bench.py and tests/golden/make_golden.py expect to find
DEFAULT_MPI = [
    # (nx, ny, conv, omp, opt)
    (20, 20, False, False, "O0"), (20, 20, False, False, "O3"), (20, 20, True, False, "O3"),
    (33, 47, False, False, "O3"), (64, 64, False, False, "O3"), (64, 64, False, True, "O3"),
    # HEAT_REF_RANKS > 1 (the reference's own block decomposition): sizes that 2x1 .. 4x3 rank grids divide
    (48, 60, False, False, "O3"), (48, 60, True, False, "O3"), (48, 60, False, True, "O3"),
    (256, 256, False, False, "O3"), (431, 431, True, False, "O3"),
    (1024, 1024, False, False, "O0"), (1024, 1024, False, False, "O3"), (1024, 1024, False, True, "O3"),
    (1024, 1024, True, True, "O3"),
    (8192, 8192, False, True, "O3"), (8192, 8192, False, True, "O0"),
    (16384, 16384, True, True, "O3"),
    # bench.py's reference arm for the 65536^2 / 32768^2 workloads: the largest grid whose whole program
    # (inidat, two prtdat walks, 5 field copies) ends inside the time-box
    (16384, 16384, False, True, "O3"),
    # bench.py --impl reference at N GPUs: the weak-scaling grid (8192*N) x 8192
    (16384, 8192, False, True, "O3"), (32768, 8192, False, True, "O3"), (65536, 8192, False, True, "O3"),
]
DEFAULT_CUDA = [
    # (nx, ny, conv, nofma)
    (20, 20, False, False), (20, 20, False, True), (20, 20, True, False),
    (33, 47, False, False), (33, 47, False, True),
    (64, 64, False, False), (64, 64, False, True), (64, 64, True, False),
    (1024, 1024, False, False), (1024, 1024, False, True), (1024, 1024, True, False),
    (8192, 8192, False, False), (16384, 16384, True, False),
]


def main(argv=None):
    if not MPI_SRC.exists():
        print(f"build_ref: {REF} not present (GPU box?) -- keeping prebuilt oracle/_ref/", file=sys.stderr)
        return 0
    OUT.mkdir(exist_ok=True)
    manifest = {}
    for cfg in DEFAULT_MPI:
        out, cmd = build_mpi(*cfg)
        manifest[out.name] = " ".join(cmd)
    if shutil.which("nvcc"):
        for cfg in DEFAULT_CUDA:
            out, cmd = build_cuda(*cfg)
            manifest[out.name] = " ".join(cmd)
    (OUT / "MANIFEST.json").write_text(json.dumps(manifest, indent=1))
    print(f"build_ref: {len(manifest)} reference binaries in {OUT}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
