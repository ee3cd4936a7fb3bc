"""Build libheat_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIB = PKG / "libheat_b200.so"
INCLUDE = PKG.parent / "include"

NVCC_FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC", "-shared"]


def sources():
    return sorted(CSRC.glob("*.cu"))


def stale():
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    deps = list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h"))
    return any(d.stat().st_mtime > t for d in deps)


DEBUG_LIB = PKG / "libheat_b200_dbg.so"


def build_debug():
    """Bounds-checked twin of the library (-DHEAT_BOUNDS_CHECK), for tests/test_bounds_debug.py."""
    deps = list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h"))
    if DEBUG_LIB.exists() and all(d.stat().st_mtime <= DEBUG_LIB.stat().st_mtime for d in deps):
        return DEBUG_LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc, *NVCC_FLAGS, "-DHEAT_BOUNDS_CHECK", "-I", str(INCLUDE), "-o", str(DEBUG_LIB), *map(str, sources()),
           "-ccbin", "/usr/bin/g++"]
    subprocess.run(cmd, check=True)
    return DEBUG_LIB


def build(force=False, verbose=False):
    """Compile every .cu under csrc/ into libheat_b200.so.  Returns the library path."""
    if not force and not stale():
        return LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: libheat_b200.so cannot be built (there is no CPU fallback)")
    cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE)]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += ["-o", str(LIB), *map(str, sources())]
    env = dict(os.environ)
    # the image exports CC/CXX=/opt/gcc/bin/*, a trimmed compiler; nvcc wants the system g++
    subprocess.run(cmd + ["-ccbin", "/usr/bin/g++"], check=True, env=env)
    return LIB


def build_variant(name, defines):
    """libheat_b200_<name>.so with extra -D switches: A/B runs of one kernel decision on the GPU box
    (HEAT_B200_LIB=<path> makes api.py, the tests and bench.py load it instead of the product library)."""
    out = PKG / f"libheat_b200_{name}.so"
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc, *NVCC_FLAGS, *defines, "-I", str(INCLUDE), "-o", str(out), *map(str, sources()), "-ccbin", "/usr/bin/g++"]
    subprocess.run(cmd, check=True)
    return out


if __name__ == "__main__":
    if "--variant" in sys.argv:
        k = sys.argv.index("--variant")
        print(build_variant(sys.argv[k + 1], [a for a in sys.argv[k + 2:] if a.startswith("-D")]))
    elif "--debug" in sys.argv:
        print(build_debug())
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
