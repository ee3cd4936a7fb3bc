#!/usr/bin/env python3
"""Regenerate tests/golden/* from the UNMODIFIED reference (oracle/_ref binaries).

Run in the build container (needs /root/reference for oracle/build_ref.py):
    python oracle/build_ref.py && python tests/golden/make_golden.py
The reference has no tests and no golden vectors of its own;
these fixtures are outputs of its MPI program, compiled unmodified against the
MPI stand-in (one rank; every even-sized case is run again as 4 ranks, 2x2 blocks, and
must give the same bits), with the exact float bits prtdat() was handed.

Files written:
  golden_small.npz   full final fields for the small cases (bit patterns)
  golden.json        sha256 of the raw float32 bytes for every case, md5 of the
                     text dumps (same values SURVEY.md section 8(c) lists), the convergence
                     step counts the reference printed, inidat centre values.
The CUDA reference cannot run here (no GPU); tests/test_reference_cuda.py runs
its prebuilt binaries on the GPU box instead.
"""
import hashlib
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import pyoracle as O  # noqa: E402

HERE = Path(__file__).resolve().parent

# (nx, ny, STEPS, keep full array?)
PLAIN = [(20, 20, 100, True), (20, 20, 0, True), (33, 47, 57, True), (64, 64, 500, True),
         (256, 256, 1000, False), (1024, 1024, 200, False), (1024, 1024, 1000, False)]
# (nx, ny, STEPS, STEP)
CONV = [(20, 20, 10000, 20), (20, 20, 10000, 7), (431, 431, 3000, 20)]
CENTRES = [20, 431, 432, 1024, 8192, 16384, 32768, 65536]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, np.float32).tobytes()).hexdigest()


def main():
    out = {"plain": [], "conv": [], "inidat_centre": {}}
    arrays = {}
    for nx, ny, steps, keep in PLAIN:
        text = (nx, ny, steps) in [(20, 20, 100), (1024, 1024, 200), (1024, 1024, 1000)]
        r = O.run_reference("mpi", nx, ny, steps, text=text)
        e = {"nx": nx, "ny": ny, "steps": steps, "updates": steps + 1,
             "initial_sha256": sha(r["initial"]), "final_sha256": sha(r["final"])}
        if text:
            e["initial_text_md5"] = r["initial_text_md5"]
            e["final_text_md5"] = r["final_text_md5"]
        if keep:
            arrays[f"final_{nx}x{ny}_s{steps}"] = r["final"].view(np.uint32)
        if nx % 2 == 0 and ny % 2 == 0:
            # the same program as 4 ranks (2x2 blocks, HEAT_REF_RANKS): its halo exchange must not change a bit
            r4 = O.run_reference("mpi", nx, ny, steps, ranks=4)
            assert sha(r4["final"]) == e["final_sha256"] and sha(r4["initial"]) == e["initial_sha256"], (nx, ny, steps)
            e["ranks_checked"] = [1, 4]
        out["plain"].append(e)
        print(e)
    for nx, ny, steps, step in CONV:
        r = O.run_reference("mpi", nx, ny, steps, conv=True, step=step)
        e = {"nx": nx, "ny": ny, "steps": steps, "step": step, "printed": r["converged_print"],
             "final_sha256": sha(r["final"])}
        if nx % 2 == 0 and ny % 2 == 0:
            r4 = O.run_reference("mpi", nx, ny, steps, conv=True, step=step, ranks=4)
            assert r4["converged_print"] == e["printed"] and sha(r4["final"]) == e["final_sha256"], (nx, ny, step)
            e["ranks_checked"] = [1, 4]
        out["conv"].append(e)
        print(e)
    for n in CENTRES:
        # M:320 in exact integers, wrapped to int32 the way the reference's `int` product wraps
        ix = iy = n // 2
        v = ix * (n - ix - 1) * iy * (n - iy - 1)
        wrapped = (v + 2**31) % 2**32 - 2**31
        out["inidat_centre"][str(n)] = float(np.float32(wrapped))
        if O.ref_binary("mpi", n, n).exists():  # confirm against the reference's own initial field
            r = O.run_reference("mpi", n, n, 0)
            assert float(r["initial"][ix, iy]) == out["inidat_centre"][str(n)], n
    (HERE / "golden.json").write_text(json.dumps(out, indent=1))
    np.savez_compressed(HERE / "golden_small.npz", **arrays)
    print("wrote", HERE / "golden.json", HERE / "golden_small.npz")


if __name__ == "__main__":
    main()
