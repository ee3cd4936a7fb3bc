"""Every load/store of the update kernels stays inside its buffer.

compute-sanitizer is closed on the GPU pool, so the library has a bounds-checked twin
(-DHEAT_BOUNDS_CHECK, parallel_heat_b200/build.py --debug): ld4/ld1/st4 compare each address with
the window of the buffer it belongs to and count violations.  This test drives that twin through
awkward shapes, all tilings, temporal blocking and strips, in a subprocess (a process loads one
build of the library), and requires zero violations and oracle-identical results."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

SCRIPT = r"""
import ctypes, sys
import numpy as np
sys.path.insert(0, %(root)r)
import parallel_heat_b200 as ph
from oracle import pyoracle as O

def strips(u, g, steps, **kw):
    ss = [ph.HeatSolver(u.shape[0], u.shape[1], rank=r, nranks=g, **kw) for r in range(g)]
    for r, s in enumerate(ss):
        s.connect_local(ss[r - 1] if r > 0 else None, ss[r + 1] if r < g - 1 else None)
    for s in ss:
        s.upload(u)
    ph.group_run(ss, steps)
    out = np.zeros_like(u)
    for s in ss:
        s.download(out)
        s.close()
    return out

rng = np.random.default_rng(0)
cases = 0
for shape in [(3, 3), (4, 5), (5, 4), (9, 33), (20, 20), (33, 47), (64, 4), (4, 64), (130, 121), (131, 250), (70, 1000),
              (300, 1030), (257, 4100), (1026, 1000), (40, 128), (41, 129), (2050, 516)]:
    u = (rng.standard_normal(shape) * 100).astype(np.float32)
    for T in (1, 2, 3, 4):
        for steps in (1, T, 2 * T + 1):
            want = O.solve(u, steps, arith=1)[0]
            got = ph.solve(u, steps, arith=1, temporal_block=T)[0]
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (shape, T, steps)
            cases += 1
    for g in (2, 3):
        if shape[0] // g >= 4:
            for T in (1, 4):
                want = O.solve(u, 2 * T + 1, arith=0)[0]
                got = strips(u, g, 2 * T + 1, arith=0, temporal_block=T)
                assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (shape, g, T)
                cases += 1
    got = ph.solve(u, 3, arith=2, conv_every=2)[0]  # residual variant
    assert np.array_equal(got.view(np.uint32), O.solve(u, 3, arith=2, conv_every=2, conv_phase=1)[0].view(np.uint32))
n = ctypes.c_uint(0)
lib = ph.load()
assert lib.heat_debug_violations(ctypes.byref(n)) == 0
print("cases", cases, "violations", n.value)
"""


def test_no_access_leaves_its_buffer(ph):
    from parallel_heat_b200 import build
    dbg = build.build_debug()
    env = dict(os.environ, HEAT_B200_LIB=str(dbg))
    for rpt, rpt_tb in (("0", "0"), ("1", "3"), ("5", "7"), ("64", "300")):
        env.update(HEAT_ROWS_PER_TILE=rpt, HEAT_ROWS_PER_TILE_TB=rpt_tb)
        pr = subprocess.run([sys.executable, "-c", SCRIPT % {"root": str(ROOT)}], env=env, capture_output=True, text=True,
                            timeout=1200)
        assert pr.returncode == 0, pr.stdout[-2000:] + pr.stderr[-4000:]
        assert pr.stdout.strip().endswith("violations 0"), pr.stdout
