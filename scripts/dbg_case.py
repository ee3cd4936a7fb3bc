import ctypes, sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import parallel_heat_b200 as ph
from oracle import pyoracle as O
rng = np.random.default_rng(0)
for shape in [(3, 3), (4, 5), (5, 4), (9, 33)]:
    u = (rng.standard_normal(shape) * 100).astype(np.float32)
    for T in (1, 2, 3, 4):
        for steps in (1, T, 2 * T + 1):
            want = O.solve(u, steps, arith=1)[0]
            got = ph.solve(u, steps, arith=1, temporal_block=T)[0]
            bad = np.argwhere(got.view(np.uint32) != want.view(np.uint32))
            print(shape, T, steps, 'bad', len(bad), bad[:6].tolist())
if os.environ.get("HEAT_B200_LIB"):
    n = ctypes.c_uint(0); ph.load().heat_debug_violations(ctypes.byref(n)); print('violations', n.value)
