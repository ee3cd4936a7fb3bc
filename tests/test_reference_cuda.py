"""Parity against the reference's OWN CUDA program, run live on the GPU box.

oracle/_ref/cuda_heat_* are cuda/cuda_heat.cu compiled unmodified (oracle/build_ref.py) with
nvcc's default -fmad=true and with -fmad=false; the exact float bits handed to its prtdat()
are captured.  F32_FMA must reproduce the first build bit for bit, F32_NOFMA the second --
this is north_star's "bit-exact when built with -fmad=false and the reference's expression
order", and it pins the oracle's two float modes as well."""
import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

REF = ROOT / "oracle" / "_ref"
needs_ref = pytest.mark.skipif(not (REF / "cuda_heat_20x20").exists(), reason="oracle/_ref/cuda_heat_* not built")


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@needs_ref
@pytest.mark.parametrize("nofma", [False, True])
@pytest.mark.parametrize("shape,steps", [((20, 20), 100), ((33, 47), 57), ((64, 64), 500), ((1024, 1024), 300)])
def test_bit_exact_vs_cuda_heat(ph, oracle, shape, steps, nofma):
    r = oracle.run_reference("cuda", shape[0], shape[1], steps, nofma=nofma)
    arith = ph.F32_NOFMA if nofma else ph.F32_FMA
    assert np.array_equal(bits(r["final"]), bits(oracle.solve(oracle.inidat(*shape), steps, arith=arith)[0])), \
        "oracle float mode is not what nvcc builds"
    got, done, _ = ph.solve(ph.inidat(*shape), steps, arith=arith)  # C:204: exactly `steps` updates
    assert done == steps
    assert np.array_equal(bits(got), bits(r["final"]))


@needs_ref
def test_convergence_vs_cuda_heat(ph, oracle):
    # cuda_heat -DCONVERGE, CHECK_INTERVAL 20 (C:16,219).  semi_reduce reads past conv_flags unless
    # NUMBER_OF_BLOCKS % 1024 == 0 (C:36 vs C:184), so the reference's own answer is only trusted
    # where that holds; elsewhere we only require that it stops no earlier than we do.
    seen = []
    for n in (20, 64):
        r = oracle.run_reference("cuda", n, n, 10000, conv=True)
        got, done, conv = ph.solve(ph.inidat(n, n), 10000, arith=ph.F32_FMA, conv_every=20, conv_phase=ph.PHASE_CUDA)
        want, wdone, wconv = oracle.solve(oracle.inidat(n, n), 10000, arith=0, conv_every=20, conv_phase=1)
        assert (done, conv) == (wdone, wconv)
        assert np.array_equal(bits(got), bits(want))
        seen.append((n, r["converged_print"], done, conv))
        if r["converged_print"] is not None and conv:
            assert r["converged_print"] >= done - 1  # C:256 prints the 0-based i
    out = ROOT / "gpurun_out"
    if out.is_dir():
        (out / "ref_cuda_conv.txt").write_text(repr(seen) + "\n")
    assert seen[0][2:] == (1981, True)  # 20x20: "Converged at 1980 steps"
