"""The CPU oracle against the reference's outputs (CPU only; no CUDA involved).

Pinning: oracle(MPI_F64MIX, steps+1 updates) must reproduce, bit for bit, what the
UNMODIFIED reference MPI program produced (tests/golden/*, made by make_golden.py), and --
where the prebuilt reference binaries are present -- what they produce right now.
"""
import hashlib

import numpy as np
import pytest

from conftest import have_reference_binaries


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, np.float32).tobytes()).hexdigest()


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_inidat_matches_reference_initial_field(oracle, golden):
    for e in golden["plain"]:
        assert sha(oracle.inidat(e["nx"], e["ny"])) == e["initial_sha256"], e


def test_inidat_int32_wrap_centres(oracle, golden):
    # SURVEY.md section 8(c) lists the same values: 432 is the first size whose product wraps
    for n, want in golden["inidat_centre"].items():
        n = int(n)
        if n > 16384:
            # one row is enough: evaluate M:320 for the centre row only
            ix = iy = n // 2
            v = (ix * (n - ix - 1) * iy * (n - iy - 1) + 2**31) % 2**32 - 2**31
            assert float(np.float32(v)) == want
        else:
            assert float(oracle.inidat(n, n)[n // 2, n // 2]) == want, n
    assert golden["inidat_centre"]["431"] > 0 > golden["inidat_centre"]["432"]


def test_inidat_ring_is_zero(oracle):
    u = oracle.inidat(37, 53)
    assert not u[0].any() and not u[-1].any() and not u[:, 0].any() and not u[:, -1].any()


def test_solve_f64mix_bit_exact_vs_golden(oracle, golden):
    for e in golden["plain"]:
        if e["nx"] * e["ny"] * e["updates"] > 3e8:
            continue  # the 1024^2 x 1001 case is covered by test_solve_1024_full below
        ini = oracle.inidat(e["nx"], e["ny"])
        fin, done, conv = oracle.solve(ini, e["steps"], arith=oracle.MPI_F64MIX, step_convention=1)
        assert done == e["updates"] and not conv
        assert sha(fin) == e["final_sha256"], e
        key = f"final_{e['nx']}x{e['ny']}_s{e['steps']}"
        if key in golden["arrays"]:
            assert np.array_equal(bits(fin), golden["arrays"][key])


def test_solve_1024_full(oracle, golden):
    # BASELINE config 1: 1024 x 1024, STEPS=1000 (1001 updates)
    e = [x for x in golden["plain"] if x["nx"] == 1024 and x["steps"] == 1000][0]
    fin, done, _ = oracle.solve(oracle.inidat(1024, 1024), 1000, arith=oracle.MPI_F64MIX, step_convention=1)
    assert done == 1001 and sha(fin) == e["final_sha256"]


def test_prtdat_text_matches_reference_md5(oracle, golden, tmp_path):
    e = [x for x in golden["plain"] if x["nx"] == 20 and x["steps"] == 100][0]
    ini = oracle.inidat(20, 20)
    oracle.prtdat(ini, tmp_path / "i.dat")
    assert hashlib.md5((tmp_path / "i.dat").read_bytes()).hexdigest() == e["initial_text_md5"]
    fin, _, _ = oracle.solve(ini, 100, arith=oracle.MPI_F64MIX, step_convention=1)
    oracle.prtdat(fin, tmp_path / "f.dat")
    assert hashlib.md5((tmp_path / "f.dat").read_bytes()).hexdigest() == e["final_text_md5"]
    # the survey's known-answer line
    line2 = (tmp_path / "f.dat").read_text().splitlines()[1]
    assert line2.split()[:6] == ["0.0", "138.5", "272.9", "399.2", "513.9", "613.9"]


def test_convergence_matches_reference_print(oracle, golden):
    for e in golden["conv"]:
        fin, done, conv = oracle.solve(oracle.inidat(e["nx"], e["ny"]), e["steps"], arith=oracle.MPI_F64MIX,
                                       step_convention=1, conv_every=e["step"], conv_phase=0)
        if e["printed"] is None:
            assert not conv and done == e["steps"] + 1
        else:
            assert conv and done - 1 == e["printed"], e  # M:302 prints the 0-based `it`
        assert sha(fin) == e["final_sha256"]


def test_cuda_phase_convention(oracle):
    # C:219 checks when i % 20 == 0, after the update: first possible stop is update 1
    ini = oracle.inidat(20, 20)
    for arith in (oracle.F32_FMA, oracle.F32_NOFMA):
        _, done, conv = oracle.solve(ini, 10000, arith=arith, conv_every=20, conv_phase=1, step_convention=0)
        assert conv and done == 1981  # "Converged at 1980 steps" (derived; checked on the GPU box against cuda_heat)
    _, done, conv = oracle.solve(np.zeros((9, 9), np.float32), 50, conv_every=20, conv_phase=1)
    assert conv and done == 1


def test_numpy_restatement_agrees(oracle):
    rng = np.random.default_rng(7)
    for shape in [(3, 3), (4, 7), (65, 130), (200, 33)]:
        u = (rng.standard_normal(shape) * 1e3).astype(np.float32)
        for arith in (oracle.F32_NOFMA, oracle.MPI_F64MIX):
            v = u
            for _ in range(4):
                a, b = oracle.step(v, arith), oracle.step_numpy(v, arith)
                assert np.array_equal(bits(a), bits(b))
                v = a


def test_modes_differ_but_stay_close(oracle):
    ini = oracle.inidat(256, 256)
    outs = {a: oracle.solve(ini, 1000, arith=a)[0] for a in (0, 1, 2)}
    scale = np.abs(outs[2]).max()
    assert not np.array_equal(bits(outs[1]), bits(outs[2]))
    for a in (0, 1):
        assert np.abs(outs[a] - outs[2]).max() / scale < 1e-5  # north_star's tolerance between modes


def test_ring_untouched_and_degenerate_shapes(oracle):
    rng = np.random.default_rng(3)
    u = rng.standard_normal((17, 23)).astype(np.float32)
    for arith in (0, 1, 2):
        v, done, _ = oracle.solve(u, 9, arith=arith)
        assert done == 9
        for sl in (np.s_[0], np.s_[-1], np.s_[:, 0], np.s_[:, -1]):
            assert np.array_equal(bits(v[sl]), bits(u[sl]))
    for shape in [(1, 1), (1, 9), (2, 2), (2, 50), (50, 2)]:  # no interior: nothing changes
        w = rng.standard_normal(shape).astype(np.float32)
        v, done, _ = oracle.solve(w, 5)
        assert done == 5 and np.array_equal(bits(v), bits(w))
    v, done, _ = oracle.solve(u, 0)  # zero updates in the CUDA convention
    assert done == 0 and np.array_equal(bits(v), bits(u))
    v, done, _ = oracle.solve(u, 0, step_convention=1)  # M:159 still does one
    assert done == 1


def test_max_principle(oracle):
    ini = oracle.inidat(64, 64)
    v, _, _ = oracle.solve(ini, 300)
    assert v.max() <= ini.max() and v.min() >= ini.min()


@pytest.mark.skipif(not have_reference_binaries(), reason="oracle/_ref not built (needs /root/reference)")
def test_live_reference_binary_agrees(oracle):
    r = oracle.run_reference("mpi", 33, 47, 31)
    fin, done, _ = oracle.solve(oracle.inidat(33, 47), 31, arith=oracle.MPI_F64MIX, step_convention=1)
    assert done == 32 and np.array_equal(bits(fin), bits(r["final"]))
    # the OpenMP build of the reference (-DOMPCH) gives the same bits
    r2 = oracle.run_reference("mpi", 64, 64, 50, omp=True)
    fin, _, _ = oracle.solve(oracle.inidat(64, 64), 50, arith=oracle.MPI_F64MIX, step_convention=1)
    assert np.array_equal(bits(fin), bits(r2["final"]))
    r3 = oracle.run_reference("mpi", 20, 20, 10000, conv=True, step=20)
    assert r3["converged_print"] == 1979


@pytest.mark.skipif(not have_reference_binaries(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("ranks", [2, 4, 6, 8, 12])
def test_reference_in_its_own_block_decomposition(oracle, ranks):
    """The unmodified MPI program run as `ranks` processes (HEAT_REF_RANKS: the stand-in forks them and carries
    the messages through shared memory): MPI_Dims_create's rank grid (2x1, 2x2, 3x2, 4x2, 4x3), row_type /
    col_type halos, persistent requests (M:51-84,130-155,160-177,263).  Same bits as the undecomposed oracle."""
    r = oracle.run_reference("mpi", 48, 60, 57, ranks=ranks)
    fin, done, _ = oracle.solve(oracle.inidat(48, 60), 57, arith=oracle.MPI_F64MIX, step_convention=1)
    assert done == 58 and np.array_equal(bits(fin), bits(r["final"]))
    assert np.array_equal(bits(oracle.inidat(48, 60)), bits(r["initial"]))
    # the OpenMP team inside every rank (M:163, -DOMPCH) changes nothing either
    r = oracle.run_reference("mpi", 48, 60, 57, ranks=ranks, omp=True, threads=2)
    assert np.array_equal(bits(fin), bits(r["final"]))


@pytest.mark.skipif(not have_reference_binaries(), reason="oracle/_ref not built (needs /root/reference)")
def test_reference_ranks_convergence_and_baseline_config(oracle, golden):
    # MPI_Allreduce(LAND) over 4 and 6 ranks (M:255): the same stop as one rank and as the oracle
    assert oracle.run_reference("mpi", 20, 20, 10000, conv=True, ranks=4)["converged_print"] == 1979
    r = oracle.run_reference("mpi", 48, 60, 100000, conv=True, ranks=6)
    fin, done, conv = oracle.solve(oracle.inidat(48, 60), 100000, arith=oracle.MPI_F64MIX, step_convention=1,
                                   conv_every=20, conv_phase=0)
    assert conv and r["converged_print"] == done - 1 and np.array_equal(bits(fin), bits(r["final"]))
    # BASELINE config[0]: 1024^2, 1000 steps, 4 ranks -- the text dump SURVEY section 8(c) pins
    e = [x for x in golden["plain"] if x["nx"] == 1024 and x["steps"] == 1000][0]
    r = oracle.run_reference("mpi", 1024, 1024, 1000, ranks=4, text=True, capture=False)
    assert r["final_text_md5"] == e["final_text_md5"] == "e2c85256585896d5183a3006279cea23"


def test_3d_oracle_against_a_numpy_restatement(oracle):
    # the 3-D mode has no reference to be pinned to (the reference is 2-D); at least its two unfused arithmetics
    # must agree bit for bit with an independent numpy statement of the same expression, and reduce to the 2-D
    # update when nothing varies along z
    rng = np.random.default_rng(3)
    u = (rng.standard_normal((9, 11, 12)) * 1e3).astype(np.float32)
    cx, cy, cz = np.float32(0.1), np.float32(0.07), np.float32(0.05)
    c = u[1:-1, 1:-1, 1:-1]
    xs = u[2:, 1:-1, 1:-1] + u[:-2, 1:-1, 1:-1]
    ys = u[1:-1, 2:, 1:-1] + u[1:-1, :-2, 1:-1]
    zs = u[1:-1, 1:-1, 2:] + u[1:-1, 1:-1, :-2]
    two = np.float32(2.0)
    want1 = u.copy()
    want1[1:-1, 1:-1, 1:-1] = ((c + cx * (xs - two * c)) + cy * (ys - two * c)) + cz * (zs - two * c)
    got1, done, _ = oracle.solve3d(u, 1, arith=oracle.F32_NOFMA, cx=cx, cy=cy, cz=cz)
    assert done == 1 and np.array_equal(got1.view(np.uint32), want1.view(np.uint32))
    cd = c.astype(np.float64)
    r = cd + np.float64(cx) * (xs.astype(np.float64) - 2.0 * cd)
    r = r + np.float64(cy) * (ys.astype(np.float64) - 2.0 * cd)
    r = r + np.float64(cz) * (zs.astype(np.float64) - 2.0 * cd)
    want2 = u.copy()
    want2[1:-1, 1:-1, 1:-1] = r.astype(np.float32)
    got2, _, _ = oracle.solve3d(u, 1, arith=oracle.MPI_F64MIX, cx=cx, cy=cy, cz=cz)
    assert np.array_equal(got2.view(np.uint32), want2.view(np.uint32))
    # the ring (all six faces) is never written
    for g in (got1, got2):
        for ax in range(3):
            for face in (0, -1):
                assert np.array_equal(np.take(g, face, axis=ax).view(np.uint32), np.take(u, face, axis=ax).view(np.uint32))
    # constant along z with cz*0: every z-slice evolves like the 2-D field
    v = (rng.standard_normal((10, 13)) * 1e2).astype(np.float32)
    v3 = np.repeat(v[:, :, None], 8, axis=2)
    a3, _, _ = oracle.solve3d(v3, 5, arith=oracle.F32_NOFMA, cz=0.0)
    a2, _, _ = oracle.solve(v, 5, arith=oracle.F32_NOFMA)
    assert np.allclose(a3[:, :, 3], a2, rtol=1e-6, atol=1e-4)  # (+ cz*0 adds a rounding-free zero: equal up to -0.0)
    assert np.array_equal(oracle.inidat3d(5, 6, 8)[2, 3], (2 * 2 * 3 * 2 * np.arange(8) * (7 - np.arange(8))).astype(np.float32))
