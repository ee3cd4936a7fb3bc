"""GPU parity tests: libheat_b200.so (through its C ABI) against the CPU oracle, the golden
vectors of the reference MPI program, and size-independent properties at full size.
Everything here is bit-exact: each arithmetic mode reproduces one build of the reference."""
import hashlib
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MODES = [0, 1, 2]  # F32_FMA, F32_NOFMA, MPI_F64MIX (same numbering in the oracle)


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, np.float32).tobytes()).hexdigest()


def assert_same_bits(a, b, what=""):
    a, b = bits(a), bits(b)
    if not np.array_equal(a, b):
        bad = np.argwhere(a != b)
        fa, fb = a.view(np.float32), b.view(np.float32)
        raise AssertionError(f"{what}: {len(bad)} of {a.size} values differ, first at {tuple(bad[0])}: "
                             f"{fa[tuple(bad[0])]!r} vs {fb[tuple(bad[0])]!r}")


def field(shape, seed=0, scale=1e3):
    return (np.random.default_rng(seed).standard_normal(shape) * scale).astype(np.float32)


# ---------------------------------------------------------------- K0 inidat
@pytest.mark.parametrize("shape", [(20, 20), (431, 431), (432, 432), (33, 47), (1024, 1024), (1000, 1026), (3, 5)])
def test_device_inidat_is_the_host_inidat(ph, oracle, shape):
    with ph.HeatSolver(*shape) as s:
        s.init_inidat()
        assert_same_bits(s.download(), oracle.inidat(*shape), f"inidat {shape}")


def test_device_inidat_8192_wraps_like_the_reference(ph, golden):
    with ph.HeatSolver(8192, 8192) as s:
        s.init_inidat()
        u = s.download()
    assert float(u[4096, 4096]) == golden["inidat_centre"]["8192"]
    rows = [0, 1, 431, 432, 4095, 8190, 8191]
    host = ph.inidat(8192, 8192)
    assert_same_bits(u[rows], host[rows], "inidat 8192 sampled rows")


# ---------------------------------------------------------------- K1: updates vs oracle
SHAPES = [(3, 3), (3, 9), (9, 3), (4, 4), (5, 5), (20, 20), (33, 47), (47, 33), (64, 4), (4, 64), (129, 257),
          (130, 1000), (1000, 1026), (1026, 1000), (257, 4100)]


@pytest.mark.parametrize("arith", MODES)
@pytest.mark.parametrize("shape", SHAPES)
def test_updates_bit_exact_vs_oracle(ph, oracle, shape, arith):
    u = field(shape, seed=shape[0] * 7 + shape[1])
    for steps in (1, 2, 7):
        want, _, _ = oracle.solve(u, steps, arith=arith)
        got, done, conv = ph.solve(u, steps, arith=arith)
        assert done == steps and not conv
        assert_same_bits(got, want, f"{shape} arith={arith} steps={steps}")


@pytest.mark.parametrize("arith", MODES)
def test_ring_and_both_buffers(ph, oracle, arith):
    # a NON-zero ring: the solver may not assume inidat's zero boundary
    u = field((37, 53), seed=5)
    with ph.HeatSolver(37, 53, arith=arith) as s:
        s.upload(u)
        s.run(3)
        newest = s.download()
        for sl in (np.s_[0], np.s_[-1], np.s_[:, 0], np.s_[:, -1]):
            assert_same_bits(newest[sl], u[sl], "ring")
        want3, _, _ = oracle.solve(u, 3, arith=arith)
        assert_same_bits(newest, want3)
        s.run(1)  # the other buffer now holds update 4; its ring must be intact too
        want4, _, _ = oracle.solve(u, 4, arith=arith)
        assert_same_bits(s.download(), want4)


@pytest.mark.parametrize("shape", [(1, 1), (1, 9), (2, 2), (2, 50), (50, 2), (2, 3)])
def test_no_interior_means_no_change(ph, shape):
    u = field(shape, seed=1)
    got, done, _ = ph.solve(u, 5)
    assert done == 5
    assert_same_bits(got, u)


def test_step_conventions(ph, oracle):
    u = ph.inidat(20, 20)
    got, done, _ = ph.solve(u, 0)
    assert done == 0 and np.array_equal(got, u)
    got, done, _ = ph.solve(u, 0, step_convention=ph.STEPS_MPI)  # M:159 `it <= STEPS`
    assert done == 1
    assert_same_bits(got, oracle.solve(u, 1)[0])


@pytest.mark.parametrize("rpt", [1, 3, 7, 64, 1000])
def test_tiling_does_not_change_bits(ph, oracle, rpt, monkeypatch):
    monkeypatch.setenv("HEAT_ROWS_PER_TILE", str(rpt))
    u = field((131, 520), seed=9)
    for arith in MODES:
        got, _, _ = ph.solve(u, 5, arith=arith)
        assert_same_bits(got, oracle.solve(u, 5, arith=arith)[0], f"rows_per_tile={rpt}")


def test_serpentine_sweep_does_not_change_bits(ph, monkeypatch):
    u = ph.inidat(300, 300)
    a, _, _ = ph.solve(u, 9)
    monkeypatch.setenv("HEAT_SERPENTINE", "0")
    b, _, _ = ph.solve(u, 9)
    assert_same_bits(a, b)


# ---------------------------------------------------------------- golden vectors of the reference
def test_golden_reference_mpi_outputs(ph, golden):
    for e in golden["plain"]:
        got, done, conv = ph.solve(ph.inidat(e["nx"], e["ny"]), e["steps"], arith=ph.MPI_F64MIX,
                                   step_convention=ph.STEPS_MPI)
        assert done == e["updates"] and not conv
        assert sha(got) == e["final_sha256"], e
        key = f"final_{e['nx']}x{e['ny']}_s{e['steps']}"
        if key in golden["arrays"]:
            assert np.array_equal(bits(got), golden["arrays"][key])


def test_golden_text_dump_md5(ph, golden, tmp_path):
    # BASELINE config 1 end to end: 1024 x 1024, STEPS=1000, final_im.dat md5 of the reference
    e = [x for x in golden["plain"] if x["nx"] == 1024 and x["steps"] == 1000][0]
    got, _, _ = ph.solve(ph.inidat(1024, 1024), 1000, arith=ph.MPI_F64MIX, step_convention=ph.STEPS_MPI)
    ph.prtdat(got, tmp_path / "final_im.dat")
    assert hashlib.md5((tmp_path / "final_im.dat").read_bytes()).hexdigest() == e["final_text_md5"]


def test_float_modes_within_tolerance_of_the_mpi_reference(ph, golden):
    # north_star: <= 1e-5 relative (max-abs / max|u|) between the float builds and the MPI build
    ini = ph.inidat(1024, 1024)
    ref, _, _ = ph.solve(ini, 1000, arith=ph.MPI_F64MIX, step_convention=ph.STEPS_MPI)
    scale = np.abs(ref).max()
    for arith in (ph.F32_FMA, ph.F32_NOFMA):
        got, _, _ = ph.solve(ini, 1000, arith=arith, step_convention=ph.STEPS_MPI)
        assert np.abs(got - ref).max() / scale <= 1e-5


# ---------------------------------------------------------------- K2: convergence
def test_convergence_matches_the_reference_print(ph, golden):
    for e in golden["conv"]:
        got, done, conv = ph.solve(ph.inidat(e["nx"], e["ny"]), e["steps"], arith=ph.MPI_F64MIX,
                                   step_convention=ph.STEPS_MPI, conv_every=e["step"], conv_phase=ph.PHASE_MPI)
        if e["printed"] is None:
            assert not conv and done == e["steps"] + 1
        else:
            assert conv and done - 1 == e["printed"], e
        assert sha(got) == e["final_sha256"]


@pytest.mark.parametrize("arith", MODES)
@pytest.mark.parametrize("phase", [0, 1])
def test_convergence_vs_oracle(ph, oracle, arith, phase):
    for shape, K in [((20, 20), 20), ((20, 20), 7), ((33, 47), 5), ((64, 200), 20)]:
        u = ph.inidat(*shape)
        want, wdone, wconv = oracle.solve(u, 10000, arith=arith, conv_every=K, conv_phase=phase)
        got, done, conv = ph.solve(u, 10000, arith=arith, conv_every=K, conv_phase=phase)
        assert (done, conv) == (wdone, wconv)
        assert conv or shape == (64, 200)  # the small grids settle well inside 10000 updates
        assert_same_bits(got, want)


def test_fused_residual_equals_separate_pass(ph, oracle):
    u = field((200, 328), seed=11)
    with ph.HeatSolver(200, 328, conv_every=1, conv_eps=0.0) as s:  # eps 0: checks every update, never stops
        s.upload(u)
        done, conv = s.run(4)
        assert done == 4 and not conv
        sep = s.residual()
    a, _, _ = oracle.solve(u, 3)
    b, _, _ = oracle.solve(u, 4)
    assert sep == oracle.max_abs_diff(a, b)


def test_run_can_be_resumed(ph, oracle):
    u = ph.inidat(64, 64)
    with ph.HeatSolver(64, 64, conv_every=20, conv_phase=ph.PHASE_MPI, arith=ph.MPI_F64MIX) as s:
        s.upload(u)
        total = 0
        for chunk in (7, 13, 50, 1000, 100000):
            d, conv = s.run(chunk)
            total += d
            if conv:
                break
        got = s.download()
    want, wdone, wconv = oracle.solve(u, 200000, arith=2, conv_every=20, conv_phase=0)
    assert conv and wconv and total == wdone
    assert_same_bits(got, want)


# ---------------------------------------------------------------- stateless entry points
@pytest.mark.parametrize("arith", MODES)
def test_stateless_step_device(ph, oracle, arith):
    import torch
    for shape in [(20, 20), (33, 47), (257, 130)]:
        u = field(shape, seed=2)
        d_in = torch.from_numpy(u).cuda()
        d_out = torch.full(shape, -7.0, device="cuda")  # ring of d_out must stay -7
        resid = torch.zeros(1, dtype=torch.int32, device="cuda")
        ph.step_residual_device(d_in.data_ptr(), d_out.data_ptr(), shape[0], shape[1], resid.data_ptr(), arith=arith,
                                stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        out = d_out.cpu().numpy()
        want = oracle.step(u, arith, out=np.full(shape, -7.0, np.float32))
        assert_same_bits(out, want, f"stateless {shape}")
        r = resid.cpu().numpy().view(np.float32)[0]
        assert r == np.abs(u[1:-1, 1:-1] - want[1:-1, 1:-1]).max()
        d_out2 = torch.full(shape, -7.0, device="cuda")
        ph.step_device(d_in.data_ptr(), d_out2.data_ptr(), shape[0], shape[1], arith=arith)
        torch.cuda.synchronize()
        assert_same_bits(d_out2.cpu().numpy(), want)


# ---------------------------------------------------------------- strips (one GPU, several strips)
def run_strips(ph, u, g, steps, **params):
    nx, ny = u.shape
    ss = [ph.HeatSolver(nx, ny, rank=r, nranks=g, **params) for r in range(g)]
    try:
        for r, s in enumerate(ss):
            s.connect_local(ss[r - 1] if r > 0 else None, ss[r + 1] if r < g - 1 else None)
        for s in ss:
            s.upload(u)
        done, conv = ph.group_run(ss, steps)
        out = np.zeros_like(u)
        for s in ss:
            s.download(out)
        return out, done, conv
    finally:
        for s in ss:
            s.close()


@pytest.mark.parametrize("g", [2, 3, 4, 8])
@pytest.mark.parametrize("arith", MODES)
def test_strip_decomposition_is_bit_identical(ph, oracle, g, arith):
    for shape in [(64, 64), (67, 130), (258, 36)]:
        if shape[0] // g < 2:
            continue
        u = field(shape, seed=g)
        want, _, _ = oracle.solve(u, 25, arith=arith)
        got, done, _ = run_strips(ph, u, g, 25, arith=arith)
        assert done == 25
        assert_same_bits(got, want, f"{g} strips {shape}")


def test_strips_convergence(ph, oracle):
    u = ph.inidat(40, 24)
    want, wdone, wconv = oracle.solve(u, 100000, arith=2, conv_every=20, conv_phase=0, step_convention=1)
    got, done, conv = run_strips(ph, u, 4, 100000, arith=2, conv_every=20, conv_phase=0, step_convention=1)
    assert wconv and conv and done == wdone
    assert_same_bits(got, want)


def test_unconnected_strip_refuses_to_run(ph):
    with ph.HeatSolver(64, 64, rank=0, nranks=2) as s:
        with pytest.raises(ph.HeatError) as ei:
            s.upload(ph.inidat(64, 64))
        assert ei.value.code == ph.api.E_STATE


# ---------------------------------------------------------------- heat_solve: upload, loop and download overlapped
@pytest.mark.parametrize("T", [1, 2, 4])
@pytest.mark.parametrize("arith", MODES)
def test_streamed_solve_is_bit_identical(ph, oracle, T, arith, monkeypatch):
    # heat_solve() streams the field through the GPU in bands of rows, pass p shifted up by p*T rows (time
    # skewing): only the ORDER of (row, update) pairs changes, so the bits must be those of the plain loop.
    # Band heights that do and do not divide the field, shorter than the total skew, a single band.
    for shape, steps in [((300, 517), 1), ((300, 517), 7), ((1000, 260), 20), ((2050, 140), 40), ((64, 64), 16), ((777, 33), 9)]:
        u = field(shape, seed=shape[0] + steps)
        want, _, _ = oracle.solve(u, steps, arith=arith)
        for band in (0, 8, 13, 64, 5000):
            if band:
                monkeypatch.setenv("HEAT_SOLVE_BAND_ROWS", str(band))
            else:
                monkeypatch.delenv("HEAT_SOLVE_BAND_ROWS", raising=False)
            got, done, conv = ph.solve(u, steps, arith=arith, temporal_block=T)
            assert (done, conv) == (steps, False)
            assert_same_bits(got, want, f"streamed {shape} steps={steps} T={T} band={band} arith={arith}")
    monkeypatch.delenv("HEAT_SOLVE_BAND_ROWS", raising=False)
    # and the same call with streaming switched off is the same field
    u = field((1000, 260), seed=5)
    a, _, _ = ph.solve(u, 20, arith=arith, temporal_block=T)
    monkeypatch.setenv("HEAT_SOLVE_STREAMED", "0")
    b, _, _ = ph.solve(u, 20, arith=arith, temporal_block=T)
    assert_same_bits(a, b, "streamed vs plain")


def test_streamed_solve_mpi_step_convention_and_ring(ph, oracle):
    # `steps`+1 updates (M:159) through the streamed path; a non-zero ring must come back untouched
    u = field((512, 300), seed=11)
    want, wdone, _ = oracle.solve(u, 30, arith=2, step_convention=1)
    got, done, _ = ph.solve(u, 30, arith=2, step_convention=ph.STEPS_MPI)
    assert done == wdone == 31
    assert_same_bits(got, want, "streamed, MPI step convention")
    assert_same_bits(got[[0, -1]], u[[0, -1]], "ring rows")
    assert_same_bits(got[:, [0, -1]], u[:, [0, -1]], "ring columns")


def test_streamed_solve_8192(ph, oracle):
    # BASELINE config[1]'s grid through heat_solve with the bench's step count: windows against the oracle
    n, k = 8192, 20
    got, done, _ = ph.solve(ph.inidat(n, n), k)
    assert done == k
    ini = oracle.inidat(n, n)
    for r0 in (0, 500, 4090, n - 60):
        band = 60
        lo, hi = max(0, r0 - k), min(n, r0 + band + k)
        want, _, _ = oracle.solve(ini[lo:hi], k, arith=0)
        assert_same_bits(got[r0:r0 + band], want[r0 - lo:r0 - lo + band], f"8192^2 streamed rows {r0}..")


@pytest.mark.parametrize("T", [1, 4])
def test_strips_from_host_rows_without_exchange(ph, oracle, T):
    # every strip solved on its own from host rows + `updates` rows of overlap (strips.solve_strip_on_host_rows):
    # the assembled field is the undecomposed one, bit for bit; strips shorter than the overlap included
    from parallel_heat_b200 import strips
    for (nx, ny), g, steps in [((640, 300), 4, 12), ((1000, 517), 3, 20), ((300, 200), 5, 40)]:
        u = field((nx, ny), seed=nx + g)
        want, _, _ = oracle.solve(u, steps, arith=1)
        got = u.copy()
        for r in range(g):
            lo, hi = strips.overlap_window(nx, g, r, steps)
            own, done = strips.solve_strip_on_host_rows(u[lo:hi].copy(), nx, g, r, steps, arith=1, temporal_block=T)  # solved in place
            r0, n = ph.strip_rows(nx, g, r)
            assert done == steps and own.shape == (n, ny)
            got[r0:r0 + n] = own
        assert_same_bits(got, want, f"{g} strips from host rows, {steps} updates, T={T}")


# ---------------------------------------------------------------- K4: temporal blocking
@pytest.mark.parametrize("T", [2, 4])
@pytest.mark.parametrize("arith", MODES)
def test_temporal_blocking_is_bit_identical(ph, oracle, T, arith):
    # T fused updates == T single updates, same bits, for any shape and any step count
    for shape in [(3, 3), (5, 9), (20, 20), (33, 47), (130, 121), (131, 250), (70, 1000), (300, 1030)]:
        u = field(shape, seed=T + shape[1])
        for steps in (T, T + 1, 3 * T + 2, 17):
            want, _, _ = oracle.solve(u, steps, arith=arith)
            got, done, _ = ph.solve(u, steps, arith=arith, temporal_block=T)
            assert done == steps
            assert_same_bits(got, want, f"T={T} {shape} steps={steps} arith={arith}")


@pytest.mark.parametrize("T", [2, 4])
def test_temporal_blocking_tile_heights(ph, oracle, T, monkeypatch):
    u = field((200, 500), seed=3)
    want, _, _ = oracle.solve(u, 2 * T, arith=0)
    for rpt in (1, 5, 64, 1000):
        monkeypatch.setenv("HEAT_ROWS_PER_TILE_TB", str(rpt))
        got, _, _ = ph.solve(u, 2 * T, arith=0, temporal_block=T)
        assert_same_bits(got, want, f"T={T} rows_per_tile_tb={rpt}")


@pytest.mark.parametrize("T", [2, 4])
@pytest.mark.parametrize("phase", [0, 1])
def test_temporal_blocking_with_convergence_check(ph, oracle, T, phase):
    # fused passes may not swallow a checked update: same stop step, same bits
    for shape, K in [((20, 20), 20), ((20, 20), 7), ((33, 47), 5), ((24, 40), 3)]:
        u = ph.inidat(*shape)
        want, wdone, wconv = oracle.solve(u, 5000, arith=2, conv_every=K, conv_phase=phase, step_convention=1)
        got, done, conv = ph.solve(u, 5000, arith=2, conv_every=K, conv_phase=phase, step_convention=1, temporal_block=T)
        assert (done, conv) == (wdone, wconv)
        assert_same_bits(got, want, f"T={T} K={K} phase={phase}")


@pytest.mark.parametrize("T", [2, 4])
@pytest.mark.parametrize("g", [2, 3, 4])
def test_temporal_blocking_in_strips(ph, oracle, T, g):
    for shape in [(64, 64), (67, 250), (258, 36)]:
        u = field(shape, seed=g * T)
        for steps in (T, 3 * T + 1, 26):
            want, _, _ = oracle.solve(u, steps, arith=1)
            got, done, _ = run_strips(ph, u, g, steps, arith=1, temporal_block=T)
            assert done == steps
            assert_same_bits(got, want, f"T={T} {g} strips {shape} steps={steps}")


def test_temporal_blocking_8192(ph, oracle):
    n, steps = 8192, 8
    with ph.HeatSolver(n, n, arith=ph.F32_FMA, temporal_block=4) as s:
        s.init_inidat()
        s.run(steps)
        got = s.download()
    want, _, _ = oracle.solve(oracle.inidat(n, n), steps, arith=0)
    assert_same_bits(got, want, "8192^2 T=4")


# ---------------------------------------------------------------- full-size checks
def test_8192_bit_exact_vs_oracle_and_properties(ph, oracle):
    # BASELINE config 2's grid; the oracle does a few updates of it in seconds
    n, steps = 8192, 6
    ini = ph.inidat(n, n)
    with ph.HeatSolver(n, n, arith=ph.F32_FMA) as s:
        s.init_inidat()
        s.run(steps)
        got = s.download()
    want, _, _ = oracle.solve(ini, steps, arith=0)
    assert_same_bits(got, want, "8192^2")
    assert got.max() <= ini.max() and got.min() >= ini.min()  # max principle
    for sl in (np.s_[0], np.s_[-1], np.s_[:, 0], np.s_[:, -1]):
        assert not got[sl].any()
    # decomposition independence at full size: 4 strips give the same bits
    parts, _, _ = run_strips(ph, ini, 4, steps, arith=ph.F32_FMA)
    assert_same_bits(parts, want, "8192^2 in 4 strips")


def test_16384_with_convergence_check(ph, oracle):
    # BASELINE config 3's grid: check every 20 updates (never converges on the wrapped field)
    n = 16384
    with ph.HeatSolver(n, n, arith=ph.F32_FMA, conv_every=20) as s:
        s.init_inidat()
        done, conv = s.run(41)
        assert done == 41 and not conv
        got = s.download()
    want, wdone, wconv = oracle.solve(oracle.inidat(n, n), 41, arith=0, conv_every=20, conv_phase=1)
    assert wdone == 41 and not wconv
    assert_same_bits(got, want, "16384^2")


# ---------------------------------------------------------------- parameters and handle API
@pytest.mark.parametrize("arith", MODES)
@pytest.mark.parametrize("T", [1, 4])
def test_anisotropic_coefficients(ph, oracle, arith, T):
    # cx weights the row+-1 difference, cy the column+-1 difference (M:168-174): with cx != cy a swap shows
    u = field((70, 260), seed=21)
    for cx, cy in [(0.05, 0.2), (0.2, 0.05), (0.0, 0.25), (0.125, 0.0)]:
        want, _, _ = oracle.solve(u, 9, arith=arith, cx=cx, cy=cy)
        got, _, _ = ph.solve(u, 9, arith=arith, cx=cx, cy=cy, temporal_block=T)
        assert_same_bits(got, want, f"cx={cx} cy={cy} arith={arith} T={T}")


def test_upload_device_set_stream_and_views(ph, oracle):
    import torch
    u = field((96, 200), seed=4)
    d = torch.from_numpy(u).cuda()
    st = torch.cuda.Stream()
    with ph.HeatSolver(96, 200, arith=ph.F32_NOFMA) as s:
        s.set_stream(st.cuda_stream)
        s.upload_device(d.data_ptr())
        s.enqueue(5)
        s.sync()
        assert 1 <= s.launch_count() <= 5  # 5 updates: one resident launch, or five separate ones
        ptr, pitch, rows, first = s.device_view(0)
        assert (pitch, rows, first) == (224, 96, 0) and ptr  # rows padded to 32 floats
        got = s.download()
        assert_same_bits(got, oracle.solve(u, 5, arith=1)[0])
        done, conv = s.run(0)
        assert (done, conv) == (0, False) and s.elapsed_ms() >= 0.0


def test_strip_errors(ph):
    with ph.HeatSolver(64, 64, rank=0, nranks=2) as a, ph.HeatSolver(64, 64, rank=1, nranks=2) as b:
        with pytest.raises(ph.HeatError) as ei:
            a.connect_local(None, None)  # strip 0 of 2 needs a south neighbour
        assert ei.value.code == ph.api.E_PEER
        with pytest.raises(ph.HeatError):
            a.connect_local(None, a)     # not adjacent
        a.connect_local(None, b)
        b.connect_local(a, None)
        a.upload(ph.inidat(64, 64))
        b.upload(ph.inidat(64, 64))
        with pytest.raises(ph.HeatError) as ei:
            a.run(3)                     # locally connected strips are driven together
        assert ei.value.code == ph.api.E_STATE
        assert ph.group_run([a, b], 3) == (3, False)
    with pytest.raises(ph.HeatError):
        ph.HeatSolver(6, 64, rank=0, nranks=2, temporal_block=4)  # 3 rows per strip cannot carry 4 ghost rows


def test_upload_and_download_of_a_strip_window(ph, oracle):
    # a rank that holds only its own window (ghost rows included) of the field
    u = field((90, 140), seed=8)
    g = 3
    ss = [ph.HeatSolver(90, 140, rank=r, nranks=g, arith=ph.F32_NOFMA, temporal_block=2) for r in range(g)]
    try:
        for r, s in enumerate(ss):
            s.connect_local(ss[r - 1] if r > 0 else None, ss[r + 1] if r < g - 1 else None)
        for s in ss:
            _, _, nloc, lo = s.device_view(0)
            s.upload_local(np.ascontiguousarray(u[lo:lo + nloc]))
        ph.group_run(ss, 7)
        out = np.zeros_like(u)
        for r, s in enumerate(ss):
            r0, n = ph.strip_rows(90, g, r)
            s.download_local(out[r0:r0 + n])
    finally:
        for s in ss:
            s.close()
    assert_same_bits(out, oracle.solve(u, 7, arith=1)[0])


@pytest.mark.parametrize("arith", MODES)
@pytest.mark.parametrize("T", [1, 4])
def test_denormals_zeros_and_huge_values(ph, oracle, arith, T):
    # denormals are kept (no flush to zero), exact zeros and values near FLT_MAX/4 behave like on the CPU
    rng = np.random.default_rng(12)
    u = (rng.standard_normal((40, 140)) * 1e-41).astype(np.float32)        # all denormal
    u[5:9, 20:60] = 0.0
    u[20:24, :] = (rng.standard_normal((4, 140)) * 1e-37).astype(np.float32)  # around the smallest normals
    u[30:34, 10:130] = (rng.standard_normal((4, 120)) * 2e37).astype(np.float32)
    assert np.isfinite(u).all()
    want, _, _ = oracle.solve(u, 9, arith=arith)
    got, _, _ = ph.solve(u, 9, arith=arith, temporal_block=T)
    assert_same_bits(got, want, f"denormals arith={arith} T={T}")


@pytest.mark.parametrize("T", [1, 4])
def test_65536_square_band_matches_oracle(ph, oracle, T):
    # BASELINE config 4's grid: 2^32 points, so every index past row 32768 needs 64-bit arithmetic (the
    # reference cannot run this size at all, C:170).  After k updates a row depends only on the k rows
    # above and below it, so a band of rows can be checked against the oracle run on that band alone.
    import torch
    free, _ = torch.cuda.mem_get_info()
    if free < 36 * 2**30:
        pytest.skip("needs 36 GiB of free device memory")
    n, k = 65536, 8
    with ph.HeatSolver(n, n, arith=ph.F32_FMA, temporal_block=T) as s:
        s.init_inidat()
        s.run(k)
        ptr, pitch, rows, first = s.device_view(0)
        assert (pitch, rows, first) == (n, n, 0)
        for lo in (0, 40000, n - 40):  # top edge, beyond the 2^31-element mark, bottom edge
            band = 40

            class _Arr:
                pass
            a = _Arr()
            a.__cuda_array_interface__ = {"shape": (band, n), "typestr": "<f4", "version": 3,
                                          "data": (ptr + 4 * pitch * lo, False)}
            got = torch.as_tensor(a, device="cuda").cpu().numpy()
            ix = np.arange(lo, lo + band, dtype=np.uint32)[:, None]
            iy = np.arange(n, dtype=np.uint32)[None, :]
            ini = (ix * (np.uint32(n) - ix - np.uint32(1)) * iy * (np.uint32(n) - iy - np.uint32(1))).view(np.int32).astype(np.float32)
            want, _, _ = oracle.solve(ini, k, arith=0)
            top = 0 if lo == 0 else k                    # rows nearer than k to a cut edge of the band are not comparable
            bot = band if lo + band == n else band - k
            assert_same_bits(got[top:bot], want[top:bot], f"65536^2 T={T} rows {lo + top}..{lo + bot}")


def test_solve_gpus_with_one_gpu_is_heat_solve(ph, oracle):
    u = ph.inidat(64, 96)
    got, done, _ = ph.solve_gpus(u, 12, 1, arith=ph.F32_NOFMA)
    assert done == 12
    assert_same_bits(got, oracle.solve(u, 12, arith=1)[0])
    with pytest.raises(ph.HeatError):
        ph.solve_gpus(u, 12, 0)


# ---------------------------------------------------------------- K5: field resident in shared memory
@pytest.mark.parametrize("arith", MODES)
def test_resident_kernel_is_bit_identical(ph, oracle, arith, monkeypatch):
    # grids that fit in the chip's shared memory run many updates in one cooperative launch
    for shape in [(3, 3), (4, 9), (20, 20), (33, 47), (150, 40), (149, 260), (300, 1030), (1024, 1024), (1500, 1500)]:
        u = field(shape, seed=shape[0])
        for steps in (4, 5, 31):
            want, _, _ = oracle.solve(u, steps, arith=arith)
            got, done, _ = ph.solve(u, steps, arith=arith)
            assert done == steps
            assert_same_bits(got, want, f"resident {shape} steps={steps} arith={arith}")
    # the same through separate launches
    monkeypatch.setenv("HEAT_RESIDENT", "0")
    u = field((300, 1030), seed=1)
    assert_same_bits(ph.solve(u, 31, arith=arith)[0], oracle.solve(u, 31, arith=arith)[0])


def test_resident_kernel_keeps_the_handle_consistent(ph, oracle):
    u = ph.inidat(200, 200)
    with ph.HeatSolver(200, 200, arith=ph.MPI_F64MIX, conv_every=20, conv_phase=ph.PHASE_MPI) as s:
        s.upload(u)
        n0 = s.launch_count()
        done, conv = s.run(61)              # 19 resident + check, three times, then one more
        assert (done, conv) == (61, False)
        assert s.launch_count() - n0 <= 9   # not one launch per update
        r = s.residual()                    # |state 61 - state 60|: the older buffer must be state 60
        got = s.download()
    a, _, _ = oracle.solve(u, 60, arith=2)
    b, _, _ = oracle.solve(u, 61, arith=2)
    assert_same_bits(got, b)
    assert r == oracle.max_abs_diff(a, b)


# ---------------------------------------------------------------- randomized differential test
def test_random_configurations_match_the_oracle(ph, oracle):
    # 160 random (shape, steps, arithmetic, coefficients, fused updates, strips, convergence) cases, fixed seed
    rng = np.random.default_rng(20261106)
    for case in range(160):
        nx, ny = int(rng.integers(3, 180)), int(rng.integers(3, 700))
        arith = int(rng.integers(0, 3))
        T = int(rng.integers(1, 5))
        steps = int(rng.integers(0, 14))
        cx, cy = (float(np.float32(v)) for v in rng.uniform(0.0, 0.24, 2))
        K = int(rng.choice([0, 0, 3, 5]))
        phase, conv_steps = int(rng.integers(0, 2)), int(rng.integers(0, 2))
        u = (rng.standard_normal((nx, ny)) * float(rng.choice([1e-3, 1.0, 1e4]))).astype(np.float32)
        kw = dict(arith=arith, cx=cx, cy=cy, conv_every=K, conv_phase=phase, step_convention=conv_steps)
        want, wdone, wconv = oracle.solve(u, steps, **kw)
        g = int(rng.integers(1, 4))
        if g > 1 and nx // g >= max(2, T) and K == 0:
            got, done, conv = run_strips(ph, u, g, steps, temporal_block=T, **kw)
        else:
            got, done, conv = ph.solve(u, steps, temporal_block=T, **kw)
        assert (done, conv) == (wdone, wconv), (case, nx, ny, kw, T, g)
        assert_same_bits(got, want, f"case {case}: {nx}x{ny} steps={steps} T={T} strips={g} {kw}")


# ---------------------------------------------------------------- 2-D blocks (the reference's topology, M:51-75)
def run_blocks(ph, u, gx, gy, steps, **params):
    nx, ny = u.shape
    bs = {(i, j): ph.HeatSolver(nx, ny, rank=i, nranks=gx, rank_y=j, nranks_y=gy, **params) for i in range(gx) for j in range(gy)}
    try:
        for (i, j), b in bs.items():
            b.connect_local(bs.get((i - 1, j)), bs.get((i + 1, j)), bs.get((i, j - 1)), bs.get((i, j + 1)))
        for b in bs.values():
            b.upload(u)
        done, conv = ph.group_run(list(bs.values()), steps)
        out = np.full_like(u, np.nan)
        for b in bs.values():
            b.download(out)
        return out, done, conv
    finally:
        for b in bs.values():
            b.close()


@pytest.mark.parametrize("arith", MODES)
@pytest.mark.parametrize("gx,gy", [(1, 2), (2, 2), (1, 3), (3, 2), (2, 4), (4, 1)])
def test_block_decomposition_is_bit_identical(ph, oracle, gx, gy, arith):
    # every block sees the same operands in the same order as the undecomposed run (SURVEY 8(e) invariant)
    for shape in [(24, 24), (67, 130), (40, 517), (130, 1030), (36, 2052)]:
        u = field(shape, seed=gx * 10 + gy)
        want, _, _ = oracle.solve(u, 21, arith=arith)
        got, done, _ = run_blocks(ph, u, gx, gy, 21, arith=arith)
        assert done == 21
        assert_same_bits(got, want, f"{gx}x{gy} blocks {shape} arith={arith}")


def test_blocks_convergence_and_inidat(ph, oracle):
    u = ph.inidat(40, 48)
    want, wdone, wconv = oracle.solve(u, 100000, arith=2, conv_every=20, conv_phase=0, step_convention=1)
    got, done, conv = run_blocks(ph, u, 2, 3, 100000, arith=2, conv_every=20, conv_phase=0, step_convention=1)
    assert wconv and conv and done == wdone
    assert_same_bits(got, want)
    # device-side inidat of a block = the block's window of the host inidat
    bs = {(i, j): ph.HeatSolver(70, 90, rank=i, nranks=2, rank_y=j, nranks_y=2) for i in range(2) for j in range(2)}
    try:
        for (i, j), b in bs.items():
            b.connect_local(bs.get((i - 1, j)), bs.get((i + 1, j)), bs.get((i, j - 1)), bs.get((i, j + 1)))
        for b in bs.values():
            b.init_inidat()
        out = np.zeros((70, 90), np.float32)
        ph.group_run(list(bs.values()), 5)
        for b in bs.values():
            b.download(out)
    finally:
        for b in bs.values():
            b.close()
    assert_same_bits(out, oracle.solve(oracle.inidat(70, 90), 5, arith=0)[0])


def test_blocks_refuse_temporal_blocking(ph):
    with pytest.raises(ph.HeatError):
        ph.HeatSolver(64, 64, rank_y=0, nranks_y=2, temporal_block=2)


def test_heat_solve_keeps_and_trims_its_buffers(ph, oracle, monkeypatch):
    # heat_solve reuses the device buffers of its previous call of the same size (stale contents must not leak into the
    # result: different fields, step counts, temporal blocking and grid types back to back), and heat_trim frees them
    import torch
    torch.cuda.synchronize()
    ph.api.trim()
    free0, _ = torch.cuda.mem_get_info()
    a, b = field((700, 517), seed=1), field((700, 517), seed=2)
    for u, steps, kw in [(a, 9, {}), (b, 3, {}), (a, 40, dict(temporal_block=4)), (b, 0, {}), (a, 7, dict(arith=2))]:
        got, done, _ = ph.solve(u, steps, **kw)
        assert done == steps
        assert_same_bits(got, oracle.solve(u, steps, arith=kw.get("arith", 0))[0], f"kept buffers, steps={steps} {kw}")
    held, _ = torch.cuda.mem_get_info()
    assert free0 - held >= 2 * 700 * 517 * 4  # the pair is still there
    got, _, _ = ph.solve(field((300, 200), seed=3), 5)  # another size replaces it
    assert_same_bits(got, oracle.solve(field((300, 200), seed=3), 5, arith=0)[0])
    ph.api.trim()
    after, _ = torch.cuda.mem_get_info()
    assert after >= free0 - (1 << 20)
    monkeypatch.setenv("HEAT_SOLVE_KEEP_BUFFERS", "0")
    ph.solve(a, 4)
    nokeep, _ = torch.cuda.mem_get_info()
    assert nokeep >= free0 - (1 << 20)
