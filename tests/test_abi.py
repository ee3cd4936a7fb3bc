"""CPU-side checks of the C ABI: the library builds for sm_100a, loads, exports every symbol
include/heat.h declares, its host helpers match the reference, and -- without a GPU -- every
compute entry point fails loudly instead of falling back to a CPU path."""
import ctypes as C
import hashlib
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]


def declared_symbols():
    text = (ROOT / "include" / "heat.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(heat_[a-z0-9_]+)\s*\(", text)) - {"heat_allreduce_max_fn"})


def test_library_exports_every_declared_symbol(ph):
    lib = C.CDLL(str(ph.api.LIB_PATH))
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/heat.h but not exported"
    # and the ctypes mirror binds exactly that set
    assert sorted(ph.api.SIGNATURES) == names


def test_params_default_is_the_reference_default(ph):
    p = ph.default_params()
    assert (p.nx, p.ny, p.steps) == (20, 20, 100)  # M:7-18
    assert p.cx == np.float32(0.1) and p.cy == np.float32(0.1)  # struct Parms, M:29-32
    assert p.conv_every == 0 and p.conv_eps == np.float32(1e-3)
    assert p.nranks == 1 and p.temporal_block == 1
    assert p.grid_type == ph.GRID_F32  # the reference's `float u` (M:46)
    assert p.nz == 1 and p.cz == np.float32(0.1)  # 2-D, like the reference
    assert C.sizeof(ph.HeatParams) == 19 * 4


def test_host_inidat_and_prtdat_match_reference(ph, golden, tmp_path):
    e = [x for x in golden["plain"] if x["nx"] == 20 and x["steps"] == 100][0]
    u = ph.inidat(20, 20)
    assert hashlib.sha256(u.tobytes()).hexdigest() == e["initial_sha256"]
    ph.prtdat(u, tmp_path / "initial_im.dat")
    assert hashlib.md5((tmp_path / "initial_im.dat").read_bytes()).hexdigest() == e["initial_text_md5"]
    big = ph.inidat(1024, 1024)
    e = [x for x in golden["plain"] if x["nx"] == 1024][0]
    assert hashlib.sha256(big.tobytes()).hexdigest() == e["initial_sha256"]
    for n, want in golden["inidat_centre"].items():
        if int(n) <= 8192:
            assert float(ph.inidat(int(n), int(n))[int(n) // 2, int(n) // 2]) == want


def test_text_dump_is_printf_for_every_kind_of_float(ph, oracle, tmp_path):
    """heat_prtdat formats "%6.1f" itself (printf costs 0.7 us per value, the dump is 3 GB at 16384^2): the file
    must be byte-identical to what fprintf writes (the oracle's prtdat, M:326-341) for random bit patterns,
    exact ties (j/4: 10*v ends in .5), both zeros, denormals, values around the fast path's limit, inf, nan."""
    rng = np.random.default_rng(20)
    bits = rng.integers(0, 2**32, size=200_000, dtype=np.uint64).astype(np.uint32)
    ties = (np.arange(-4000, 4000, dtype=np.float32) * 2 + 1) / 4
    near = np.float32(1e14) * np.array([0.5, 0.99, 0.999999, 1, 1.000001, 1.01, 2, 9.99, 10, 100], np.float32)
    special = np.array([0.0, -0.0, 0.04, -0.04, 0.05, -0.05, 0.95, -0.95, 9.95, 99.95, 99999.95, 1e-45, -1e-45, 1e-38,
                        3.4028235e38, -3.4028235e38, np.inf, -np.inf, np.nan, 2147483648.0, -2147483648.0, 16777216.0,
                        8388607.5, 8388608.5, 0.25, 0.75, 1.25, 1.35, 1.45], np.float32)
    small = (rng.standard_normal(50_000) * 1000).astype(np.float32)
    v = np.concatenate([bits.view(np.float32), ties, near, -near, special, small])
    ny = 37  # not a multiple of the formatter's band of 16 text rows
    v = np.resize(v, (v.size // ny + 1) * ny).reshape(-1, ny).copy()
    ph.prtdat(v, tmp_path / "fast.dat")
    oracle.prtdat(v, tmp_path / "printf.dat")
    assert (tmp_path / "fast.dat").read_bytes() == (tmp_path / "printf.dat").read_bytes()
    one = np.array([[1.25]], np.float32)  # a single value, a single row
    ph.prtdat(one, tmp_path / "one.dat")
    assert (tmp_path / "one.dat").read_bytes() == b"   1.2\n"


def test_binary_dump_round_trip(ph, tmp_path):
    u = ph.inidat(33, 47)
    ph.prtdat_bin(u, tmp_path / "u.bin")
    assert (tmp_path / "u.bin").stat().st_size == 16 + u.nbytes
    assert np.array_equal(ph.readdat_bin(tmp_path / "u.bin").view(np.uint32), u.view(np.uint32))
    (tmp_path / "bad.bin").write_bytes(b"not a dump")
    with pytest.raises(ph.HeatError):
        ph.readdat_bin(tmp_path / "bad.bin")


def test_strip_rows_partition(ph):
    for nx, g in [(10, 3), (65536, 8), (16384, 4), (7, 7), (1026, 5)]:
        parts = [ph.strip_rows(nx, g, r) for r in range(g)]
        assert parts[0][0] == 0 and sum(n for _, n in parts) == nx
        for (a, n), (b, _) in zip(parts, parts[1:]):
            assert a + n == b
        assert max(n for _, n in parts) - min(n for _, n in parts) <= 1
    with pytest.raises(ph.HeatError):
        ph.strip_rows(4, 5, 0)
    with pytest.raises(ph.HeatError):
        ph.strip_rows(10, 2, 2)


def test_bad_arguments_are_rejected_with_a_message(ph):
    lib = ph.load()
    assert lib.heat_inidat(0, 5, None) == ph.api.E_INVALID
    assert b"heat_inidat" in lib.heat_last_error()
    assert lib.heat_solve(None, None, None, None) == ph.api.E_INVALID
    p = ph.default_params(nx=0)
    g = np.zeros((4, 4), np.float32)
    assert lib.heat_solve(C.byref(p), g.ctypes.data_as(C.POINTER(C.c_float)), None, None) == ph.api.E_INVALID
    for bad in (dict(arith=7), dict(temporal_block=5), dict(rank=2, nranks=2), dict(steps=-1), dict(conv_every=-2)):
        h = C.c_void_p()
        p = ph.default_params(**bad)
        assert lib.heat_create(C.byref(p), C.byref(h)) == ph.api.E_INVALID, bad
        assert not h.value


@pytest.mark.skipif(__import__("torch").cuda.is_available(), reason="only meaningful on a box without a GPU")
def test_no_cpu_fallback_without_a_gpu(ph):
    assert ph.device_count() == 0
    with pytest.raises(ph.HeatError) as ei:
        ph.solve(ph.inidat(20, 20), 5)
    assert ei.value.code == ph.api.E_CUDA
    with pytest.raises(ph.HeatError):
        ph.HeatSolver(20, 20)


def test_missing_library_fails_loudly(ph, monkeypatch, tmp_path):
    monkeypatch.setattr(ph.api, "_lib", None)
    monkeypatch.setattr(ph.api, "LIB_PATH", tmp_path / "nope.so")
    with pytest.raises(ph.HeatError, match="no CPU fallback"):
        ph.api.load()


def test_product_never_imports_the_oracle():
    # parity claims are void if the shipped path can reach oracle/ (tier rule 3)
    for f in list((ROOT / "parallel_heat_b200").rglob("*.py")) + list((ROOT / "parallel_heat_b200").rglob("*.cu*")) \
            + list((ROOT / "src").rglob("*.c")) + [ROOT / "include" / "heat.h"]:
        text = f.read_text()
        assert "pyoracle" not in text and "heat_oracle" not in text and "oracle/" not in text.replace("oracle/_ref", ""), f


def test_host_inidat_twins_match_the_oracle(ph, oracle):
    # heat_inidat_f64 / heat_inidat3d: the int product of M:320 (wrap included), stored as double / with a third factor pair
    assert np.array_equal(ph.inidat_f64(500, 433), oracle.inidat_f64(500, 433))
    assert np.array_equal(ph.inidat_f64(500, 433).astype(np.float32).view(np.uint32), ph.inidat(500, 433).view(np.uint32))
    a, b = ph.inidat3d(33, 40, 52), oracle.inidat3d(33, 40, 52)
    assert a.shape == (33, 40, 52) and np.array_equal(a.view(np.uint32), b.view(np.uint32))
    assert (a[0] == 0).all() and (a[:, 0] == 0).all() and (a[:, :, -1] == 0).all()  # the ring starts at zero


def test_parameter_combinations_are_refused_before_any_device_is_touched(ph):
    # validate() runs first in heat_create, so these answers do not need a GPU
    bad = [dict(nz=6), dict(nz=-1), dict(nz=8, temporal_block=2), dict(nz=8, rank_y=0, nranks_y=2),
           dict(nz=8, grid_type=ph.GRID_F64), dict(grid_type=ph.GRID_F64, temporal_block=2),
           dict(grid_type=ph.GRID_F64, rank_y=0, nranks_y=2), dict(grid_type=7), dict(temporal_block=5),
           dict(rank=2, nranks=2), dict(arith=3), dict(conv_every=-1)]
    for kw in bad:
        with pytest.raises(ph.HeatError) as ei:
            ph.HeatSolver(64, 64, **kw)
        assert ei.value.code == ph.api.E_INVALID, kw
    with pytest.raises(ph.HeatError) as ei:
        ph.HeatSolver(64, 1000000, nz=4096)  # ny*nz must fit an int
    assert ei.value.code == ph.api.E_INVALID and "fit an int" in str(ei.value)
