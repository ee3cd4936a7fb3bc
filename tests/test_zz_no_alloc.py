"""heat_step_residual_noalloc(): the update + residual for callers that cannot allocate.

Some users of this library never get to call cudaMalloc: their application handed the whole
GPU to its own memory pool at start-up, or they sit in a time loop where the device-wide stall
of an allocation is not acceptable.  For them the library carries its reduction workspace
inside its own image.  This test calls that entry point through the C ABI, twice in a row (the
second call meets whatever the first one left in the shared workspace), and requires the very
same bits -- field and residual -- as the ordinary entry point heat_step_residual_device().

Kept in a file that sorts last so that a `pytest -x` run has already executed every other test
before this one can stop it.  It must pass on its own in a fresh process as well.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("arith", [0, 1, 2])
def test_two_calls_in_a_row_match_the_ordinary_entry_point(ph, arith):
    import torch
    for shape in [(20, 20), (33, 47), (300, 1030), (2050, 517)]:
        nx, ny = shape
        u = torch.from_numpy((np.random.default_rng(nx + ny).standard_normal(shape) * 1e3).astype(np.float32)).cuda()
        stream = torch.cuda.current_stream().cuda_stream
        src_new, src_old = u, u
        for call in range(2):
            # ordinary entry point: the caller owns (and zeroes) the reduction word
            out_old = torch.full(shape, -3.0, device="cuda")
            word = torch.zeros(1, dtype=torch.int32, device="cuda")
            ph.step_residual_device(src_old.data_ptr(), out_old.data_ptr(), nx, ny, word.data_ptr(), arith=arith, stream=stream)
            # entry point under test: only input, output and the 4-byte result are passed in;
            # the result slot is deliberately dirty to show that nothing needs clearing
            out_new = torch.full(shape, -3.0, device="cuda")
            res = torch.full((1,), 123.0, device="cuda")
            ph.api.step_residual_noalloc(src_new.data_ptr(), out_new.data_ptr(), nx, ny, res.data_ptr(), arith=arith, stream=stream)
            torch.cuda.synchronize()
            assert np.array_equal(bits(out_new.cpu().numpy()), bits(out_old.cpu().numpy())), (shape, call, "field")
            assert res.cpu().numpy().view(np.uint32)[0] == word.cpu().numpy().view(np.uint32)[0], (shape, call, "residual")
            assert float(res.item()) > 0.0
            src_new, src_old = out_new, out_old  # second call: feed the first call's result back in


@pytest.mark.parametrize("arith", [0, 1, 2])
def test_field_and_residual_match_the_cpu_oracle(ph, oracle, arith):
    # the same entry point against the CPU restatement of the reference's update (M:168-174 / C:155-161) and of
    # its convergence quantity max|old-new| (M:245, C:67) -- so this file does not lean on any other GPU path
    import torch
    for shape in [(20, 20), (33, 47), (300, 1030)]:
        nx, ny = shape
        u = (np.random.default_rng(7 * nx + ny).standard_normal(shape) * 1e3).astype(np.float32)
        want = oracle.step(u, arith=arith, out=np.full(shape, -3.0, np.float32))
        inner = (slice(1, -1), slice(1, -1))
        want_res = oracle.max_abs_diff(np.ascontiguousarray(u[inner]), np.ascontiguousarray(want[inner]))
        d_in = torch.from_numpy(u).cuda()
        d_out = torch.full(shape, -3.0, device="cuda")
        res = torch.full((1,), 123.0, device="cuda")
        for call in range(2):
            ph.api.step_residual_noalloc(d_in.data_ptr(), d_out.data_ptr(), nx, ny, res.data_ptr(), arith=arith,
                                         stream=torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            assert np.array_equal(bits(d_out.cpu().numpy()), bits(want)), (shape, call, "field vs oracle")
            assert res.cpu().numpy().view(np.uint32)[0] == np.float32(want_res).view(np.uint32), (shape, call, "residual vs oracle")


def test_no_interior_gives_zero_residual(ph):
    import torch
    u = torch.ones((2, 9), device="cuda")
    out = torch.full((2, 9), -3.0, device="cuda")
    res = torch.full((1,), 5.0, device="cuda")
    ph.api.step_residual_noalloc(u.data_ptr(), out.data_ptr(), 2, 9, res.data_ptr())
    torch.cuda.synchronize()
    assert float(res.item()) == 0.0 and bool((out == -3.0).all())


def test_rows_longer_than_the_fixed_workspace_are_refused(ph):
    import torch
    small = torch.zeros(64, device="cuda")
    with pytest.raises(ph.HeatError) as ei:
        ph.api.step_residual_noalloc(small.data_ptr(), small.data_ptr(), 3, ph.api.NOALLOC_MAX_NY + 1, small.data_ptr())
    assert ei.value.code == ph.api.E_INVALID and "HEAT_NOALLOC_MAX_NY" in str(ei.value)
    torch.cuda.synchronize()  # nothing was launched, nothing faulted
