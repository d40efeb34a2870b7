"""qexb200_norm2_noalloc against qexb200_field_norm2, through the C ABI.

The no-allocation entry point exists for callers that cannot obtain device memory at the moment they need a norm:
a job that has filled the GPU with resident links (or shares it with another backend's memory pool), code that must
not trigger the allocator's device-wide synchronisation while other streams are busy, and tools that hold raw
device pointers but no qexb200 context.  Its reduction scratch is a static array inside the library that every call
on a GPU shares and that keeps the previous call's contents -- so the test calls it twice in a row, with DIFFERENT
inputs, and each result must match the allocating entry point on the same numbers.

The file is named so that pytest collects it last: `pytest -x` stops at the first failure, and a failure here must
not hide the parity tests.  It runs alone in a fresh process (it brings up its own context for the comparison side)
and equally after the rest of the suite (whatever context a previous test left is replaced).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NORM2_TOL = 1e-12      # the relative bar tests/test_gpu_parity.py holds norm2 to
MAX_REALS = 134217728  # QEXB200_NOALLOC_MAX_REALS, include/qex_b200.h


def _noalloc(qb, x_dev, out_dev, nreals=None):
    """One call of the new entry point on torch-owned device memory (the TEST allocates; the entry point must not)."""
    import torch
    n = x_dev.numel() if nreals is None else nreals
    rc = qb.lib().qexb200_norm2_noalloc(C.c_void_p(x_dev.data_ptr()), n, C.c_void_p(out_dev.data_ptr()),
                                        C.c_void_p(torch.cuda.current_stream().cuda_stream))
    return rc


@pytest.fixture()
def ctx(qb):
    if qb.is_initialized():
        qb.finalize()
    qb.init((8, 8, 8, 16))
    qb.set_host_layout((1, 1, 1, 1))
    yield qb
    if qb.is_initialized():
        qb.finalize()


def test_two_calls_in_a_row_match_the_allocating_norm2(ctx):
    import torch
    qb = ctx
    nsites, _ = qb.local_sites()
    rng = np.random.default_rng(20261106)
    out = torch.full((1,), -1.0, dtype=torch.float64, device="cuda")
    for call, scale in enumerate((1.0, 3.0)):
        host = np.ascontiguousarray(scale * rng.standard_normal((nsites, 3, 2)))
        f = qb.DeviceField(host)
        want = f.norm2(qb.ALL)                       # existing entry point: context scratch from cudaMalloc
        f.free()
        x = torch.from_numpy(host).to("cuda")        # same numbers in memory the caller owns
        assert _noalloc(qb, x, out) == 0, qb.lib().qexb200_last_error().decode()
        torch.cuda.synchronize()
        got = float(out.item())
        assert abs(got - want) <= NORM2_TOL * want, (call, got, want)
        # and both agree with the plain host sum
        assert abs(got - float((host ** 2).sum())) <= NORM2_TOL * want
        # same input again: the fixed summation order makes the result bit-identical
        assert _noalloc(qb, x, out) == 0
        torch.cuda.synchronize()
        assert float(out.item()) == got


def test_works_without_a_context_and_on_ragged_sizes(qb):
    """No qexb200_init: the entry point stands alone.  Sizes around the 8192-real tile, odd counts, empty input."""
    import torch
    if qb.is_initialized():
        qb.finalize()
    rng = np.random.default_rng(7)
    host = rng.standard_normal(3 * 8192 + 5)
    x = torch.from_numpy(host).to("cuda")
    out = torch.full((1,), -1.0, dtype=torch.float64, device="cuda")
    for n in (0, 1, 2, 7, 8191, 8192, 8193, 2 * 8192, 3 * 8192 + 5):
        assert _noalloc(qb, x, out, n) == 0, qb.lib().qexb200_last_error().decode()
        torch.cuda.synchronize()
        want = float((host[:n] ** 2).sum())
        assert abs(float(out.item()) - want) <= NORM2_TOL * max(want, 1e-300), (n, float(out.item()), want)


def test_rejects_what_it_cannot_do(qb):
    import torch
    x = torch.zeros(16, dtype=torch.float64, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    assert _noalloc(qb, x, out, MAX_REALS + 1) != 0
    assert "QEXB200_NOALLOC_MAX_REALS" in qb.lib().qexb200_last_error().decode()
    assert _noalloc(qb, x, out, -1) != 0
    host = np.zeros(16)
    rc = qb.lib().qexb200_norm2_noalloc(C.c_void_p(host.ctypes.data), 16, C.c_void_p(out.data_ptr()), None)
    assert rc != 0 and "device memory" in qb.lib().qexb200_last_error().decode()
    rc = qb.lib().qexb200_norm2_noalloc(C.c_void_p(x.data_ptr() + 8), 8, C.c_void_p(out.data_ptr()), None)
    assert rc != 0 and "aligned" in qb.lib().qexb200_last_error().decode()
    # a refused call must leave the next good one intact
    y = torch.ones(10, dtype=torch.float64, device="cuda")
    assert _noalloc(qb, y, out) == 0
    torch.cuda.synchronize()
    assert float(out.item()) == 10.0
