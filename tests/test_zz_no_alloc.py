"""GPU: spb_colsum_static - the column sum for callers that cannot give the library scratch memory (a jax.ffi handler
registered without a scratch result, or a process whose own pool already holds all of HBM).  Its partial sums live in
a fixed array inside the library, so the call itself allocates nothing.  The output must agree with the ordinary
spb_colsum (caller-provided scratch) within the tolerance the other parity tests use, on two calls in a row - the
second call sees whatever the first one left in the shared array.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import max_rel_err
from oracle.metrics import TOL_BF16, TOL_F32

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _lib():
    from solvingpapers_b200 import _lib as L
    return L, L.load()


def _both(x2d):
    """(static-scratch result, ordinary result) for one matrix; every buffer is created BEFORE the calls."""
    L, lib = _lib()
    from solvingpapers_b200 import ops
    N, cols = x2d.shape
    out = torch.full((cols,), float("nan"), dtype=torch.float32, device=DEV)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    dt = L.BF16 if x2d.dtype == torch.bfloat16 else L.F32
    rc = lib.spb_colsum_static(dt, C.c_void_p(x2d.data_ptr()), x2d.stride(0), N, cols, C.c_void_p(out.data_ptr()), st)
    assert rc == 0, lib.spb_last_error()
    ref = ops.colsum(x2d)
    torch.cuda.synchronize()
    return out, ref


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_static_scratch_colsum_matches_ordinary_colsum_twice_in_a_row(dtype):
    tol = TOL_BF16 if dtype == torch.bfloat16 else TOL_F32
    rng = np.random.default_rng(21)
    # first call: wide enough that the row-chunk count is capped by the fixed array (8192 columns -> 128 chunks)
    a = torch.from_numpy(rng.standard_normal((6000, 8192)).astype(np.float32)).to(dtype).to(DEV)
    got, ref = _both(a)
    assert torch.isfinite(got).all()
    assert max_rel_err(got, ref) < tol
    assert max_rel_err(got, a.double().sum(0)) < tol
    # second call straight after, different shape and data, a strided column slice: stale partials must not leak
    b = torch.from_numpy((rng.standard_normal((777, 500)) + 3.0).astype(np.float32)).to(dtype).to(DEV)
    got2, ref2 = _both(b[:, 37:437])
    assert max_rel_err(got2, ref2) < tol
    assert max_rel_err(got2, b[:, 37:437].double().sum(0)) < tol
    # and once more with the first input: same answer as before
    got3, _ = _both(a)
    assert torch.equal(got3, got)


def test_static_scratch_colsum_rejects_sizes_beyond_its_fixed_array():
    L, lib = _lib()
    x = torch.zeros((4, 8200), dtype=torch.float32, device=DEV)
    out = torch.zeros(8200, dtype=torch.float32, device=DEV)
    rc = lib.spb_colsum_static(L.F32, C.c_void_p(x.data_ptr()), 8200, 4, 8200, C.c_void_p(out.data_ptr()), None)
    assert rc == -4  # SPB_E_LIMIT
    assert b"at most 8192 columns" in lib.spb_last_error()
