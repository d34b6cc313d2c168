"""examg_dot_dense_noalloc (scratch in a file-scope device array, no allocation inside the call) against examg_dot,
through the C ABI.  Collected last on purpose (file name).

The entry point has no counterpart in the reference
(include/examg.h).  Tolerance: 1e-14 relative, the bar of the other reduction tests (both sides accumulate in
double-double; only the final rounding can differ)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL_SUM = 1e-14


def _dense_dot(L, check, a_t, b_t, out_t, stream=None):
    check(L.examg_dot_dense_noalloc(a_t.data_ptr(), b_t.data_ptr(), a_t.numel(), out_t.data_ptr(), stream))


def test_dot_dense_noalloc_matches_examg_dot_twice():
    import torch
    from exastencils_b200 import runtime as rt
    L = rt.lib()
    lvl = 6  # 65^3 nodes, 63^3 loop points: several thread blocks, a ragged last chunk
    ctx = rt.Context(3, lvl, lvl, block_cells=(1, 1, 1))
    shape = ctx.shape(lvl)
    fa, fb = ctx.field_create("a", lvl, 1, 1), ctx.field_create("b", lvl, 1, 1)
    out = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
    rng = np.random.default_rng(20240607)
    # two calls in a row on different data: the second one finds the first one's partials and ticket in the array
    for trial in range(2):
        a, b = rng.uniform(-1.0, 1.0, size=shape), rng.uniform(-1.0, 1.0, size=shape)
        ctx.upload(fa, a)
        ctx.upload(fb, b)
        want = ctx.dot(fa, fb)  # loop points = all nodes but the physical boundary
        a_t = torch.from_numpy(np.ascontiguousarray(a[1:-1, 1:-1, 1:-1])).cuda()
        b_t = torch.from_numpy(np.ascontiguousarray(b[1:-1, 1:-1, 1:-1])).cuda()
        torch.cuda.synchronize()
        _dense_dot(L, rt.check, a_t, b_t, out)
        torch.cuda.synchronize()
        got = float(out.item())
        assert np.isfinite(got)
        assert abs(got - want) <= RTOL_SUM * max(abs(want), 1.0), (trial, got, want)
        # and against a float128-free host check: math.fsum of the rounded products is the exact sum, rounded once
        import math
        exact = math.fsum((a[1:-1, 1:-1, 1:-1] * b[1:-1, 1:-1, 1:-1]).ravel().tolist())
        assert abs(got - exact) <= RTOL_SUM * max(abs(exact), 1.0)
    # same input again: identical bits (the shared array's leftovers do not leak into the result)
    _dense_dot(L, rt.check, a_t, b_t, out)
    torch.cuda.synchronize()
    assert float(out.item()) == got
    ctx.close()


def test_dot_dense_noalloc_on_a_stream_and_small_sizes():
    import torch
    from exastencils_b200 import runtime as rt
    L = rt.lib()
    stream = torch.cuda.Stream()
    rng = np.random.default_rng(7)
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    for n in (0, 1, 255, 65536, 65537, 300001):
        a, b = rng.uniform(-1.0, 1.0, size=n), rng.uniform(-1.0, 1.0, size=n)
        a_t, b_t = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
        torch.cuda.synchronize()
        with torch.cuda.stream(stream):
            _dense_dot(L, rt.check, a_t, b_t, out, C.c_void_p(stream.cuda_stream))
        stream.synchronize()
        import math
        exact = math.fsum((a * b).tolist())
        assert abs(float(out.item()) - exact) <= RTOL_SUM * max(abs(exact), 1.0), n


def test_dot_dense_noalloc_rejects_oversize_and_null():
    import torch
    from exastencils_b200 import runtime as rt
    L = rt.lib()
    x = torch.zeros(8, dtype=torch.float64, device="cuda")
    rc = L.examg_dot_dense_noalloc(x.data_ptr(), x.data_ptr(), (1 << 28) + 1, x.data_ptr(), None)
    assert rc == -1
    msg = L.examg_last_error().decode()
    assert "exceed the maximum" in msg and "EXAMG_NOALLOC_MAX_N" in msg
    assert L.examg_dot_dense_noalloc(None, x.data_ptr(), 8, x.data_ptr(), None) == -1
    assert L.examg_dot_dense_noalloc(x.data_ptr(), x.data_ptr(), -1, x.data_ptr(), None) == -1
