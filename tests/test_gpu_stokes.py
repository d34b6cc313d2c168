"""Staggered-grid Stokes on the GPU (Testing/Application/ExaStokes_2D.exa4, SURVEY 8f-3: face-centred fields and the
mixed-localisation 5 x 5 block smoother): the program through libexamg's examg_stokes_* entry points prints the
reference's golden stdout and equals the CPU oracle bit for bit, loop by loop and over whole cycles."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "Application_ExaStokes_2D.results")


def _fields_equal(gpu, ref, lvl, which):
    from oracle import stokes as os_
    for k in which:
        a, b = gpu.field(lvl, k), ref.field(lvl, k)
        assert np.array_equal(a, b), "field %d on level %d: %d entries differ" % (k, lvl, int(np.sum(a != b)))
    del os_


def test_golden_stdout(oracle):
    from exastencils_b200 import StokesProgram
    g = StokesProgram()
    out = g.solve(oracle.print_reduced_prec)
    g.close()
    want = open(GOLDEN).read().split("\n")
    assert len(out) == 34 and out == want[:34]


def test_loops_against_the_oracle(oracle):
    """Every loop on its own, from random fields (ghost layers included): apply bc on the finest and on a coarser
    level, each of the nine colours, the residuals and their norms, restriction, SetSolution, correction."""
    from exastencils_b200 import StokesProgram
    from oracle import stokes
    lo, hi = 2, 5
    g, r = StokesProgram(lo, hi), stokes.Solver(lo, hi)
    rng = np.random.default_rng(12)
    for lvl in (hi, hi - 1):
        for k in range(9):
            a = rng.uniform(-1.0, 1.0, size=r.field(lvl, k).shape)
            r.field(lvl, k)[...] = a
            g.set_field(lvl, k, a)
    everything = list(range(9))
    for lvl in (hi, hi - 1):
        g.apply_bc(lvl)
        r.apply_bc(lvl)
        _fields_equal(g, r, lvl, [stokes.U, stokes.V, stokes.P])
    for c1 in range(3):
        for c0 in range(3):
            g.smooth_color(hi, c0, c1)
            r.smooth_color(hi, c0, c1)
            _fields_equal(g, r, hi, [stokes.U, stokes.V, stokes.P])
    ng = g.norms(hi)
    r.residual(hi)
    nr = r.norms(hi)
    _fields_equal(g, r, hi, [stokes.RES_U, stokes.RES_V, stokes.RES_P])
    for a, b in zip(ng, nr):
        assert abs(a - b) <= 1e-14 * b
    g.restrict(hi)
    r.restrict(hi)
    _fields_equal(g, r, hi - 1, [stokes.RHS_U, stokes.RHS_V, stokes.RHS_P])
    g.correct(hi)
    r.correct(hi)
    _fields_equal(g, r, hi, [stokes.U, stokes.V, stokes.P])
    g.set_solution(hi - 1, 0.25)
    r.set_solution(hi - 1, 0.25)
    _fields_equal(g, r, hi - 1, everything)
    g.close()
    r.close()


@pytest.mark.parametrize("hi", [4, 6])
def test_cycles_against_the_oracle(hi, oracle):
    """Three V(3,3) cycles of the lid-driven cavity: velocities and pressure bit-identical to the oracle after every
    cycle on every level, norms to 1e-14 (double-double sums in a different order)."""
    from exastencils_b200 import StokesProgram
    from oracle import stokes
    g, r = StokesProgram(2, hi), stokes.Solver(2, hi)
    g.apply_bc(hi)
    r.apply_bc(hi)
    first = None
    for cyc in range(3):
        g.vcycle()
        r.vcycle()
        for lvl in range(2, hi + 1):
            _fields_equal(g, r, lvl, [stokes.U, stokes.V, stokes.P])
        ng = g.norms(hi)
        r.residual(hi)
        nr = r.norms(hi)
        for a, b in zip(ng, nr):
            assert abs(a - b) <= 1e-14 * b
        first = first or max(nr)
    assert max(nr) < 0.05 * first
    g.close()
    r.close()


def test_argument_checks():
    from exastencils_b200 import runtime as rt
    from exastencils_b200 import StokesProgram
    g = StokesProgram(2, 3)
    with pytest.raises(rt.ExaMGError, match="level"):
        g.smooth(7)
    with pytest.raises(rt.ExaMGError, match="coarsest"):
        g.restrict(2)
    with pytest.raises(rt.ExaMGError, match="colour"):
        g.smooth_color(3, 3, 0)
    g.close()
    import ctypes as C
    ctx = rt.Context(3, 1, 2)
    d = rt.StokesDesc(3, 3, 10, 0.02)
    assert not ctx.L.examg_stokes_create(ctx.h, C.byref(d))
    assert b"2D" in ctx.L.examg_last_error()
    ctx.close()
