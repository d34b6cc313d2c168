"""Staggered-grid Stokes on the GPU (Testing/Application/ExaStokes_2D.exa4 and ExaStokes_3D.exa4, SURVEY 8f-3: face-centred
fields and the mixed-localisation block smoother): the programs through libexamg's examg_stokes_* entry points print the
reference's golden stdout and equal the CPU oracle bit for bit, loop by loop and over whole cycles."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "Application_ExaStokes_%dD.results")


def _fields_equal(gpu, ref, lvl, which):
    for k in which:
        a, b = gpu.field(lvl, k), ref.field(lvl, k)
        assert np.array_equal(a, b), "field %d on level %d: %d entries differ" % (k, lvl, int(np.sum(a != b)))


@pytest.mark.parametrize("dim", [2, 3])
def test_golden_stdout(dim, oracle):
    from exastencils_b200 import StokesProgram
    g = StokesProgram(dim=dim)
    out = g.solve(oracle.print_reduced_prec)
    g.close()
    want = [w for w in open(GOLDEN % dim).read().split("\n") if w]
    assert len(out) == (34 if dim == 2 else 45) and out == want


@pytest.mark.parametrize("dim,hi", [(2, 5), (3, 3)])
def test_loops_against_the_oracle(dim, hi, oracle):
    """Every loop on its own, from random fields (ghost layers included): apply bc on the finest and on a coarser
    level, each of the 3^dim colours, the residuals and their norms, restriction, SetSolution, correction."""
    from exastencils_b200 import StokesProgram
    from oracle import stokes
    lo = 2 if dim == 2 else 1
    g, r = StokesProgram(lo, hi, dim=dim), stokes.Solver(lo, hi, dim=dim)
    rng = np.random.default_rng(12 + dim)
    for lvl in (hi, hi - 1):
        for k in r.field_ids():
            a = rng.uniform(-1.0, 1.0, size=r.field(lvl, k).shape)
            r.field(lvl, k)[...] = a
            g.set_field(lvl, k, a)
    unknowns = r.unknown_ids()
    for lvl in (hi, hi - 1):
        g.apply_bc(lvl)
        r.apply_bc(lvl)
        _fields_equal(g, r, lvl, unknowns)
    for c2 in range(3 if dim == 3 else 1):
        for c1 in range(3):
            for c0 in range(3):
                g.smooth_color(hi, c0, c1, c2)
                r.smooth_color(hi, c0, c1, c2)
                _fields_equal(g, r, hi, unknowns)
    ng = g.norms(hi)
    r.residual(hi)
    nr = r.norms(hi)
    _fields_equal(g, r, hi, [stokes.RES + d for d in range(dim)] + [stokes.RES_P])
    assert len(ng) == len(nr) == dim + 1
    for a, b in zip(ng, nr):
        assert abs(a - b) <= 1e-14 * b
    g.restrict(hi)
    r.restrict(hi)
    _fields_equal(g, r, hi - 1, [stokes.RHS + d for d in range(dim)] + [stokes.RHS_P])
    g.correct(hi)
    r.correct(hi)
    _fields_equal(g, r, hi, unknowns)
    g.set_solution(hi - 1, 0.25)
    r.set_solution(hi - 1, 0.25)
    _fields_equal(g, r, hi - 1, r.field_ids())
    g.close()
    r.close()


@pytest.mark.parametrize("dim,hi", [(2, 4), (2, 6), (3, 4)])
def test_cycles_against_the_oracle(dim, hi, oracle):
    """Three V(3,3) cycles of the lid-driven cavity: velocities and pressure bit-identical to the oracle after every
    cycle on every level, norms to 1e-14 (double-double sums in a different order)."""
    from exastencils_b200 import StokesProgram
    from oracle import stokes
    lo = 2 if dim == 2 else 1
    g, r = StokesProgram(lo, hi, dim=dim), stokes.Solver(lo, hi, dim=dim)
    g.apply_bc(hi)
    r.apply_bc(hi)
    first = None
    for cyc in range(3):
        g.vcycle()
        r.vcycle()
        for lvl in range(lo, hi + 1):
            _fields_equal(g, r, lvl, r.unknown_ids())
        ng = g.norms(hi)
        r.residual(hi)
        nr = r.norms(hi)
        for a, b in zip(ng, nr):
            assert abs(a - b) <= 1e-14 * b
        first = first or max(nr)
    assert max(nr) < 0.1 * first
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
    with pytest.raises(rt.ExaMGError, match="colour"):
        g.smooth_color(3, 0, 0, 1)   # 2D: no third colour index
    with pytest.raises(rt.ExaMGError, match="field"):
        g.field(3, 2)                # w of a 2D program
    g.close()
    import ctypes as C
    ctx = rt.Context(2, 1, 2, block_cells=(1, 2))
    d = rt.StokesDesc(3, 3, 10, 0.02)
    assert not ctx.L.examg_stokes_create(ctx.h, C.byref(d))
    assert b"same number of cells" in ctx.L.examg_last_error()
    ctx.close()
