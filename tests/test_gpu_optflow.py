"""Optical-flow path: CUDA vs oracle through the C ABI.

Kernels: fields bit-exact (shared arithmetic contract, oracle/examg_oracle_optflow.c header), sums within 1e-14
relative.  Program: the printed history equals the reference's golden Testing/Application/OpticalFlow2D.results and
matches the oracle within 1e-12 relative; u, v within 1e-12 of the oracle's (the coarse CG's step lengths come
from sums, so fields are not bit-identical after a whole cycle).
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL_SUM = 1e-14
RTOL_NORM = 1e-12


def _rand(shape, seed):
    return np.random.default_rng(seed).uniform(-1.0, 1.0, size=shape)


def _pad(a):
    """dense (ny, nx) -> oracle layout (ny+2, nx+2) with Neumann ghost cells"""
    from oracle import optflow as of
    out = np.zeros((a.shape[0] + 2, a.shape[1] + 2))
    out[1:-1, 1:-1] = a
    return of.neumann(out)


def _inner(a):
    return np.ascontiguousarray(a[1:-1, 1:-1])


CASES = [((1, 1), 3), ((3, 5), 3), ((1, 1), 6), ((5, 3), 5), ((1, 1), 9), ((1, 2), 1)]


@pytest.mark.parametrize("cells,lvl", CASES)
@pytest.mark.parametrize("relax", [1.0, 0.8])
def test_system2_kernels(cells, lvl, relax, oracle):
    from exastencils_b200 import runtime as rt
    from oracle import optflow as of
    ctx = rt.Context(2, max(lvl - 1, 0), lvl, block_cells=cells + (1,))
    nx, ny, _ = ctx.cells(lvl)
    seed = 100 * lvl + 10 * cells[0] + cells[1]
    h = ctx.width(lvl)
    L5 = of.laplace(h)
    a2 = 1.3 * 1.3
    names = ["u", "v", "fu", "fv", "ru", "rv", "cuu", "cuv", "cvv"]
    host = {n: _rand((ny, nx), seed + k) for k, n in enumerate(names)}
    host["cuu"] = np.abs(host["cuu"]) * 50.0
    host["cvv"] = np.abs(host["cvv"]) * 50.0
    host["cuv"] = host["cuv"] * 5.0
    f = {}
    for n in names:
        f[n] = ctx.field_create_cell(n, lvl)
        ctx.upload(f[n], host[n][None])
    for n in ("u", "v", "ru", "rv"):
        ctx.field_set_bc(f[n], rt.BC_NEUMANN)
        ctx.apply_bc(f[n])
    offs = [(0, 0, 0), (-1, 0, 0), (1, 0, 0), (0, -1, 0), (0, 1, 0)]
    lap = ctx.stencil_const(lvl, offs, list(L5))
    sys = ctx.system2_create(lap, a2, f["cuu"], f["cuv"], f["cvv"])
    P = {n: _pad(host[n]) for n in names}

    def dl(n):
        return ctx.download_cells(f[n], lvl)[0]

    # residual (+ norm)
    n2 = ctx.system2_residual(sys, f["u"], f["v"], f["fu"], f["fv"], f["ru"], f["rv"], norm=True)
    ru, rv = of.residual(L5, a2, P["cuu"], P["cuv"], P["cvv"], P["u"], P["v"], P["fu"], P["fv"])
    assert np.array_equal(dl("ru"), _inner(ru)) and np.array_equal(dl("rv"), _inner(rv))
    n2_ref = of.dot2(ru, ru, rv, rv)
    assert abs(n2 - n2_ref) <= RTOL_SUM * n2_ref
    # red-black block smoother, single colours then whole sweeps (ghost cells refreshed by the kernel itself)
    u, v = P["u"].copy(), P["v"].copy()
    for color in (0, 1, 1, 0):
        ctx.system2_smooth_rb_color(sys, f["u"], f["v"], f["fu"], f["fv"], relax, color)
        of.smooth_color(L5, a2, P["cuu"], P["cuv"], P["cvv"], u, v, P["fu"], P["fv"], relax, color)
        of.neumann(u)
        of.neumann(v)
        assert np.array_equal(dl("u"), _inner(u)) and np.array_equal(dl("v"), _inner(v)), "colour %d" % color
    ctx.system2_smooth_rb(sys, f["u"], f["v"], f["fu"], f["fv"], relax, 3)
    for _ in range(3):
        for color in (0, 1):
            of.smooth_color(L5, a2, P["cuu"], P["cuv"], P["cvv"], u, v, P["fu"], P["fv"], relax, color)
            of.neumann(u)
            of.neumann(v)
    assert np.array_equal(dl("u"), _inner(u)) and np.array_equal(dl("v"), _inner(v))
    # the residual after smoothing reads the ghost cells the smoother wrote
    ctx.system2_residual(sys, f["u"], f["v"], f["fu"], f["fv"], f["ru"], f["rv"])
    ru, rv = of.residual(L5, a2, P["cuu"], P["cuv"], P["cvv"], u, v, P["fu"], P["fv"])
    assert np.array_equal(dl("ru"), _inner(ru)) and np.array_equal(dl("rv"), _inner(rv))
    # cell restriction / prolongation
    if lvl - 1 >= ctx.min_level and lvl >= 1:
        cl = lvl - 1
        c1, c2 = ctx.field_create_cell("c1", cl), ctx.field_create_cell("c2", cl)
        ctx.restrict_cell([f["ru"], f["rv"]], [c1, c2])
        of.neumann(ru)
        of.neumann(rv)
        e1, e2 = of.restrict(ru), of.restrict(rv)
        g1, g2 = ctx.download_cells(c1, cl)[0], ctx.download_cells(c2, cl)[0]
        assert np.array_equal(g1, _inner(e1)) and np.array_equal(g2, _inner(e2))
        ctx.prolong_add_cell([c1, c2], [f["u"], f["v"]])
        of.prolong_add(e1, u)
        of.prolong_add(e2, v)
        of.neumann(u)
        of.neumann(v)
        assert np.array_equal(dl("u"), _inner(u)) and np.array_equal(dl("v"), _inner(v))
        # ... and the ghost cells the prolongation refreshed
        ctx.system2_residual(sys, f["u"], f["v"], f["fu"], f["fv"], f["ru"], f["rv"])
        ru, rv = of.residual(L5, a2, P["cuu"], P["cuv"], P["cvv"], u, v, P["fu"], P["fv"])
        assert np.array_equal(dl("ru"), _inner(ru)) and np.array_equal(dl("rv"), _inner(rv))
    ctx.close()


def _oracle_solver(F, lo, hi, **kw):
    from oracle import optflow as of
    s = of.Solver(F, lo, hi, **kw)
    s.init_images()
    s.init_fields()
    return s


def _gpu_program(F, lo, hi, images=None, **kw):
    from exastencils_b200 import OpticalFlowProgram
    from exastencils_b200.optflow import analytic_images
    p = OpticalFlowProgram(F, lo, hi, **kw)
    p.set_images(*(images if images is not None else analytic_images(*p.cells())))
    return p


def test_golden_stdout(oracle):
    """Testing/Application/OpticalFlow2D: levels 2..8, V(3,3); printed history = the reference's .results"""
    p = _gpu_program((1, 1), 2, 8)
    out = [oracle.print_reduced_prec(r) for r in p.solve()]
    p.close()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "Application_OpticalFlow2D.results")
    assert out == open(path).read().split()


@pytest.mark.parametrize("F,lo,hi,kw", [
    ((1, 1), 2, 8, dict()),
    ((3, 2), 1, 6, dict(alpha=0.7, relax=0.9, nu_pre=2, nu_post=1)),
    ((1, 1), 2, 10, dict(nu_pre=2, nu_post=2)),
])
def test_program_vs_oracle(F, lo, hi, kw, oracle):
    from exastencils_b200 import optflow as gof
    from oracle import optflow as of
    ref = _oracle_solver(F, lo, hi, **kw)
    # the same frames on both sides (libm's and numpy's sin/cos may differ in the last bit)
    gpu = _gpu_program(F, lo, hi, images=(_inner(ref.field(hi, of.IMG0)), _inner(ref.field(hi, of.IMG1))), **kw)
    # InitFields: coefficient hierarchy and rhs bit-exact
    for lvl in range(lo, hi + 1):
        for which in (gof.IXIX, gof.IXIY, gof.IYIY):
            assert np.array_equal(gpu.download(which, lvl), _inner(ref.field(lvl, which))), (lvl, which)
    for which in (gof.RHS_U, gof.RHS_V):
        assert np.array_equal(gpu.download(which), _inner(ref.field(hi, which)))
    r_ref, r_gpu = ref.update_residual_norm(), gpu.residual_norm()
    assert abs(r_ref - r_gpu) <= RTOL_NORM * r_ref
    for it in range(4):
        ref.vcycle()
        r_ref = ref.update_residual_norm()
        r_gpu = gpu.cycles(1)
        assert abs(r_ref - r_gpu) <= RTOL_NORM * max(r_ref, 1e-300), "cycle %d: %r vs %r" % (it, r_ref, r_gpu)
        assert gpu.last_cg_iters() == ref.last_cg_iters()
        for which in (gof.U, gof.V):
            e = _inner(ref.field(hi, which))
            g = gpu.download(which)
            assert np.max(np.abs(g - e)) <= 1e-12 * max(np.max(np.abs(e)), 1e-300), (it, which)
    ref.close()
    gpu.close()


def test_coarse_cg(oracle):
    """mgCycle@coarsest alone on a 12 x 8 grid with random data"""
    from exastencils_b200 import runtime as rt
    from oracle import optflow as of
    lvl = 2
    ctx = rt.Context(2, lvl, lvl, block_cells=(3, 2, 1))
    nx, ny, _ = ctx.cells(lvl)
    L5 = of.laplace(ctx.width(lvl))
    names = ["u", "v", "fu", "fv", "ru", "rv", "pu", "pv", "apu", "apv", "cuu", "cuv", "cvv"]
    host = {n: _rand((ny, nx), 7 + k) for k, n in enumerate(names)}
    host["cuu"], host["cvv"] = np.abs(host["cuu"]) * 20, np.abs(host["cvv"]) * 20
    f = {}
    for n in names:
        f[n] = ctx.field_create_cell(n, lvl)
        ctx.upload(f[n], host[n][None])
    for n in ("u", "v"):
        ctx.field_set_bc(f[n], rt.BC_NEUMANN)
        ctx.apply_bc(f[n])
    lap = ctx.stencil_const(lvl, [(0, 0, 0), (-1, 0, 0), (1, 0, 0), (0, -1, 0), (0, 1, 0)], list(L5))
    sys = ctx.system2_create(lap, 1.0, f["cuu"], f["cuv"], f["cvv"])
    its = ctx.system2_cg_coarse(sys, [f[n] for n in names[:10]], 128, 1e-3)
    # oracle: a one-level solver whose mgCycle@coarsest is the CG
    s = of.Solver((3, 2), lvl, lvl, alpha=1.0)
    for n, which in (("u", of.U), ("v", of.V), ("fu", of.RHS_U), ("fv", of.RHS_V), ("cuu", of.CUU), ("cuv", of.CUV),
                     ("cvv", of.CVV)):
        s.field(lvl, which)[...] = _pad(host[n])
    s.vcycle()
    assert its == s.last_cg_iters()
    for n, which in (("u", of.U), ("v", of.V), ("ru", of.RES_U), ("rv", of.RES_V)):
        e = _inner(s.field(lvl, which))
        g = ctx.download_cells(f[n], lvl)[0]
        assert np.max(np.abs(g - e)) <= 1e-12 * np.max(np.abs(e)), n
    s.close()
    ctx.close()
