"""The 3D optical-flow program on the GPU (Testing/Application/OpticalFlow3D.exa4: Vec3 unknown, 3 x 3 block inverse per
cell): through libexamg's examg_optflow3_* entry points it prints the reference's golden stdout; every pointwise loop
equals the CPU oracle bit for bit, whole cycles (whose CG sums run in a different order) to 1e-12."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "Application_OpticalFlow3D.results")


def _pair(lo, hi, F=(1, 1, 1), **kw):
    from exastencils_b200 import OpticalFlow3DProgram
    from oracle import optflow3
    return OpticalFlow3DProgram(lo, hi, block_cells=F, **kw), optflow3.Solver(F, lo, hi, **kw)


def _same_images(g, r):
    """both back-ends start from the oracle's images (libm sin / cos on the host)"""
    r.init_images()
    from oracle import optflow3
    for which in (optflow3.IMG0, optflow3.IMG1):
        g.set_field(g.max_level, which, r.field(r.max_level, which))


def _equal(g, r, lvl, ids, what):
    for k in ids:
        a, b = g.field(lvl, k), r.field(lvl, k)
        assert np.array_equal(a, b), "%s: array %d on level %d: %d entries differ" % (what, k, lvl, int(np.sum(a != b)))


def test_golden_stdout(oracle):
    """levels 2..6 (64^3 cells), V(3,3): the five printed residual norms of OpticalFlow3D.results"""
    from exastencils_b200 import OpticalFlow3DProgram
    g = OpticalFlow3DProgram()
    g.init_images()
    g.init_fields()
    out = []
    its = g.solve(lambda v: out.append(oracle.print_reduced_prec(v)))
    g.close()
    assert its == 4 and out == open(GOLDEN).read().split()


@pytest.mark.parametrize("fused", [0, 1])
@pytest.mark.parametrize("F,lo,hi", [((1, 1, 1), 2, 5), ((3, 2, 1), 1, 3), ((5, 3, 2), 1, 4), ((1, 1, 1), 2, 6)])
def test_loops_against_the_oracle(F, lo, hi, fused, oracle, request):
    """InitFields, each colour, whole sweeps, residual + norm, restriction, correction: bit-identical arrays (ghost cells
    included).  fused = 1: the opt-in one-pass sweep k3_sweep (EXAMG_OF3_FUSED is read once per process: it runs in a
    process of its own; ragged tiles and several z segments at 80 x 48 x 32)."""
    import subprocess
    import sys
    if fused and os.environ.get("EXAMG_OF3_FUSED") != "1":
        env = dict(os.environ, EXAMG_OF3_FUSED="1")
        node = "%s::%s" % (__file__, request.node.name)
        out = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", node], env=env, capture_output=True,
                             text=True, timeout=600)
        assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-2000:]
        return
    from oracle import optflow3 as o3
    g, r = _pair(lo, hi, F)
    _same_images(g, r)
    g.init_fields()
    r.init_fields()
    allf = list(range(o3.NFIELDS))
    for lvl in range(lo, hi + 1):
        _equal(g, r, lvl, range(o3.G, o3.G + 6), "initGradients")
    _equal(g, r, hi, range(o3.RHS, o3.RHS + 3), "rhs")
    # a coefficient array replaced after InitFields: the split-row copies the colour passes read must follow
    gxx = r.field(hi, o3.G) + 0.25
    r.field(hi, o3.G)[...] = gxx
    g.set_field(hi, o3.G, gxx)
    # random state on two levels (ghost cells consistent: the oracle's smoother applies bc itself)
    rng = np.random.default_rng(5)
    for lvl in (hi, hi - 1):
        for k in list(range(o3.FLOW, o3.FLOW + 3)) + list(range(o3.RHS, o3.RHS + 3)):
            a = rng.uniform(-1.0, 1.0, size=r.field(lvl, k).shape)
            if k < o3.RHS:  # Neumann ghost cells
                a = np.pad(a[1:-1, 1:-1, 1:-1], 1, mode="edge")
            r.field(lvl, k)[...] = a
            g.set_field(lvl, k, a)
    for color in (0, 1, 1, 0):
        g.smooth_color(hi, color)
        r.smooth_color(hi, color)
        _equal(g, r, hi, range(o3.FLOW, o3.FLOW + 3), "colour %d" % color)
    g.smooth(hi, 2)
    r.smooth(hi, 2)
    _equal(g, r, hi, range(o3.FLOW, o3.FLOW + 3), "two sweeps")
    g.smooth(hi, 1)  # an odd number: flow now lives in its second buffers
    r.smooth(hi, 1)
    _equal(g, r, hi, range(o3.FLOW, o3.FLOW + 3), "third sweep")
    g.smooth_color(hi, 1)
    r.smooth_color(hi, 1)
    _equal(g, r, hi, range(o3.FLOW, o3.FLOW + 3), "a colour after the one-pass sweeps")
    ng, nr = g.update_residual_norm(hi), r.update_residual_norm(hi)
    _equal(g, r, hi, range(o3.RES, o3.RES + 3), "residual")
    assert abs(ng - nr) <= 1e-14 * nr
    g.restrict(hi)
    r.restrict(hi)
    _equal(g, r, hi - 1, list(range(o3.RHS, o3.RHS + 3)) + list(range(o3.FLOW, o3.FLOW + 3)), "restriction")
    a = rng.uniform(-1.0, 1.0, size=r.field(hi - 1, 0).shape)
    for k in range(3):
        r.field(hi - 1, k)[...] = a * (k + 1)
        g.set_field(hi - 1, k, a * (k + 1))
    g.correct(hi)
    r.correct(hi)
    _equal(g, r, hi, range(o3.FLOW, o3.FLOW + 3), "correction")
    assert len(allf) == 15
    g.close()
    r.close()


def test_cycles_against_the_oracle(oracle):
    """three V(3,3) cycles at 32^3: norms within 1e-12, the flow within 1e-12 of its size, CG iteration counts equal"""
    from oracle import optflow3 as o3
    g, r = _pair(2, 5)
    _same_images(g, r)
    g.init_fields()
    r.init_fields()
    assert abs(g.update_residual_norm() - r.update_residual_norm()) <= 1e-14 * r.update_residual_norm()
    for _ in range(3):
        g.vcycle()
        r.vcycle()
        assert g.last_cg_iters() == r.last_cg_iters()
        ng, nr = g.update_residual_norm(), r.update_residual_norm()
        assert abs(ng - nr) <= 1e-12 * nr
    for k in range(3):
        a, b = g.field(5, k), r.field(5, k)
        assert np.max(np.abs(a - b)) <= 1e-12 * np.max(np.abs(b))
    g.close()
    r.close()


def test_graph_and_kernel_by_kernel_cycles_agree(oracle):
    """mgCycle@finest as one CUDA graph = the same kernels launched one by one (level-wise entry points)"""
    from exastencils_b200 import OpticalFlow3DProgram
    from oracle import optflow3 as o3
    a, b = OpticalFlow3DProgram(2, 4), OpticalFlow3DProgram(2, 4)
    for p in (a, b):
        p.init_images()
        p.init_fields()
    a.vcycle()
    a.vcycle()

    def rec(p, lvl):
        if lvl == p.min_level:
            p.cg()
            return
        p.smooth(lvl, 3)
        p.residual(lvl)
        p.restrict(lvl)
        rec(p, lvl - 1)
        p.correct(lvl)
        p.smooth(lvl, 3)

    rec(b, 4)
    rec(b, 4)
    for k in range(o3.FLOW, o3.FLOW + 3):
        assert np.array_equal(a.field(4, k), b.field(4, k))
    a.close()
    b.close()


def test_argument_checks():
    from exastencils_b200 import OpticalFlow3DProgram
    from exastencils_b200 import runtime as rt
    g = OpticalFlow3DProgram(2, 3)
    with pytest.raises(rt.ExaMGError):
        g.smooth_color(3, 2)
    with pytest.raises(rt.ExaMGError):
        g.restrict(2)
    with pytest.raises(rt.ExaMGError):
        g.smooth(7, 1)
    with pytest.raises(rt.ExaMGError):
        g.field(3, 17)
    g.close()
    with pytest.raises(rt.ExaMGError):
        OpticalFlow3DProgram(0, 3)  # one cell per dimension on the coarsest level
