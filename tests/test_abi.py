"""CPU-side checks of the drop-in boundary: libexamg.so builds, loads, exports every
symbol include/examg.h declares, the ctypes binding covers all of them, and the
product path fails loudly (no CPU fallback) when there is no CUDA device."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    import exastencils_b200
    exastencils_b200.build()
    return exastencils_b200


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "examg.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(examg_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(built):
    from exastencils_b200 import runtime as rt
    syms = _declared_symbols()
    assert len(syms) >= 45
    L = ctypes.CDLL(rt.LIB_PATH)
    missing = [s for s in syms if not hasattr(L, s)]
    assert not missing, "declared in examg.h but not exported: %s" % missing


def test_binding_covers_header(built):
    from exastencils_b200 import runtime as rt
    assert sorted(rt.SIGNATURES) == _declared_symbols()
    assert rt.lib().examg_abi_version() == 1


def test_header_cites_reference():
    text = open(os.path.join(ROOT, "include", "examg.h")).read()
    assert text.count(".scala:") >= 15


def test_no_cpu_fallback(built):
    """Without a GPU the product path must raise, not compute on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from exastencils_b200 import runtime as rt, MultigridProgram
    with pytest.raises(rt.ExaMGError, match="no CUDA device"):
        rt.Context(3, 1, 3)
    with pytest.raises(rt.ExaMGError):
        MultigridProgram(3, (1, 1, 1), 1, 3)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "exastencils_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, fn)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle|#include\s+\".*oracle|dlopen.*oracle|CDLL.*oracle",
                                     src, flags=re.M), fn


def test_laplace_stencil_matches_oracle(oracle):
    """Generation-time stencil constants of the host side equal the oracle's."""
    from exastencils_b200 import laplace_stencil
    for dim, F, lvl, op, order in [(3, (1, 1, 1), 5, oracle.OP_LAPLACE_SCALED, oracle.ORDER_BENCH),
                                   (2, (4, 4), 6, oracle.OP_LAPLACE_SCALED, oracle.ORDER_TESTING),
                                   (3, (9, 9, 9), 2, oracle.OP_LAPLACE_UNSCALED, oracle.ORDER_TESTING)]:
        s = oracle.Solver(dim, F, lvl, lvl, op=op, order=order)
        st = s.stencil(lvl)
        h = [1.0 / (F[d] << lvl) if d < dim else 1.0 for d in range(3)]
        offs, coefs = laplace_stencil(dim, h, scaled=(op == oracle.OP_LAPLACE_SCALED),
                                      order="bench" if order == oracle.ORDER_BENCH else "testing")
        assert offs == st.offsets()
        assert coefs == st.coeffs()
        s.close()


def _rpn(prog, X, Y, Z, comps):
    """Reference interpreter of the postfix programs of examg_field_eval / examg_field_set_bc_expr (numpy, host libm)."""
    import math
    import numpy as np
    from exastencils_b200 import runtime as rt
    st = []
    for op, val in prog:
        if op == rt.OP_CONST: st.append(np.float64(val))
        elif op == rt.OP_X: st.append(X)
        elif op == rt.OP_Y: st.append(Y)
        elif op == rt.OP_Z: st.append(Z)
        elif op == rt.OP_COMP: st.append(comps[int(val)])
        elif op == rt.OP_NEG: st.append(-st.pop())
        elif op == rt.OP_EXP: st.append(np.vectorize(math.exp)(st.pop()))  # the host libm, not numpy's SIMD exp
        elif op == rt.OP_SIN: st.append(np.sin(st.pop()))
        elif op == rt.OP_COS: st.append(np.cos(st.pop()))
        elif op == rt.OP_SINH: st.append(np.sinh(st.pop()))
        elif op == rt.OP_SQRT: st.append(np.sqrt(st.pop()))
        else:
            b, a = st.pop(), st.pop()
            st.append(a + b if op == rt.OP_ADD else a - b if op == rt.OP_SUB else a * b if op == rt.OP_MUL else a / b)
    assert len(st) == 1
    return st[0]


def test_varcoeff27_programs_match_oracle(oracle):
    """The postfix programs the host side hands to examg_field_eval for BASELINE config 3's operator
    (exastencils_b200.program.varcoeff27_programs), interpreted with numpy, give the oracle's coefficients bit for bit
    (same association; numpy and the oracle share the host libm) -- the device differs only by its exp."""
    import numpy as np
    from exastencils_b200.program import varcoeff27_programs
    lvl, kappa = 3, 10.0
    s = oracle.Solver(3, (1, 1, 1), lvl, lvl, op=oracle.OP_VARCOEFF27, order=oracle.ORDER_TESTING, kappa=kappa,
                      smoother=oracle.SMOOTH_JACOBI)
    want = s.coeff(lvl)
    x, y, z = s.node_positions(lvl)
    X, Y, Z = np.broadcast_arrays(x[None, None, :], y[None, :, None], z[:, None, None])
    h = 1.0 / (1 << lvl)
    comps = {}
    for comp, prog in varcoeff27_programs((h, h, h), kappa):
        assert len(prog) <= 64
        comps[comp] = np.broadcast_to(_rpn(prog, X, Y, Z, comps), X.shape)
    for comp in range(27):
        assert np.array_equal(comps[comp], want[comp]), comp
    s.close()


def test_cycle_desc_from_options(built):
    """`generate solver for ... with { ... }` options / the solver_* Knowledge keys -> examg_cycle_desc (SURVEY 8f-2;
    C/s/e/solver/l3/L3_SolverForEquation.scala:84-94, defaults C/s/e/config/Knowledge.scala:203-262).  Pure host code:
    runs without a GPU."""
    from exastencils_b200 import runtime as rt
    d, ncol = rt.cycle_desc_from_options("solver_smoother_coloring = \"red-black\"", 3)   # everything else: defaults
    assert (d.smoother, d.nu_pre, d.nu_post, d.omega, d.coarse_solver, d.cg_max_it, d.cg_rel_tol, ncol) == \
           (rt.SMOOTH_RBGS, 3, 3, 1.0, rt.CGS_CG, 512, 0.001, 0)
    # Examples/lib/solver_singleScalarEquation.exa3-style option block
    d, ncol = rt.cycle_desc_from_options("""
        solver_smoother_jacobiType = true
        solver_smoother_numPre = 2
        solver_smoother_numPost = 1
        solver_smoother_damping = 0.8
        solver_cgs = "BiCGStab"
        solver_cgs_maxNumIts = 128
        solver_cgs_targetResReduction = 1e-3
        solver_targetResReduction = 1e-10
        solver_maxNumIts = 20
        solver_silent = true""", 2)
    assert (d.smoother, d.nu_pre, d.nu_post, d.omega, d.coarse_solver, d.cg_max_it, d.cg_rel_tol) == \
           (rt.SMOOTH_JACOBI, 2, 1, 0.8, rt.CGS_BICGSTAB, 128, 1e-3)
    for text, dim, want in [("solver_smoother_coloring = 4-way", 2, (rt.SMOOTH_COLORED_2, 2)),
                            ("solver_smoother_coloring = '8-way'", 3, (rt.SMOOTH_COLORED_2, 2)),
                            ("solver_smoother_coloring = 9-way; solver_cgs = MinRes", 2, (rt.SMOOTH_COLORED_3, 3)),
                            ("solver_smoother_coloring = 27-way, solver_cgs = ConjugateResidual", 3, (rt.SMOOTH_COLORED_3, 3))]:
        d, ncol = rt.cycle_desc_from_options(text, dim)
        assert (d.smoother, ncol) == want
    d, _ = rt.cycle_desc_from_options("solver_smoother_coloring = rb; solver_cgs = Smoother", 3)
    assert d.coarse_solver == rt.CGS_SMOOTHER and d.use_fas == 0
    d, _ = rt.cycle_desc_from_options("solver_smoother_coloring = rb; solver_useFAS = true", 3)
    assert d.use_fas == 1
    for bad, match in [("solver_useFAS = maybe", "solver_useFAS"), ("solver_smoother_coloring = None", "lexicographic"),
                       ("solver_smoother_coloring = 4-way", "4-way in 3D"), ("solver_smoother_coloring = rb; solver_cgs = GMRES", "solver_cgs = gmres"),
                       ("solver_smoother_coloring = rb; solver_bogus = 1", "unknown option solver_bogus"),
                       ("solver_smoother_coloring = rb; solver_smoother_numPre = two", "not a number"),
                       ("", "lexicographic")]:
        with pytest.raises(rt.ExaMGError, match=match):
            rt.cycle_desc_from_options(bad, 3)
