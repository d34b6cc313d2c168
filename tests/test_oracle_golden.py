"""Pins the CPU oracle against the reference's own golden vectors
(Testing/**/*.results, copied to tests/golden/ by tests/golden/make_golden.py).

The reference's comparator (Testing/run_test.py:12-43) accepts a line if the
strings are equal or the numbers differ by <= 1e-6 absolute; we require string
equality of every printed token (4 significant digits), which is stricter.
"""
import numpy as np
import pytest

import programs
from oracle_driver import make_oracle_solver


def _run(prog, og):
    s = make_oracle_solver(prog.cfg)
    prog.setup(s)
    s.sync_slots()
    out = programs.run_solve(prog, s, og.print_reduced_prec, og.max_abs_diff)
    s.close()
    return out


@pytest.mark.parametrize("prog", [programs.POISSON_2D, programs.POISSON_2D_FV, programs.SISC_2D_CONST, programs.SISC_2D_VAR,
                                  programs.BC_2D_POLY, programs.BC_2D_TRIG, programs.COMMBASIC_2D,
                                  programs.FMG_2D_CONST, programs.FMG_2D_POLY, programs.BC_2D_PERIODIC],
                         ids=lambda p: p.name)
def test_golden_2d(prog, oracle):
    assert _run(prog, oracle) == prog.expected()


def test_identical_golden_files():
    """Reference programs whose golden stdout is byte-identical to one the oracle and the GPU path reproduce: other
    communication strategies of the Jac program (Testing/CommBasic), the same 2D Poisson solver entered at layer 2 / 3
    instead of 4 (Testing/Poisson/2D_FD_Poisson_fromL{2,3}).  The reference's own CUDA target differs from its CPU
    build in the last printed digit (Testing/CUDA/2D_VarCoeff vs Testing/SISC/2D_VarCoeff): bit-for-bit identity with
    a generated binary is not something the reference itself offers."""
    g = programs.golden
    for name in ("CommBasic_Hybrid", "CommBasic_Strategy26", "CommBasic_Summarize"):
        assert g(name) == g("Smoothers_Jac"), name
    assert g("CommBasic_PureOMP") == g("CommBasic_PureMPI")
    for name in ("Poisson_2D_FD_Poisson_fromL2", "Poisson_2D_FD_Poisson_fromL3"):
        assert g(name) == g("Poisson_2D_FD_Poisson_fromL4"), name
    assert g("Poisson_2D_FV_Poisson_fromL2") == g("Poisson_2D_FV_Poisson_fromL4")
    cuda, cpu = g("CUDA_2D_VarCoeff"), g("SISC_2D_VarCoeff")
    assert cuda != cpu and len(cuda) == len(cpu)
    assert all(abs(float(a) - float(b)) <= 1e-3 * abs(float(b)) for a, b in zip(cuda, cpu))


@pytest.mark.slow
@pytest.mark.parametrize("prog", [programs.COMMBASIC_PURE, programs.SMOOTHERS_RBGS, programs.SMOOTHERS_JAC, programs.SISC_3D_CONST,
                                  programs.SISC_3D_VAR, programs.BC_3D_POLY, programs.BC_3D_TRIG, programs.OPTS_SEQ,
                                  programs.OPTS_PAR, programs.FMG_3D_TRIG, programs.FMG_3D_VAR,
                                  programs.BC_3D_PERIODIC], ids=lambda p: p.name)
def test_golden_3d(prog, oracle):
    assert _run(prog, oracle) == prog.expected()


def test_print_reduced_prec(oracle):
    p = oracle.print_reduced_prec
    assert p(6.714e7 + 1.0) == "6.714e+07"
    assert p(2812.2) == "2812"
    assert p(129.83) == "129.8"
    assert p(0.00078212) == "0.0007821"
    assert p(5e-13) == "EFFECTIVELY ZERO"
    assert p(5.55e-12) == "6e-12"
    assert p(5.55e-11) == "5.5e-11" or p(5.55e-11) == "5.6e-11"
    assert p(5.554e-10) == "5.55e-10"


def optflow_history(solver, fmt, max_it=128, reduction=1e-5):
    """doSolve@finest of Testing/Application/OpticalFlow2D.exa4: printed residual norms."""
    r0 = solver.update_residual_norm()
    out, r, it = [fmt(r0)], r0, 0
    while it < max_it and r > reduction * r0:
        it += 1
        solver.vcycle()
        r = solver.update_residual_norm()
        out.append(fmt(r))
    return out


def test_golden_optical_flow_2d(oracle):
    """Testing/Application/OpticalFlow2D.{exa4,knowledge,results}: levels 2..8, V(3,3) red-black 2x2 block smoother,
    CG on 4x4 cells; pins examg_oracle_optflow.c."""
    import os
    from oracle import optflow
    s = optflow.Solver((1, 1), 2, 8, alpha=1.0, relax=1.0, nu_pre=3, nu_post=3, cg_max_it=128, cg_rel_tol=1e-3)
    s.init_images()
    s.init_fields()
    got = optflow_history(s, oracle.print_reduced_prec)
    s.close()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "Application_OpticalFlow2D.results")
    assert got == open(path).read().split()


def test_golden_optical_flow_3d(oracle):
    """Testing/Application/OpticalFlow3D.{exa4,knowledge,results}: Vec3 flow on 64^3 cells, levels 2..6, V(3,3) red-black
    sweeps with the 3x3 diagonal block inverted per cell, CG on 4^3 cells; pins examg_oracle_optflow3.c."""
    import os
    from oracle import optflow3
    s = optflow3.Solver((1, 1, 1), 2, 6)
    s.init_images()
    s.init_fields()
    got = optflow_history(s, oracle.print_reduced_prec)
    s.close()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "Application_OpticalFlow3D.results")
    assert got == open(path).read().split()


def test_golden_cell_based(oracle):
    """Testing/CellBased/{2D,3D}_{Basic,Neumann}: cell-centred FV Poisson, Dirichlet ghost cells 2g - u / Neumann ghost
    cells, cell restriction and correction, CG with `apply bc to VecP`, mean subtraction of the pure Neumann problem."""
    import cell_programs as cp
    for prog in cp.PROGRAMS:
        b = cp.OracleCells(prog)
        assert cp.solve(b, oracle.print_reduced_prec) == prog.expected(), prog.name


def test_golden_fas_2d_basic(oracle):
    """Testing/NonLinear/FAS_2D_Basic.{exa4,knowledge,results}: the nonlinear full approximation scheme program (SURVEY
    8f-2) on 1152^2 cells, levels 0..7 -- pins og_loop_eval and the FAS transfer statements (tests/fas_program.py)."""
    import fas_program as fp
    b = fp.OracleFas(fp.FasProgram(), oracle)
    assert fp.solve(b, oracle.print_reduced_prec) == fp.expected()


def test_fas_cycle_of_a_linear_operator(oracle):
    """solver_useFAS of the layer-3 generator (og_config.use_fas): for a linear operator the full approximation scheme
    is the correction scheme in other words -- same iterates up to rounding, from a non-zero first guess."""
    rng = np.random.default_rng(3)
    sols = []
    for fas in (False, True):
        s = oracle.Solver(2, (3, 2), 1, 5, smoother=oracle.SMOOTH_JACOBI, omega=0.8, nu_pre=2, nu_post=2, use_fas=fas)
        if not sols:
            u0, f0 = rng.uniform(-1, 1, s.solution(5).shape), rng.uniform(-1, 1, s.rhs(5).shape)
        s.solution(5)[...] = u0
        s.rhs(5)[...] = f0
        s.sync_slots()
        hist = [s.update_residual_norm()]
        for _ in range(4):
            s.vcycle()
            hist.append(s.update_residual_norm())
        sols.append((hist, s.solution(5).copy()))
        s.close()
    (h0, u_cs), (h1, u_fas) = sols
    assert h0[-1] < 1e-3 * h0[0]
    assert np.allclose(h0, h1, rtol=1e-9)
    assert np.max(np.abs(u_cs - u_fas)) <= 1e-11 * np.max(np.abs(u_cs))


@pytest.mark.parametrize("dim", [2, 3])
def test_golden_stokes(dim, oracle):
    """Testing/Application/ExaStokes_{2D,3D}.{exa4,knowledge,results}: staggered grid (velocity component d at Face_d, p
    at cell centres), 3^dim-colour (2 dim + 1)^2 local solves, default 'linear' transfer operators per localisation,
    levels 2..8 / 1..6 -- pins examg_oracle_stokes.c: all 34 / 45 printed residual norms."""
    import os
    from oracle import stokes
    s = stokes.Solver(dim=dim)
    got = stokes.solve(s, oracle.print_reduced_prec)
    s.close()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "Application_ExaStokes_%dD.results" % dim)
    want = [w for w in open(path).read().split("\n") if w]
    assert got == want and len(got) == (34 if dim == 2 else 45)
