"""The reference's regression programs restated as data: configuration, initial
fields, what is printed per iteration and the golden stdout (4 s.f.) copied from
the reference's `.results` files (see tests/golden/README.md).

Used by the oracle-vs-golden tests (CPU) and the CUDA-vs-oracle parity tests (GPU),
so both sides run exactly the same program.
"""
import math
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden(name):
    with open(os.path.join(GOLDEN_DIR, name + ".results")) as f:
        return f.read().split()


def _grid(pos, dim):
    """Broadcastable node-position arrays X, Y[, Z] for arrays shaped (nz, ny, nx)."""
    x, y, z = pos
    X = x[None, None, :]
    Y = y[None, :, None]
    Z = z[:, None, None]
    return X, Y, Z


def _set_boundary(u, vals, dim, periodic=(0, 0, 0)):
    """`apply bc`: Dirichlet value on physical-boundary nodes only
    (C/s/e/boundary/ir/IR_DirichletBC.scala:37-40); a periodic dimension has no physical boundary."""
    vals = np.broadcast_to(vals, u.shape)
    per = list(periodic) + [0] * (3 - len(periodic))
    if not per[0]:
        u[:, :, 0] = vals[:, :, 0]
        u[:, :, -1] = vals[:, :, -1]
    if not per[1]:
        u[:, 0, :] = vals[:, 0, :]
        u[:, -1, :] = vals[:, -1, :]
    if dim == 3 and not per[2]:
        u[0, :, :] = vals[0, :, :]
        u[-1, :, :] = vals[-1, :, :]


def _interior(dim, periodic=(0, 0, 0)):
    """Loop points of `loop over F`: boundary nodes excluded, except in periodic dimensions."""
    per = list(periodic) + [0] * (3 - len(periodic))
    sl = [slice(None) if per[d] else slice(1, -1) for d in (2, 1, 0)]  # (z, y, x)
    if dim == 2:
        sl[0] = slice(None)
    return tuple(sl)


class Program:
    """One reference test program."""

    def __init__(self, name, golden_name, cfg, rhs=None, bc=None, exact=None, init=None,
                 tol=1e-5, strict=True, prints="res", print_count=True, max_it=100, fmg=False):
        self.name, self.golden_name, self.cfg, self.fmg = name, golden_name, cfg, fmg
        self.rhs, self.bc, self.exact, self.init = rhs, bc, exact, init
        self.tol, self.strict, self.prints, self.print_count, self.max_it = tol, strict, prints, print_count, max_it

    @property
    def dim(self):
        return self.cfg["dim"]

    def expected(self):
        return golden(self.golden_name)

    def setup(self, solver):
        """Fill RHS@finest, Solution@finest (interior + Dirichlet nodes) of any
        solver exposing .max_level, .node_positions(lvl), .solution(lvl), .rhs(lvl)."""
        lvl = solver.max_level
        X, Y, Z = _grid(solver.node_positions(lvl), self.dim)
        u = solver.solution(lvl)
        f = solver.rhs(lvl)
        per = self.cfg.get("periodic", (0, 0, 0))
        from oracle import oracle as og
        og.set_periodic(per)  # the oracle's loop ranges (also used for NormError on GPU results) are process-wide state
        inner = _interior(self.dim, per)
        if self.rhs is not None:
            f[inner] = np.broadcast_to(self.rhs(X, Y, Z), f.shape)[inner]
        if self.init is not None:
            u[inner] = self.init(u[inner].shape)
        if self.bc is not None:
            _set_boundary(u, self.bc(X, Y, Z), self.dim, per)
        self._exact = None
        if self.exact is not None:
            self._exact = np.ascontiguousarray(np.broadcast_to(self.exact(X, Y, Z), u.shape))

    def converged(self, res, init_res):
        return res < self.tol * init_res if self.strict else res <= self.tol * init_res


PI = math.pi

# Examples/Poisson/2D_FD_Poisson_fromL4.exa4 run by Testing/Poisson/2D_FD_Poisson_fromL4.knowledge
# (2x2 blocks x 2x2 fragments, levels 0-8  ==  global grid F=4, levels 0-8)
POISSON_2D = Program(
    "Poisson2D_fromL4", "Poisson_2D_FD_Poisson_fromL4",
    dict(dim=2, F=(4, 4), min_level=0, max_level=8, op="scaled", order="bench", smoother="rbgs", omega=0.8,
         nu_pre=3, nu_post=3, restrict_scale=1.0, cg_max_it=128, cg_rel_tol=1e-3),
    rhs=lambda X, Y, Z: PI ** 2 * np.cos(PI * X) - 4.0 * PI ** 2 * np.sin(2.0 * PI * Y),
    bc=lambda X, Y, Z: np.cos(PI * X) - np.sin(2.0 * PI * Y),
    exact=lambda X, Y, Z: np.cos(PI * X) - np.sin(2.0 * PI * Y),
    tol=1e-10, strict=False, prints="err+res", print_count=False)

# Examples/Poisson/2D_FV_Poisson_fromL4.exa4 (Testing/Poisson/2D_FV_Poisson_fromL{2,4}): the vertex-centred finite-volume
# form of the same problem -- stencil entries vf_cellWidth_y / (0.5 * (vf_cellWidth_x + vf_cellWidth_x@[-1,0])) = 1 on the
# uniform grid (centre 1+1+1+1), RHS scaled by vf_cellVolume, residual restricted with the integral weights
# (NodeRestrictionForRes = 4 x full weighting).  Errors equal the FD program's, residuals are h^2 times smaller.
def _fv_rhs(X, Y, Z):
    vol = (X[0, 0, 1] - X[0, 0, 0]) * (Y[0, 1, 0] - Y[0, 0, 0])
    return vol * (PI ** 2 * np.cos(PI * X) - 4.0 * PI ** 2 * np.sin(2.0 * PI * Y))


POISSON_2D_FV = Program(
    "Poisson2D_FV_fromL4", "Poisson_2D_FV_Poisson_fromL4",
    dict(POISSON_2D.cfg, op="unscaled", restrict_scale=4.0),
    rhs=_fv_rhs, bc=POISSON_2D.bc, exact=POISSON_2D.exact, tol=1e-10, strict=False, prints="err+res", print_count=False)

_quad_bc = lambda X, Y, Z: ((X * X) - ((0.5 * Y) * Y)) - ((0.5 * Z) * Z)

# Testing/Smoothers/RBGS.{exa4,knowledge}: 3x3x3 blocks x 3x3x3 fragments, levels 0-6 => F=9
SMOOTHERS_RBGS = Program(
    "Smoothers_RBGS", "Smoothers_RBGS",
    dict(dim=3, F=(9, 9, 9), min_level=0, max_level=6, op="unscaled", order="testing", smoother="rbgs",
         omega=1.0, nu_pre=3, nu_post=3, restrict_scale=4.0, cg_max_it=512, cg_rel_tol=1e-3),
    bc=_quad_bc)

# Testing/Smoothers/Jac.{exa4,knowledge}
SMOOTHERS_JAC = Program(
    "Smoothers_Jac", "Smoothers_Jac",
    dict(dim=3, F=(9, 9, 9), min_level=0, max_level=6, op="unscaled", order="testing", smoother="jacobi",
         omega=0.8, nu_pre=3, nu_post=3, restrict_scale=4.0, cg_max_it=512, cg_rel_tol=1e-3),
    bc=_quad_bc)

# Testing/CommBasic/{PureMPI,PureOMP}: the Jac program on 3x3x3 blocks (or fragments) of one fragment => F=3.
# CommBasic/{Hybrid,Strategy26,Summarize}.results are byte-identical to Smoothers/Jac.results (same program, other
# communication strategies), see tests/test_oracle_golden.py::test_identical_golden_files.
COMMBASIC_PURE = Program("CommBasic_PureMPI", "CommBasic_PureMPI", dict(SMOOTHERS_JAC.cfg, F=(3, 3, 3)), bc=_quad_bc)

KAPPA = 10.0
_b3 = lambda X, Y, Z: ((X - X ** 2) * (Y - Y ** 2)) * (Z - Z ** 2)
_b2 = lambda X, Y, Z: (X - X ** 2) * (Y - Y ** 2)
_rhs3 = lambda X, Y, Z: (2.0 * KAPPA) * ((((X - X ** 2) * (Y - Y ** 2)) + ((X - X ** 2) * (Z - Z ** 2)))
                                         + ((Y - Y ** 2) * (Z - Z ** 2)))
_rhs2 = lambda X, Y, Z: (2.0 * KAPPA) * ((X - X ** 2) + (Y - Y ** 2))

# Testing/SISC/3D_ConstCoeff: 2^3 blocks x 2^3 frags x fragLen 2 => F=8, levels 0-5 (256^3 cells)
SISC_3D_CONST = Program(
    "SISC_3D_ConstCoeff", "SISC_3D_ConstCoeff",
    dict(dim=3, F=(8, 8, 8), min_level=0, max_level=5, op="scaled", order="testing", smoother="jacobi", omega=0.85,
         nu_pre=3, nu_post=3, restrict_scale=1.0, cg_max_it=1024, cg_rel_tol=1e-3),
    rhs=_rhs3, bc=lambda X, Y, Z: KAPPA * _b3(X, Y, Z), exact=lambda X, Y, Z: KAPPA * _b3(X, Y, Z),
    prints="err")

SISC_3D_VAR = Program(
    "SISC_3D_VarCoeff", "SISC_3D_VarCoeff",
    dict(dim=3, F=(8, 8, 8), min_level=0, max_level=5, op="varcoeff", order="testing", kappa=KAPPA,
         smoother="jacobi", omega=0.85, nu_pre=3, nu_post=3, restrict_scale=1.0, cg_max_it=1024, cg_rel_tol=1e-3),
    rhs=_rhs3, bc=lambda X, Y, Z: 1.0 - np.exp((-1.0 * KAPPA) * _b3(X, Y, Z)),
    exact=lambda X, Y, Z: 1.0 - np.exp((-1.0 * KAPPA) * _b3(X, Y, Z)), prints="err")

# Testing/SISC/2D_*: 2x2 blocks x 2x2 frags x fragLen 2 => F=8, levels 2-7
SISC_2D_CONST = Program(
    "SISC_2D_ConstCoeff", "SISC_2D_ConstCoeff",
    dict(dim=2, F=(8, 8), min_level=2, max_level=7, op="scaled", order="testing", smoother="jacobi", omega=0.79,
         nu_pre=3, nu_post=3, restrict_scale=1.0, cg_max_it=1024, cg_rel_tol=1e-3),
    rhs=_rhs2, bc=lambda X, Y, Z: KAPPA * _b2(X, Y, Z), exact=lambda X, Y, Z: KAPPA * _b2(X, Y, Z),
    prints="err")

SISC_2D_VAR = Program(
    "SISC_2D_VarCoeff", "SISC_2D_VarCoeff",
    dict(dim=2, F=(8, 8), min_level=2, max_level=7, op="varcoeff", order="testing", kappa=KAPPA,
         smoother="jacobi", omega=0.79, nu_pre=3, nu_post=3, restrict_scale=1.0, cg_max_it=1024, cg_rel_tol=1e-3),
    rhs=_rhs2, bc=lambda X, Y, Z: 1.0 - np.exp((-1.0 * KAPPA) * _b2(X, Y, Z)),
    exact=lambda X, Y, Z: 1.0 - np.exp((-1.0 * KAPPA) * _b2(X, Y, Z)), prints="err")

# Testing/BC/{2D,3D}_{Polynomial,Trigonometric}.{exa4,knowledge}: the Smoothers/Jac cycle (unscaled stencil,
# omega = 0.8 Jacobi, V(3,3), CG) on 2^d blocks x 2^d fragments => F=4, levels 0-6, with a non-trivial Dirichlet
# expression on the finest level; prints the initial residual, then NormError_0 (max |u - exact|) per cycle.
_BC_CFG3 = dict(dim=3, F=(4, 4, 4), min_level=0, max_level=6, op="unscaled", order="testing", smoother="jacobi",
                omega=0.8, nu_pre=3, nu_post=3, restrict_scale=4.0, cg_max_it=512, cg_rel_tol=1e-3)
_BC_CFG2 = dict(_BC_CFG3, dim=2, F=(4, 4))
_poly2 = lambda X, Y, Z: (X * X) - (Y * Y) + 0.0 * Z
_trig2 = lambda X, Y, Z: np.sin(PI * X) * np.sinh(PI * Y) + 0.0 * Z
_trig3 = lambda X, Y, Z: (np.sin(PI * X) * np.sin(PI * Y)) * np.sinh((math.sqrt(2.0) * PI) * Z)
BC_2D_POLY = Program("BC_2D_Polynomial", "BC_2D_Polynomial", _BC_CFG2, bc=_poly2, exact=_poly2, prints="err")
BC_2D_TRIG = Program("BC_2D_Trigonometric", "BC_2D_Trigonometric", _BC_CFG2, bc=_trig2, exact=_trig2, prints="err")
BC_3D_POLY = Program("BC_3D_Polynomial", "BC_3D_Polynomial", _BC_CFG3, bc=_quad_bc, exact=_quad_bc, prints="err")
BC_3D_TRIG = Program("BC_3D_Trigonometric", "BC_3D_Trigonometric", _BC_CFG3, bc=_trig3, exact=_trig3, prints="err")
# Testing/CommBasic/2D.{exa4,knowledge}: the same 2D cycle on 3x3 blocks x 3x3 fragments (F=9), prints the residual
COMMBASIC_2D = Program("CommBasic_2D", "CommBasic_2D", dict(_BC_CFG2, F=(9, 9)), bc=_poly2)
# Testing/BC/{2D,3D}_Periodic: Dirichlet data x^2 in x, periodic in y (domain_rect_periodic_y), RHS = -2, scaled stencil
BC_2D_PERIODIC = Program("BC_2D_Periodic", "BC_2D_Periodic",
                         dict(_BC_CFG2, op="scaled", restrict_scale=1.0, periodic=(0, 1, 0)),
                         rhs=lambda X, Y, Z: -2.0 + 0.0 * (X + Y + Z), bc=lambda X, Y, Z: X * X + 0.0 * (Y + Z),
                         exact=lambda X, Y, Z: X * X + 0.0 * (Y + Z), prints="err")
BC_3D_PERIODIC = Program("BC_3D_Periodic", "BC_3D_Periodic",
                         dict(_BC_CFG3, op="scaled", restrict_scale=1.0, periodic=(0, 1, 1)),
                         rhs=lambda X, Y, Z: -2.0 + 0.0 * (X + Y + Z), bc=lambda X, Y, Z: X * X + 0.0 * (Y + Z),
                         exact=lambda X, Y, Z: X * X + 0.0 * (Y + Z), prints="err")
BC_PROGRAMS = [BC_2D_POLY, BC_2D_TRIG, BC_3D_POLY, BC_3D_TRIG]

# Testing/Opts/{base,tempBlock}.exa4 with the seq_* knowledge files (one fragment, levels 0-6) and
# {base_par,tempBlock_par}.exa4 with par_* (2x2x2 blocks, levels 0-8 => F=2, 513^3 nodes): the Smoothers/Jac cycle,
# Dirichlet 0, RHS 0, Solution@finest filled with (double)std::rand()/RAND_MAX by a sequential loop.  ALL knowledge
# variants (address precalculation, unrolling, vectorisation, polyhedral tiling) and the temporally blocked program
# must print the same history -- the reference's own statement that these optimisations leave results unchanged.
def _opts_init_seq(shape):
    from oracle import oracle as og
    return og.glibc_rand(int(np.prod(shape))).reshape(shape)


def _opts_init_par(shape):
    """Every MPI rank draws the same rand() sequence for its own loop points, then `apply bc` resets the physical
    boundary and the first `communicate` copies the lower block's value onto the shared nodes
    (C/s/e/communication/ir/IR_PackInfoDuplicate.scala:15-39).  Pinned empirically by par.results: the sequence
    covers ALL 257 local nodes per dimension (boundary nodes included) -- that reproduces all seven printed values,
    256 loop points per dimension (interface included, boundary excluded) gives 2.161e+04 / 257.2 / ... instead."""
    from oracle import oracle as og
    nz, ny, nx = shape  # interior of the 513^3 global grid: 511 per dimension
    m = (nx + 1) // 2 + 1   # 257 local nodes per block and dimension
    r = og.glibc_rand(m ** 3).reshape((m, m, m))
    g = np.zeros((nz + 2, ny + 2, nx + 2))
    for bz in (1, 0):
        for by in (1, 0):
            for bx in (1, 0):
                z0, y0, x0 = bz * (m - 1), by * (m - 1), bx * (m - 1)
                g[z0:z0 + m, y0:y0 + m, x0:x0 + m] = r
    return g[1:-1, 1:-1, 1:-1]


_OPTS_CFG = dict(dim=3, F=(1, 1, 1), min_level=0, max_level=6, op="unscaled", order="testing", smoother="jacobi",
                 omega=0.8, nu_pre=3, nu_post=3, restrict_scale=4.0, cg_max_it=512, cg_rel_tol=1e-3)
OPTS_SEQ = Program("Opts_seq", "Opts_seq", _OPTS_CFG, init=_opts_init_seq)
OPTS_PAR = Program("Opts_par", "Opts_par", dict(_OPTS_CFG, F=(2, 2, 2), max_level=8), init=_opts_init_par)

# BASELINE config 3: "3D 27-point variable-coefficient Poisson, damped Jacobi with temporally blocked sweeps".  The
# reference ships no such program; this is Testing/SISC/3D_VarCoeff (same coefficient, RHS, boundary data,
# omega) with the 27-point operator of oracle/examg_oracle_solver.c:og_init_varcoeff27 and V(2,2), on one fragment.
# No golden vector: parity is oracle <-> GPU only ("parity unpinned by the reference").
CONFIG3_VARCOEFF27 = Program(
    "Config3_VarCoeff27", None,
    dict(dim=3, F=(1, 1, 1), min_level=1, max_level=6, op="varcoeff27", order="testing", kappa=KAPPA,
         smoother="jacobi", omega=0.85, nu_pre=2, nu_post=2, restrict_scale=1.0, cg_max_it=1024, cg_rel_tol=1e-3),
    rhs=_rhs3, bc=lambda X, Y, Z: 1.0 - np.exp((-1.0 * KAPPA) * _b3(X, Y, Z)),
    exact=lambda X, Y, Z: 1.0 - np.exp((-1.0 * KAPPA) * _b3(X, Y, Z)), prints="res")

# Testing/FMG/*.{exa4,knowledge}: the SISC / BC programs (scaled stencils, restriction unscaled) started by a full
# multigrid sweep from the coarsest level
FMG_2D_CONST = Program("FMG_2D_ConstCoeff", "FMG_2D_ConstCoeff", SISC_2D_CONST.cfg, rhs=_rhs2, bc=SISC_2D_CONST.bc,
                       exact=SISC_2D_CONST.exact, prints="err", fmg=True)
FMG_3D_VAR = Program("FMG_3D_VarCoeff", "FMG_3D_VarCoeff", SISC_3D_VAR.cfg, rhs=_rhs3, bc=SISC_3D_VAR.bc,
                     exact=SISC_3D_VAR.exact, prints="err", fmg=True)
FMG_2D_POLY = Program("FMG_2D_Polynomial", "FMG_2D_Polynomial", dict(_BC_CFG2, op="scaled", restrict_scale=1.0),
                      bc=_poly2, exact=_poly2, prints="err", fmg=True)
FMG_3D_TRIG = Program("FMG_3D_Trigonometric", "FMG_3D_Trigonometric", dict(_BC_CFG3, op="scaled", restrict_scale=1.0),
                      bc=_trig3, exact=_trig3, prints="err", fmg=True)
FMG_PROGRAMS = [FMG_2D_CONST, FMG_2D_POLY, FMG_3D_TRIG, FMG_3D_VAR]

ALL = [POISSON_2D, SMOOTHERS_RBGS, SMOOTHERS_JAC, SISC_3D_CONST, SISC_3D_VAR, SISC_2D_CONST, SISC_2D_VAR]


# ---- full multigrid start-up (Testing/FMG/*.exa4: Function FMG@(all but finest)) -------------------------------
def _is_gpu(solver):
    return hasattr(solver, "ctx")


def _put_solution(solver, lvl, arr):
    """Solution<0>@lvl and Solution<1>@lvl := arr (the interior of the non-active slot is dead data)."""
    if _is_gpu(solver):
        from exastencils_b200 import runtime as rt
        f = solver.f_solution[lvl]
        for slot in range(rt.check(solver.ctx.L.examg_field_num_slots(solver.ctx.h, f))):
            solver.ctx.upload(f, arr, slot)
    else:
        solver.solution(lvl)[...] = arr
        solver.sync_slots()


def _get_solution(solver, lvl):
    return solver.fetch_solution(lvl) if _is_gpu(solver) else solver.solution(lvl).copy()


def run_fmg(prog, solver):
    """FMG@coarsest: on every level below the finest, coarsest first: SetFuncDir (analytic Dirichlet data on the
    boundary nodes of both slots), InitRHS, VCycle@lvl, Correction@finer (Solution@finer += P * Solution@lvl),
    ResetBC (the level's declared boundary value 0 again)."""
    dim = prog.dim
    for lvl in range(solver.min_level, solver.max_level):
        X, Y, Z = _grid(solver.node_positions(lvl), dim)
        u = _get_solution(solver, lvl)
        _set_boundary(u, prog.bc(X, Y, Z), dim)
        _put_solution(solver, lvl, u)
        f = np.zeros_like(u)
        if prog.rhs is not None:
            inner = _interior(dim)
            f[inner] = np.broadcast_to(prog.rhs(X, Y, Z), f.shape)[inner]
        if _is_gpu(solver):
            solver.ctx.upload(solver.f_rhs[lvl], f)
        else:
            solver.rhs(lvl)[...] = f
        solver.vcycle(lvl)
        if _is_gpu(solver):
            solver.ctx.prolong_add(solver.f_solution[lvl], solver.f_solution[lvl + 1])
        else:
            from oracle import oracle as og
            og.prolong_add(dim, solver.solution(lvl), solver.solution(lvl + 1))
        u = _get_solution(solver, lvl)
        _set_boundary(u, np.zeros_like(u), dim)
        _put_solution(solver, lvl, u)


def run_solve(prog, solver, printer, max_abs_err=None):
    """Solve@finest of the program: returns the list of printed tokens.

    solver: .update_residual_norm(), .vcycle(), .solution(lvl), .max_level
    printer: printWithReducedPrec restatement (value -> str)
    max_abs_err: callable(solution_array, exact_array) -> float (NormError / PrintError)
    """
    out = []
    init_res = solver.update_residual_norm()
    out.append(printer(init_res))
    res = init_res
    if prog.fmg:
        run_fmg(prog, solver)
    it = 0
    while not (it >= prog.max_it or prog.converged(res, init_res)):
        it += 1
        solver.vcycle()
        # a GPU solver downloads its Solution (get<name>()); the oracle aliases its own array
        fetch = getattr(solver, "fetch_solution", None) or (lambda: solver.solution(solver.max_level))
        if prog.prints == "err+res":  # PrintError precedes the residual update in 2D_FD_Poisson_fromL4
            out.append(printer(max_abs_err(fetch(), prog._exact)))
        res = solver.update_residual_norm()
        if prog.prints == "err":
            out.append(printer(max_abs_err(fetch(), prog._exact)))
        else:
            out.append(printer(res))
    if prog.print_count:
        out.append(str(it))
    return out
