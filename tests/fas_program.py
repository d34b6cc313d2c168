"""Testing/NonLinear/FAS_2D_Basic.exa4 restated (full approximation scheme): the nonlinear problem
-Laplace u + gam * u * exp(u) = f on the unit square, 3x3 blocks of 3x3 fragments (F = 9 cells per dimension on level 0),
levels 0..7, V(3,3) with a Newton-damped Jacobi smoother, 30 smoother steps as the coarsest-level solve.

The program text (VCycle, Smoother, UpResidual, Restriction, Correction, Solve) is written once against a backend: the CPU
oracle (OracleFas: og.apply / og.loop_eval / og.restrict / og.prolong_add on numpy arrays) or libexamg (GpuFas:
examg_apply / examg_loop_eval / examg_restrict / examg_prolong_add ...).  A stencil convolution is one call
(tmp = Laplace * Solution); the pointwise nonlinear part of every loop is one postfix program over the fields at the node.
Golden stdout: tests/golden/NonLinear_FAS_2D_Basic.results (4 significant digits).
"""
import math
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "NonLinear_FAS_2D_Basic.results")
GAM, OMEGA, PI = 10.0, 0.8, math.pi
# EXAMG_OP_* (include/examg.h)
CONST, X, Y, Z, ADD, SUB, MUL, DIV, NEG, EXP, SIN, COS, SINH, SQRT, COMP, FIELD = range(16)
OFFSETS = [(0, 0, 0), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0)]  # Stencil Laplace@all, declaration order


def expected():
    with open(GOLDEN) as f:
        return f.read().split()


def laplace_coeffs(h):
    return [2.0 / (h * h) + 2.0 / (h * h)] + [-1.0 / (h * h)] * 4


def c(v):
    return [(CONST, v)]


def fld(k):
    return [(FIELD, float(k))]


def op(code):
    return [(code, 0.0)]


def gam_term(k):
    """gamSten * F_k = ( gam * exp ( F_k ) ) * F_k"""
    return c(GAM) + fld(k) + op(EXP) + op(MUL) + fld(k) + op(MUL)


def smoother_program(diag):
    """sources (Solution<active>, RHS, tmp = Laplace * Solution<active>):
    Solution + 0.8 * ( ( RHS - ( tmp + gamSten * Solution ) ) / ( diag + gam * ( 1.0 + Solution ) * exp ( Solution ) ) )"""
    num = fld(1) + fld(2) + gam_term(0) + op(ADD) + op(SUB)
    den = c(diag) + c(GAM) + c(1.0) + fld(0) + op(ADD) + op(MUL) + fld(0) + op(EXP) + op(MUL) + op(ADD)
    return fld(0) + c(OMEGA) + num + den + op(DIV) + op(MUL) + op(ADD)


def residual_program():
    """sources (Solution, RHS, tmp): RHS - ( tmp + gamSten * Solution )"""
    return fld(1) + fld(2) + gam_term(0) + op(ADD) + op(SUB)


def coarse_rhs_program():
    """sources (RHS@coarser, tmp = Laplace@coarser * Approximation@coarser, Approximation@coarser):
    RHS + ( tmp + gamSten * Approximation )"""
    return fld(0) + fld(1) + gam_term(2) + op(ADD) + op(ADD)


def sol_program():
    """solFct: ( x ** 2 - x ** 3 ) * sin ( 3.0 * PI * y )"""
    bump = op(X) + op(X) + op(MUL) + op(X) + op(X) + op(MUL) + op(X) + op(MUL) + op(SUB)
    return bump + c(3.0 * PI) + op(Y) + op(MUL) + op(SIN) + op(MUL)


def rhs_program():
    """rhsFct: ( ( 9.0 * PI ** 2 + gam * exp ( solFct ) ) * ( x ** 2 - x ** 3 ) + 6.0 * x - 2.0 ) * sin ( 3.0 * PI * y )"""
    bump = op(X) + op(X) + op(MUL) + op(X) + op(X) + op(MUL) + op(X) + op(MUL) + op(SUB)
    first = c(9.0 * (PI * PI)) + c(GAM) + sol_program() + op(EXP) + op(MUL) + op(ADD) + bump + op(MUL)
    return first + c(6.0) + op(X) + op(MUL) + op(ADD) + c(2.0) + op(SUB) + c(3.0 * PI) + op(Y) + op(MUL) + op(SIN) + op(MUL)


class FasProgram:
    def __init__(self, max_level=7, F=9):
        self.min_level, self.max_level, self.F = 0, max_level, F

    def h(self, lvl):
        return 1.0 / float(self.F << lvl)

    def nodes(self, lvl):
        return (self.F << lvl) + 1


# ---- the program text --------------------------------------------------------------------------------------------------

def smoother(b, lvl):
    b.apply(lvl, "u", "tmp")
    b.eval(lvl, "u", ["u", "rhs", "tmp"], smoother_program(laplace_coeffs(b.prog.h(lvl))[0]), to_next=True)
    b.advance(lvl)


def up_residual(b, lvl):
    b.apply(lvl, "u", "tmp")
    b.eval(lvl, "res", ["u", "rhs", "tmp"], residual_program())


def restriction(b, lvl):
    b.restrict(lvl, "res", "rhs")            # RHS@coarser = RestrictionNode * Residual
    b.restrict(lvl, "u", "approx")           # Approximation@coarser = RestrictionNode * Solution
    b.copy(lvl - 1, "approx", "u")           # Solution<active>@coarser = Approximation@coarser
    b.apply(lvl - 1, "approx", "tmp")        # f = r + A v
    b.eval(lvl - 1, "rhs", ["rhs", "tmp", "approx"], coarse_rhs_program())


def correction(b, lvl):
    b.axpy(lvl - 1, -1.0, "approx", "u")     # e = u - v
    b.prolong_add(lvl)


def vcycle(b, lvl):
    if lvl == b.prog.min_level:
        for _ in range(30):
            smoother(b, lvl)
        return
    for _ in range(3):
        smoother(b, lvl)
    up_residual(b, lvl)
    restriction(b, lvl)
    vcycle(b, lvl - 1)
    correction(b, lvl)
    for _ in range(3):
        smoother(b, lvl)


def solve(b, printer):
    """Application + Solve: the printed tokens (initial residual, the error after every cycle, the cycle count)"""
    L = b.prog.max_level
    b.init_finest()
    out = []
    up_residual(b, L)
    res0 = b.norm(L)
    res = res0
    out.append(printer(res0))
    num_it = 0
    while not (res < 1.0e-5 * res0 or num_it >= 100):
        num_it += 1
        vcycle(b, L)
        up_residual(b, L)
        res = b.norm(L)
        out.append(printer(b.max_error(L)))
    out.append(str(num_it))
    return out


# ---- backends ----------------------------------------------------------------------------------------------------------

class OracleFas:
    def __init__(self, prog, og):
        self.prog, self.og = prog, og
        og.set_periodic((0, 0, 0))
        self.f = {}
        self.cur = {}
        self.st = {}
        for lvl in range(prog.min_level, prog.max_level + 1):
            n = prog.nodes(lvl)
            for name in ("res", "rhs", "approx", "tmp", "exact"):
                self.f[(name, lvl)] = np.zeros((1, n, n))
            self.f[("u", lvl)] = [np.zeros((1, n, n)), np.zeros((1, n, n))]
            self.cur[lvl] = 0
            self.st[lvl] = og.Stencil.make(OFFSETS, laplace_coeffs(prog.h(lvl)))

    def close(self):
        pass

    def arr(self, name, lvl, nxt=False):
        if name == "u":
            return self.f[("u", lvl)][self.cur[lvl] ^ (1 if nxt else 0)]
        return self.f[(name, lvl)]

    def init_finest(self):
        L = self.prog.max_level
        h = [self.prog.h(L)] * 2
        self.og.loop_eval(self.arr("rhs", L), [], rhs_program(), h)
        self.og.loop_eval(self.arr("exact", L), [], sol_program(), h)

    def apply(self, lvl, src, dst):
        self.f[(dst, lvl)][...] = self.og.apply(self.st[lvl], self.arr(src, lvl))

    def eval(self, lvl, dst, srcs, prog, to_next=False):
        self.og.loop_eval(self.arr(dst, lvl, nxt=to_next), [self.arr(s, lvl) for s in srcs], prog, [self.prog.h(lvl)] * 2)

    def advance(self, lvl):
        self.cur[lvl] ^= 1

    def restrict(self, lvl, src, dst):
        out = self.arr(dst, lvl - 1)
        r = self.og.restrict(2, self.arr(src, lvl), out.shape, 1.0)
        self.og.copy_interior(r.reshape(out.shape), out)

    def copy(self, lvl, src, dst):
        self.og.copy_interior(self.arr(src, lvl), self.arr(dst, lvl))

    def axpy(self, lvl, alpha, x, y):
        self.og.axpy(alpha, self.arr(x, lvl), self.arr(y, lvl))

    def prolong_add(self, lvl):
        self.og.prolong_add(2, self.arr("u", lvl - 1), self.arr("u", lvl))

    def norm(self, lvl):
        r = self.arr("res", lvl)
        return math.sqrt(self.og.dot(r, r))

    def max_error(self, lvl):
        return self.og.max_abs_diff(self.arr("u", lvl), self.arr("exact", lvl))

    def solution(self, lvl):
        return self.arr("u", lvl)[0]


class GpuFas:
    def __init__(self, prog):
        from exastencils_b200 import runtime as rt
        self.prog, self.rt = prog, rt
        self.ctx = ctx = rt.Context(2, prog.min_level, prog.max_level, block_cells=(prog.F, prog.F))
        self.f, self.st = {}, {}
        for lvl in range(prog.min_level, prog.max_level + 1):
            for name in ("res", "rhs", "approx", "tmp", "exact"):
                self.f[(name, lvl)] = ctx.field_create(name, lvl)
            self.f[("u", lvl)] = ctx.field_create("Solution", lvl, 2)
            for name in ("u", "res", "approx"):  # Field ...< global, BasicComm, 0.0 >
                ctx.field_set_bc(self.f[(name, lvl)], rt.BC_DIRICHLET_CONST, 0.0)
            self.st[lvl] = ctx.stencil_const(lvl, OFFSETS, laplace_coeffs(prog.h(lvl)))

    def close(self):
        self.ctx.close()

    def init_finest(self):
        L = self.prog.max_level
        self.ctx.loop_eval(self.f[("rhs", L)], [], rhs_program())
        self.ctx.loop_eval(self.f[("exact", L)], [], sol_program())

    def apply(self, lvl, src, dst):
        self.ctx.apply(self.st[lvl], self.f[(src, lvl)], self.f[(dst, lvl)])

    def eval(self, lvl, dst, srcs, prog, to_next=False):
        d = self.f[(dst, lvl)]
        slot = (self.ctx.field_active_slot(d) + 1) % 2 if to_next else -1
        self.ctx.loop_eval(d, [self.f[(s, lvl)] for s in srcs], prog, dst_slot=slot)

    def advance(self, lvl):
        self.ctx.field_advance(self.f[("u", lvl)])

    def restrict(self, lvl, src, dst):
        self.ctx.restrict(self.f[(src, lvl)], self.f[(dst, lvl - 1)], 1.0)

    def copy(self, lvl, src, dst):
        self.ctx.copy(self.f[(src, lvl)], self.f[(dst, lvl)])

    def axpy(self, lvl, alpha, x, y):
        self.ctx.axpy(alpha, self.f[(x, lvl)], self.f[(y, lvl)])

    def prolong_add(self, lvl):
        self.ctx.prolong_add(self.f[("u", lvl - 1)], self.f[("u", lvl)])

    def norm(self, lvl):
        r = self.f[("res", lvl)]
        return math.sqrt(self.ctx.dot(r, r))

    def max_error(self, lvl):
        return self.ctx.max_abs_diff(self.f[("u", lvl)], self.f[("exact", lvl)])

    def solution(self, lvl):
        return self.ctx.download(self.f[("u", lvl)], lvl)
