"""Testing/CellBased/{2D,3D}_{Basic,Neumann}.exa4 restated: cell-centred finite-volume Poisson.

Fields are cell-localised with one ghost layer; `apply bc` writes the ghost cells (Dirichlet: 2 * value - inner cell,
C/s/e/boundary/ir/IR_DirichletBC.scala:42-50; Neumann(1): inner cell, IR_NeumannBC.scala:70-80) before every sweep;
restriction = mean of the 2^d children, correction = parent's value (L4_DefaultRestriction.scala:40-47).
The cycle text below is the programs' (VCycle, VCycle_0 = CG with `apply bc to VecP`, Smoother, UpResidual, Restriction,
Correction, Solve); it runs on a backend -- the CPU oracle (OracleCells) or libexamg (GpuCells, tests/test_gpu_cells.py).
"""
import math
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
PI = math.pi
OMEGA = 0.8


class CellProgram:
    """kind "cellbased": Testing/CellBased (scaled stencil, trigonometric data, prints NormError_0);
    kind "hybridcell": Testing/CommBasic/{HybridCell,Strategy26Cell} (3x3x3 blocks x 3x3x3 fragments => F = 9, unscaled
    stencil [6, -1 ...], 4.0 * RestrictionStencil, Dirichlet data x^2 - y^2/2 - z^2/2, prints the residual)."""

    def __init__(self, name, dim, neumann, max_level=6, F=4, kind="cellbased", golden=None):
        self.name, self.dim, self.neumann, self.max_level, self.F, self.min_level = name, dim, neumann, max_level, F, 0
        self.kind = kind
        self.scaled = kind == "cellbased"
        self.restrict_scale = 1.0 if kind == "cellbased" else 4.0
        self.prints = "err" if kind == "cellbased" else "res"
        self.golden = golden or "CellBased_" + name

    def expected(self):
        with open(os.path.join(GOLDEN_DIR, self.golden + ".results")) as f:
            return f.read().split()

    def stencil_coeffs(self, lvl):
        """Stencil Laplace@all in declaration order: centre, +x, -x, +y, -y[, +z, -z]"""
        if not self.scaled:
            return [2.0 * self.dim] + [-1.0] * (2 * self.dim)
        h = self.h(lvl)
        diag = 2.0 / (h * h) + 2.0 / (h * h)
        if self.dim == 3:
            diag = diag + 2.0 / (h * h)
        return [diag] + [-1.0 / (h * h)] * (2 * self.dim)

    def cells(self, lvl):
        return self.F << lvl

    def h(self, lvl):
        return 1.0 / float(self.cells(lvl))

    def centres(self, lvl):
        """vf_cellCenter = vf_nodePosition + 0.5 * vf_cellWidth"""
        h = self.h(lvl)
        return (np.arange(self.cells(lvl), dtype=np.float64) * h) + 0.5 * h

    # analytic data (Field declarations, InitRHS, NormError_0)
    def exact(self, X, Y, Z):
        if self.kind == "hybridcell":
            return ((X * X) - ((0.5 * Y) * Y)) - ((0.5 * Z) * Z)
        if self.neumann:
            e = np.cos((2.0 * PI) * X) * np.cos((2.0 * PI) * Y)
            return e * np.cos((2.0 * PI) * Z) if self.dim == 3 else e
        if self.dim == 3:
            return (np.sin(PI * X) * np.sin(PI * Y)) * np.sinh((math.sqrt(2.0) * PI) * Z)
        return np.sin(PI * X) * np.sinh(PI * Y)

    def rhs(self, X, Y, Z):
        if not self.neumann:
            return 0.0 * (X + Y + Z)
        if self.dim == 3:
            return ((((12.0 * PI) * PI) * np.cos((2.0 * PI) * X)) * np.cos((2.0 * PI) * Y)) * np.cos((2.0 * PI) * Z)
        return (((8.0 * PI) * PI) * np.cos((2.0 * PI) * X)) * np.cos((2.0 * PI) * Y) + 0.0 * Z

    def grid(self, lvl):
        c = self.centres(lvl)
        Z = c[:, None, None] if self.dim == 3 else np.zeros((1, 1, 1))
        return c[None, None, :], c[None, :, None], Z

    def face_values(self, lvl):
        """Dirichlet data on the six faces (vf_boundaryCoord: the face position in the face's dimension, cell centres in
        the others), face f = 2*d + side, shaped like the face's cells."""
        c = self.centres(lvl)
        n = c.size
        out = []
        for d in range(self.dim):
            for side in (0, 1):
                b = float(side)
                if self.dim == 2:
                    X, Y = (np.full(n, b), c) if d == 0 else (c, np.full(n, b))
                    out.append(np.ascontiguousarray(self.exact(X, Y, 0.0).reshape(1, n)))
                else:
                    A, B = np.meshgrid(c, c, indexing="xy")  # A varies fastest (second index)
                    if d == 0:
                        X, Y, Z = np.full_like(A, b), A, B      # (nz, ny)
                    elif d == 1:
                        X, Y, Z = A, np.full_like(A, b), B      # (nz, nx)
                    else:
                        X, Y, Z = A, B, np.full_like(A, b)      # (ny, nx)
                    out.append(np.ascontiguousarray(self.exact(X, Y, Z)))
        while len(out) < 6:
            out.append(None)
        return out


PROGRAMS = [CellProgram("2D_Basic", 2, False), CellProgram("2D_Neumann", 2, True),
            CellProgram("3D_Basic", 3, False), CellProgram("3D_Neumann", 3, True)]
# 576^3 cells: GPU only (tests/test_gpu_cells.py); the oracle checks the same program at a lower finest level
HYBRID_CELL = CellProgram("HybridCell", 3, False, F=9, kind="hybridcell", golden="CommBasic_HybridCell")


# ---- the programs' functions, on a backend -----------------------------------------------------------------------
def smoother(B, lvl):
    B.apply_bc("u", lvl)   # apply bc to Solution<active>; communicate ghost
    B.jacobi(lvl)          # Solution<next> = Solution<active> + ((1/diag) * 0.8) * (RHS - Laplace * Solution<active>); advance


def up_residual(B, lvl):
    B.apply_bc("u", lvl)
    B.residual(lvl)


def vcycle_0(B):
    lvl = B.prog.min_level
    up_residual(B, lvl)
    B.apply_bc("r", lvl)
    res = B.norm(lvl)
    initial = res
    B.copy("r", "p", lvl)
    for _ in range(512):
        B.apply_bc("p", lvl)
        B.apply("p", "ap", lvl)
        alpha_denom = B.dot("p", "ap", lvl)
        alpha = (res * res) / alpha_denom
        B.axpy(alpha, "p", "u", lvl)
        B.axpy(-alpha, "ap", "r", lvl)
        next_res = B.norm(lvl)
        if next_res <= 0.001 * initial:
            return
        beta = (next_res * next_res) / (res * res)
        B.xpay("r", beta, "p", lvl)
        res = next_res


def vcycle(B, lvl):
    if lvl == B.prog.min_level:
        return vcycle_0(B)
    for _ in range(3):
        smoother(B, lvl)
    up_residual(B, lvl)
    B.apply_bc("r", lvl)          # Restriction: apply bc to Residual; communicate ghost
    B.restrict(lvl)               # RHS@coarser = RestrictionStencil * Residual  (HybridCell: 4.0 * RestrictionStencil)
    B.set("u", lvl - 1, 0.0)      # SetSolution@coarser ( 0 )
    vcycle(B, lvl - 1)
    B.apply_bc("u", lvl - 1)      # Correction: apply bc to Solution@coarser; communicate
    B.correct(lvl)
    for _ in range(3):
        smoother(B, lvl)


def solve(B, printer):
    """Function Solve: prints the initial residual, NormError_0 after every cycle, the cycle count."""
    P = B.prog
    L = P.max_level
    out = []
    up_residual(B, L)
    res0 = B.norm(L)
    out.append(printer(res0))
    res, it = res0, 0
    while not (res < 1.0e-5 * res0 or it >= 100):
        it += 1
        vcycle(B, L)
        if P.neumann:  # the solution of the pure Neumann problem is fixed by its mean
            integral = B.sum("u", L) / float(P.cells(L) ** P.dim)
            B.shift("u", L, -integral)
        up_residual(B, L)
        res = B.norm(L)
        out.append(printer(B.max_error(L) if P.prints == "err" else res))
    out.append(str(it))
    return out


# ---- CPU oracle backend ----------------------------------------------------------------------------------------------
class OracleCells:
    """Padded arrays (cells + 2 per dimension): the oracle's node loops run over exactly the cells (oracle/examg_oracle.c)."""

    def __init__(self, prog):
        from oracle import oracle as og
        self.og, self.prog = og, prog
        og.set_periodic((0, 0, 0))
        P = prog
        self.a, self.st, self.faces = {}, {}, {}
        offs = [(0, 0, 0), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0)] + ([(0, 0, 1), (0, 0, -1)] if P.dim == 3 else [])
        for lvl in range(P.min_level, P.max_level + 1):
            n, h = P.cells(lvl), P.h(lvl)
            shape = ((n + 2,) * 3) if P.dim == 3 else (1, n + 2, n + 2)
            for k in ("u", "f", "r"):
                self.a[(k, lvl)] = np.zeros(shape)
            self.st[lvl] = og.Stencil.make(offs, P.stencil_coeffs(lvl))
        for k in ("p", "ap"):
            self.a[(k, P.min_level)] = np.zeros_like(self.a[("u", P.min_level)])
        L = P.max_level
        self.faces[L] = None if P.neumann else P.face_values(L)
        X, Y, Z = P.grid(L)
        inner = self._inner()
        self.a[("f", L)][inner] = np.broadcast_to(P.rhs(X, Y, Z), self.a[("f", L)][inner].shape)
        self.exact = np.zeros_like(self.a[("u", L)])
        self.exact[inner] = np.broadcast_to(P.exact(X, Y, Z), self.exact[inner].shape)

    def _inner(self):
        return (slice(1, -1),) * 3 if self.prog.dim == 3 else (slice(None), slice(1, -1), slice(1, -1))

    def apply_bc(self, k, lvl):
        kind = self.og.BC_CELL_NEUMANN if self.prog.neumann else self.og.BC_CELL_DIRICHLET
        faces = self.faces.get(lvl) if k == "u" else None
        self.og.cell_apply_bc(self.prog.dim, self.a[(k, lvl)], kind, faces)

    def jacobi(self, lvl):
        self.a[("u", lvl)] = self.og.smooth_jacobi(self.st[lvl], OMEGA, self.a[("u", lvl)], self.a[("f", lvl)])

    def residual(self, lvl):
        self.a[("r", lvl)] = self.og.residual(self.st[lvl], self.a[("u", lvl)], self.a[("f", lvl)])

    def norm(self, lvl):
        r = self.a[("r", lvl)]
        return math.sqrt(self.og.dot(r, r))

    def dot(self, x, y, lvl):
        return self.og.dot(self.a[(x, lvl)], self.a[(y, lvl)])

    def apply(self, x, y, lvl):
        self.a[(y, lvl)] = self.og.apply(self.st[lvl], self.a[(x, lvl)])

    def copy(self, src, dst, lvl):
        self.a[(dst, lvl)][self._inner()] = self.a[(src, lvl)][self._inner()]

    def set(self, k, lvl, v):
        self.a[(k, lvl)][self._inner()] = v

    def axpy(self, alpha, x, y, lvl):
        self.og.axpy(alpha, self.a[(x, lvl)], self.a[(y, lvl)])

    def xpay(self, x, beta, y, lvl):
        self.og.xpay(self.a[(x, lvl)], beta, self.a[(y, lvl)])

    def restrict(self, lvl):
        self.a[("f", lvl - 1)] = self.og.cell_restrict(self.prog.dim, self.a[("r", lvl)], self.a[("f", lvl - 1)].shape)
        if self.prog.restrict_scale != 1.0:
            self.a[("f", lvl - 1)] *= self.prog.restrict_scale  # a power of two: exact

    def correct(self, lvl):
        self.og.cell_prolong_add(self.prog.dim, self.a[("u", lvl - 1)], self.a[("u", lvl)])

    def sum(self, k, lvl):
        ones = np.zeros_like(self.a[(k, lvl)])
        ones[self._inner()] = 1.0
        return self.og.dot(self.a[(k, lvl)], ones)

    def shift(self, k, lvl, c):
        ones = np.zeros_like(self.a[(k, lvl)])
        ones[self._inner()] = 1.0
        self.og.axpy(c, ones, self.a[(k, lvl)])

    def max_error(self, lvl):
        return self.og.max_abs_diff(self.a[("u", lvl)], self.exact)

    def interior(self, k, lvl):
        return np.ascontiguousarray(self.a[(k, lvl)][self._inner()])

    def close(self):
        pass


# ---- libexamg backend (through the C ABI) -------------------------------------------------------------------------
class GpuCells:
    """Cell-localised fields of libexamg: examg_field_create_cell, Dirichlet / Neumann ghost-cell updates by
    examg_apply_bc, the generic stencil kernels on the cells, examg_restrict_cell / examg_prolong_add_cell."""

    def __init__(self, prog, num_blocks=None, rank=0, device=0, dist=None):
        """num_blocks / rank / dist: this process owns block `rank` of a num_blocks grid (x fastest) and wires the peers
        through the process group `dist` once every field exists (tests/mgpu/cell_worker.py)."""
        from exastencils_b200 import runtime as rt
        self.rt, self.prog = rt, prog
        P = prog
        nb = tuple(num_blocks) if num_blocks else (1,) * P.dim
        assert all(P.F % v == 0 for v in nb)
        self.nb, self.rank = nb + (1,) * (3 - len(nb)), rank
        self.ctx = ctx = rt.Context(P.dim, P.min_level, P.max_level, block_cells=tuple(P.F // v for v in nb),
                                    num_blocks=self.nb, rank=rank, device=device)
        self.f, self.A = {}, {}
        offs = [(0, 0, 0), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0)] + ([(0, 0, 1), (0, 0, -1)] if P.dim == 3 else [])
        kind = rt.BC_NEUMANN if P.neumann else rt.BC_DIRICHLET_CONST
        for lvl in range(P.min_level, P.max_level + 1):
            self.A[lvl] = ctx.stencil_const(lvl, offs, P.stencil_coeffs(lvl))
            self.f[("u", lvl)] = ctx.field_create_cell("Solution", lvl, 2)
            if P.restrict_scale != 1.0:
                self.f[("zero", lvl)] = ctx.field_create_cell("Zero", lvl)
            self.f[("f", lvl)] = ctx.field_create_cell("RHS", lvl)
            self.f[("r", lvl)] = ctx.field_create_cell("Residual", lvl)
            ctx.field_set_bc(self.f[("u", lvl)], kind, 0.0)
            ctx.field_set_bc(self.f[("r", lvl)], kind, 0.0)
        self.f[("p", P.min_level)] = ctx.field_create_cell("VecP", P.min_level)
        self.f[("ap", P.min_level)] = ctx.field_create_cell("VecGradP", P.min_level)
        ctx.field_set_bc(self.f[("p", P.min_level)], kind, 0.0)
        L = P.max_level
        if not P.neumann:  # Field Solution< global, ..., sin(PI x) * sinh(PI y) >@finest, x = vf_boundaryCoord_x
            if P.kind == "hybridcell":
                prog_bc = [(rt.OP_X, 0), (rt.OP_X, 0), (rt.OP_MUL, 0),
                           (rt.OP_CONST, 0.5), (rt.OP_Y, 0), (rt.OP_MUL, 0), (rt.OP_Y, 0), (rt.OP_MUL, 0), (rt.OP_SUB, 0),
                           (rt.OP_CONST, 0.5), (rt.OP_Z, 0), (rt.OP_MUL, 0), (rt.OP_Z, 0), (rt.OP_MUL, 0), (rt.OP_SUB, 0)]
            elif P.dim == 2:
                prog_bc = [(rt.OP_CONST, PI), (rt.OP_X, 0), (rt.OP_MUL, 0), (rt.OP_SIN, 0),
                           (rt.OP_CONST, PI), (rt.OP_Y, 0), (rt.OP_MUL, 0), (rt.OP_SINH, 0), (rt.OP_MUL, 0)]
            else:
                prog_bc = [(rt.OP_CONST, PI), (rt.OP_X, 0), (rt.OP_MUL, 0), (rt.OP_SIN, 0),
                           (rt.OP_CONST, PI), (rt.OP_Y, 0), (rt.OP_MUL, 0), (rt.OP_SIN, 0), (rt.OP_MUL, 0),
                           (rt.OP_CONST, math.sqrt(2.0) * PI), (rt.OP_Z, 0), (rt.OP_MUL, 0), (rt.OP_SINH, 0), (rt.OP_MUL, 0)]
            ctx.field_set_bc_expr(self.f[("u", L)], prog_bc)
        X, Y, Z = P.grid(L)
        shape = ctx.cell_shape(L)
        gshape = (shape[0] * self.nb[2], shape[1] * self.nb[1], shape[2] * self.nb[0])  # (z, y, x)
        self.exact = ctx.field_create_cell("Exact", L)
        self.ones = ctx.field_create_cell("Ones", L)
        if dist is not None:
            from exastencils_b200.distributed import wire_peers
            wire_peers(ctx, dist, device="cuda:%d" % device)
        sl = self.block_slice(shape)
        ctx.upload(self.f[("f", L)], np.ascontiguousarray(np.broadcast_to(P.rhs(X, Y, Z), gshape)[sl]))
        ctx.upload(self.exact, np.ascontiguousarray(np.broadcast_to(P.exact(X, Y, Z), gshape)[sl]))
        ctx.upload(self.ones, np.ones(shape))

    def block_slice(self, shape):
        """this block's part of a global cell array shaped (z, y, x); blocks are numbered x fastest"""
        b = [self.rank % self.nb[0], (self.rank // self.nb[0]) % self.nb[1], self.rank // (self.nb[0] * self.nb[1])]
        return tuple(slice(b[2 - k] * shape[k], (b[2 - k] + 1) * shape[k]) for k in range(3))

    def apply_bc(self, k, lvl):
        self.ctx.apply_bc(self.f[(k, lvl)])

    def jacobi(self, lvl):
        self.ctx.smooth_jacobi(self.A[lvl], self.f[("u", lvl)], self.f[("f", lvl)], OMEGA, 1)

    def residual(self, lvl):
        self.ctx.residual(self.A[lvl], self.f[("u", lvl)], self.f[("f", lvl)], self.f[("r", lvl)])

    def norm(self, lvl):
        r = self.f[("r", lvl)]
        return math.sqrt(self.ctx.dot(r, r))

    def dot(self, x, y, lvl):
        return self.ctx.dot(self.f[(x, lvl)], self.f[(y, lvl)])

    def apply(self, x, y, lvl):
        self.ctx.apply(self.A[lvl], self.f[(x, lvl)], self.f[(y, lvl)])

    def copy(self, src, dst, lvl):
        self.ctx.copy(self.f[(src, lvl)], self.f[(dst, lvl)])

    def set(self, k, lvl, v):
        self.ctx.set(self.f[(k, lvl)], v)

    def axpy(self, alpha, x, y, lvl):
        self.ctx.axpy(alpha, self.f[(x, lvl)], self.f[(y, lvl)])

    def xpay(self, x, beta, y, lvl):
        self.ctx.xpay(self.f[(x, lvl)], beta, self.f[(y, lvl)])

    def restrict(self, lvl):
        self.ctx.restrict_cell([self.f[("r", lvl)]], [self.f[("f", lvl - 1)]])
        if self.prog.restrict_scale != 1.0:  # RHS@coarser = 0 + scale * RHS@coarser (a power of two: exact)
            self.ctx.xpay(self.f[("zero", lvl - 1)], self.prog.restrict_scale, self.f[("f", lvl - 1)])

    def correct(self, lvl):
        self.ctx.prolong_add_cell([self.f[("u", lvl - 1)]], [self.f[("u", lvl)]])

    def sum(self, k, lvl):
        return self.ctx.dot(self.f[(k, lvl)], self.ones)

    def shift(self, k, lvl, c):
        self.ctx.axpy(c, self.ones, self.f[(k, lvl)])

    def max_error(self, lvl):
        return self.ctx.max_abs_diff(self.f[("u", lvl)], self.exact)

    def interior(self, k, lvl):
        return self.ctx.download_cells(self.f[(k, lvl)], lvl)

    def close(self):
        self.ctx.close()
