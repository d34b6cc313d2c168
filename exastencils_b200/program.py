"""Host side of the generated multigrid solver, on top of libexamg.

This mirrors, function by function, what the ExaStencils generator prints for the
L4 programs of the hot path (Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4,
Examples/Poisson/2D_FD_Poisson_fromL4.exa4, Testing/Smoothers/*.exa4,
Testing/SISC/*.exa4): the leveled functions `Solve@finest`, `mgCycle@lvl`,
`ResNorm@lvl`, `UpResidual@lvl` keep their names and meaning, but every
`loop over` / `communicate` / `apply bc` is a call into libexamg instead of an
inlined loop nest.  Two execution modes:

  mode="library" : mgCycle@finest is one examg_solver_vcycle call (CUDA graph)
  mode="calls"   : mgCycle@lvl issues one libexamg call per L4 statement, exactly
                   the sequence the new generator back-end emits

Both give bit-identical fields.  No numerics happen in Python.
"""
import ctypes as C
import math

import numpy as np

from . import runtime as rt


def laplace_stencil(dim, h, scaled=True, order="bench"):
    """Stencil Laplace@lvl of the L4 programs: offsets in declaration order and the
    generation-time constants `2/hx**2 + 2/hy**2 [+ 2/hz**2]`, `-1/h**2`
    (Benchmark/Poisson3D/...exa4:36-44; unscaled: Testing/Smoothers/RBGS.exa4)."""
    offsets = [(0, 0, 0)]
    for d in range(dim):
        first = -1 if order == "bench" else 1
        for s in (first, -first):
            o = [0, 0, 0]
            o[d] = s
            offsets.append(tuple(o))
    if not scaled:
        return offsets, [2.0 * dim] + [-1.0] * (2 * dim)
    diag = 0.0
    for d in range(dim):
        t = 2.0 / (h[d] * h[d])
        diag = t if d == 0 else diag + t
    coeffs = [diag]
    for d in range(dim):
        coeffs += [-1.0 / (h[d] * h[d])] * 2
    return offsets, coeffs


def varcoeff27_offsets():
    """27 entries, lexicographic (dx fastest, dz slowest); the centre is entry 13."""
    return [(dx, dy, dz) for dz in (-1, 0, 1) for dy in (-1, 0, 1) for dx in (-1, 0, 1)]


VARCOEFF27_WEIGHTS = (0.0, 14.0 / 30.0, 3.0 / 30.0, 1.0 / 30.0)  # faces, edges, corners of the compact 27-point Laplacian


def varcoeff27_programs(hs, kappa):
    """Postfix programs (examg_field_eval) that assemble BASELINE config 3's operator, the 27-point analogue of
    InitLaplace in Testing/SISC/3D_VarCoeff.exa4:212-240: entry d != 0 is  ((-1 * w_|d|) * a(x + d*h/2)) / (h*h)
    with getCoefficient a = exp(kappa * (((x - x*x) * (y - y*y)) * (z - z*z))), the centre is -(sum of the others,
    in entry order).  Returns [(component, program)] in the order they have to run."""
    out = []
    h = hs[0]  # scaling 1/hx^2 (cubic cells in all measured configurations); midpoints use each dimension's width
    width = {rt.OP_X: hs[0], rt.OP_Y: hs[1], rt.OP_Z: hs[2]}

    def pos(op, d):  # x + (0.5 * d) * h
        return [(op, 0.0), (rt.OP_CONST, (0.5 * d) * width[op]), (rt.OP_ADD, 0.0)]

    def bump(op, d):  # X - X * X
        return pos(op, d) + pos(op, d) + pos(op, d) + [(rt.OP_MUL, 0.0), (rt.OP_SUB, 0.0)]

    for e, (dx, dy, dz) in enumerate(varcoeff27_offsets()):
        if e == 13:
            continue
        w = VARCOEFF27_WEIGHTS[(dx != 0) + (dy != 0) + (dz != 0)]
        prog = [(rt.OP_CONST, -1.0 * w), (rt.OP_CONST, kappa)]
        prog += bump(rt.OP_X, dx) + bump(rt.OP_Y, dy) + [(rt.OP_MUL, 0.0)] + bump(rt.OP_Z, dz) + [(rt.OP_MUL, 0.0)]
        prog += [(rt.OP_MUL, 0.0), (rt.OP_EXP, 0.0), (rt.OP_MUL, 0.0), (rt.OP_CONST, h * h), (rt.OP_DIV, 0.0)]
        out.append((e, prog))
    others = [e for e in range(27) if e != 13]
    prog = [(rt.OP_COMP, float(others[0]))]
    for e in others[1:]:
        prog += [(rt.OP_COMP, float(e)), (rt.OP_ADD, 0.0)]
    out.append((13, prog + [(rt.OP_NEG, 0.0)]))
    return out


class MultigridProgram:
    """One generated solver instance (one block on one GPU)."""

    def __init__(self, dim, block_cells, min_level, max_level, op="scaled", order="bench", kappa=0.0,
                 smoother="rbgs", omega=0.8, nu_pre=3, nu_post=3, restrict_scale=1.0, cg_max_it=128,
                 cg_rel_tol=1e-3, num_blocks=(1, 1, 1), rank=0, device=0, mode="library", stream=None,
                 domain_lo=(0.0, 0.0, 0.0), domain_hi=(1.0, 1.0, 1.0), periodic=(0, 0, 0), coarse_solver="cg", use_fas=False):
        assert mode in ("library", "calls")
        assert not use_fas or mode == "library", "solver_useFAS: the library cycle"
        assert coarse_solver == "cg" or mode == "library", "the call-by-call mirror spells out the CG of the program text"
        self.coarse_solver = {"cg": rt.CGS_CG, "bicgstab": rt.CGS_BICGSTAB, "cr": rt.CGS_CR, "minres": rt.CGS_MINRES,
                              "smoother": rt.CGS_SMOOTHER}[coarse_solver]
        self.dim, self.min_level, self.max_level = dim, min_level, max_level
        self.op, self.order, self.kappa = op, order, kappa
        self.smoother, self.omega, self.nu_pre, self.nu_post = smoother, omega, nu_pre, nu_post
        self.restrict_scale, self.cg_max_it, self.cg_rel_tol = restrict_scale, cg_max_it, cg_rel_tol
        self.mode = mode
        # initGlobals / initDomain / setupBuffers
        self.ctx = rt.Context(dim, min_level, max_level, block_cells=block_cells, num_blocks=num_blocks, rank=rank,
                              device=device, domain_lo=domain_lo, domain_hi=domain_hi, periodic=periodic)
        if stream is not None:
            self.ctx.set_stream(stream)
        ctx = self.ctx
        self.levels = list(range(min_level, max_level + 1))
        self.coeff_field = {}
        self.stencil = {}
        self._host = {}
        for lvl in self.levels:
            h = ctx.width(lvl)
            offsets, coeffs = laplace_stencil(dim, h, scaled=(op != "unscaled"), order=order)
            if op == "varcoeff27":
                assert dim == 3, "varcoeff27: 3D"
                offsets = varcoeff27_offsets()
                cf = ctx.field_create("LaplaceCoeff", lvl, 1, 27)
                self.coeff_field[lvl] = cf
                self.stencil[lvl] = ctx.stencil_field(lvl, offsets, cf)
                for comp, prog in varcoeff27_programs(h, kappa):  # InitLaplace@lvl, on the device
                    ctx.field_eval(cf, prog, comp)
            elif op == "varcoeff":
                cf = ctx.field_create("LaplaceCoeff", lvl, 1, len(offsets))
                self.coeff_field[lvl] = cf
                self.stencil[lvl] = ctx.stencil_field(lvl, offsets, cf)
            else:
                self.stencil[lvl] = ctx.stencil_const(lvl, offsets, coeffs)
            self._offsets = offsets
        desc = rt.CycleDesc()
        desc.smoother = {"rbgs": rt.SMOOTH_RBGS, "jacobi": rt.SMOOTH_JACOBI, "colored2": rt.SMOOTH_COLORED_2,
                         "colored3": rt.SMOOTH_COLORED_3}[smoother]
        desc.omega, desc.nu_pre, desc.nu_post = omega, nu_pre, nu_post
        desc.restrict_scale, desc.cg_max_it, desc.cg_rel_tol = restrict_scale, cg_max_it, cg_rel_tol
        desc.coarse_solver = self.coarse_solver
        desc.use_fas = 1 if use_fas else 0
        arr = (C.c_int * len(self.levels))(*[self.stencil[l] for l in self.levels])
        desc.stencils = C.cast(arr, C.POINTER(C.c_int))
        self._solver = ctx.L.examg_solver_create(ctx.h, C.byref(desc))
        if not self._solver:
            raise rt.ExaMGError("examg_solver_create: " + ctx.L.examg_last_error().decode())
        self.f_solution = {l: self._field(l, rt.F_SOLUTION) for l in self.levels}
        self.f_rhs = {l: self._field(l, rt.F_RHS) for l in self.levels}
        self.f_residual = {l: self._field(l, rt.F_RESIDUAL) for l in self.levels}
        self.f_cgtmp0 = self._field(min_level, rt.F_CGTMP0)
        self.f_cgtmp1 = self._field(min_level, rt.F_CGTMP1)
        # Field Solution< global, NodeWithComm, <expr> >@finest: boundary values come from the host arrays
        ctx.field_set_bc(self.f_solution[max_level], rt.BC_DIRICHLET_HOST)

    def _field(self, lvl, which):
        return rt.check(self.ctx.L.examg_solver_field(self._solver, lvl, which))

    def close(self):
        if getattr(self, "_solver", None):
            self.ctx.L.examg_solver_destroy(self._solver)
            self._solver = None
        if getattr(self, "ctx", None):
            self.ctx.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- host staging arrays (what initFields / the Application function fill) ----
    def dims(self, lvl):
        return self.ctx.dims(lvl)

    def node_positions(self, lvl):
        """vf_nodePosition_d = i*h (C/s/e/grid/ir/IR_VF_NodePosition.scala:109-111), global index."""
        n = self.dims(lvl)
        h = self.ctx.width(lvl)
        nb = self.ctx.cfg.num_blocks
        out = []
        r = self.ctx.cfg.rank
        for d in range(3):
            if d < self.dim:
                b = r % nb[d]
                r //= nb[d]
                g0 = b * (n[d] - 1)
                out.append((np.arange(n[d], dtype=np.float64) + g0) * h[d])
            else:
                out.append(np.zeros(1))
        return out

    def _staging(self, key, lvl, lead=None):
        if key not in self._host:
            shp = self.ctx.shape(lvl)
            if lead:
                shp = (lead,) + shp
            self._host[key] = np.zeros(shp, dtype=np.float64)
        return self._host[key]

    def solution(self, lvl):
        return self._staging(("u", lvl), lvl)

    def rhs(self, lvl):
        return self._staging(("f", lvl), lvl)

    def coeff(self, lvl):
        if self.op not in ("varcoeff", "varcoeff27"):
            return None
        return self._staging(("c", lvl), lvl, lead=len(self._offsets))

    def commit(self):
        """Upload every staged host array (set<name>() of the external-field interface)."""
        ctx = self.ctx
        for (kind, lvl), a in self._host.items():
            if kind == "u":
                f = self.f_solution[lvl]
                for slot in range(rt.check(ctx.L.examg_field_num_slots(ctx.h, f))):
                    ctx.upload(f, a, slot)  # apply bc to Solution<0>; apply bc to Solution<1>
            elif kind == "f":
                ctx.upload(self.f_rhs[lvl], a)
            else:
                ctx.upload(self.coeff_field[lvl], a)

    def sync_slots(self):
        self.commit()

    def fetch_solution(self, lvl=None):
        lvl = self.max_level if lvl is None else lvl
        return self.ctx.download(self.f_solution[lvl], lvl)

    def fetch_residual(self, lvl=None):
        lvl = self.max_level if lvl is None else lvl
        return self.ctx.download(self.f_residual[lvl], lvl)

    def fetch_rhs(self, lvl):
        return self.ctx.download(self.f_rhs[lvl], lvl)

    # ---- leveled functions of the L4 program -------------------------------------
    def UpResidual(self, lvl):
        """Residual = RHS - Laplace * Solution (…fromL4.exa4:123-125)."""
        self.ctx.residual(self.stencil[lvl], self.f_solution[lvl], self.f_rhs[lvl], self.f_residual[lvl])

    def ResNorm(self, lvl):
        """sqrt(sum Residual^2) (…fromL4.exa4:113-119)."""
        r = self.f_residual[lvl]
        return math.sqrt(self.ctx.dot(r, r))

    def update_residual_norm(self, lvl=None):
        """UpResidual@lvl + ResNorm@lvl with the reduction fused into the residual kernel."""
        lvl = self.max_level if lvl is None else lvl
        if self.mode == "library" and lvl == self.max_level:
            out = C.c_double()
            rt.check(self.ctx.L.examg_solver_residual_norm(self._solver, C.byref(out)))
            return out.value
        n2 = self.ctx.residual(self.stencil[lvl], self.f_solution[lvl], self.f_rhs[lvl], self.f_residual[lvl],
                               norm=True)
        return math.sqrt(n2)

    def Smoother(self, lvl, n):
        ctx = self.ctx
        A, u, f = self.stencil[lvl], self.f_solution[lvl], self.f_rhs[lvl]
        if self.smoother == "rbgs":
            for _ in range(n):  # color with { (i0+i1+i2)%2, ... }
                for color in (0, 1):
                    ctx.smooth_rb_color(A, u, f, self.omega, color)
                    ctx.apply_bc(u)
        else:
            for _ in range(n):
                # communicate ghost; loop; advance -- no `apply bc` in the Jacobi programs (Testing/SISC, Testing/Smoothers/Jac,
                # Testing/FMG: their FMG start-up relies on the boundary nodes keeping what SetFuncDir wrote)
                ctx.smooth_jacobi(A, u, f, self.omega, 1)

    def mgCycle(self, lvl):
        ctx = self.ctx
        if lvl == self.min_level:  # mgCycle@coarsest
            ctx.cg_coarse(self.stencil[lvl], self.f_solution[lvl], self.f_rhs[lvl], self.f_residual[lvl],
                          self.f_cgtmp0, self.f_cgtmp1, self.cg_max_it, self.cg_rel_tol, want_iters=False)
            return
        self.Smoother(lvl, self.nu_pre)
        self.UpResidual(lvl)
        ctx.apply_bc(self.f_residual[lvl])
        ctx.restrict(self.f_residual[lvl], self.f_rhs[lvl - 1], self.restrict_scale)
        ctx.set(self.f_solution[lvl - 1], 0.0)
        ctx.apply_bc(self.f_solution[lvl - 1])
        self.mgCycle(lvl - 1)
        ctx.exchange(self.f_solution[lvl - 1])
        ctx.prolong_add(self.f_solution[lvl - 1], self.f_solution[lvl])
        if self.smoother == "rbgs":  # Benchmark/Poisson3D/3D_FD_Poisson_fromL4.exa4:233-237
            ctx.apply_bc(self.f_solution[lvl])
        self.Smoother(lvl, self.nu_post)

    def vcycle(self, lvl=None):
        lvl = self.max_level if lvl is None else lvl
        if self.mode == "library":
            rt.check(self.ctx.L.examg_solver_vcycle(self._solver, lvl))
        else:
            self.mgCycle(lvl)

    def cycles(self, n):
        """n x mgCycle@finest, then residual + norm (one host synchronisation)."""
        out = C.c_double()
        if self.mode == "library":
            rt.check(self.ctx.L.examg_solver_cycles(self._solver, n, C.byref(out)))
            return out.value
        for _ in range(n):
            self.mgCycle(self.max_level)
        return self.update_residual_norm()

    def last_cg_iters(self):
        out = C.c_int()
        rt.check(self.ctx.L.examg_solver_last_cg_iters(self._solver, C.byref(out)))
        return out.value

    def Solve(self, max_it=100, target_reduction=1e-6, strict=False):
        """Solve@finest (…fromL4.exa4:121-150): returns the residual history."""
        init = self.update_residual_norm()
        hist = [init]
        res, it = init, 0
        while it < max_it and not (res < target_reduction * init if strict else res <= target_reduction * init):
            it += 1
            self.vcycle()
            res = self.update_residual_norm()
            hist.append(res)
        return hist
