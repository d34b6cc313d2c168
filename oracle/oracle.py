"""ctypes loader for oracle/_ref/libexamg_oracle.so -- TEST INFRASTRUCTURE ONLY.

Thin numpy-facing wrappers around the C restatement in examg_oracle.c /
examg_oracle_solver.c (see examg_oracle.h for the reference file:line each
function follows).  Arrays are numpy float64, C-contiguous, shape (nz, ny, nx)
(x fastest), which is the oracle's dense layout.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libexamg_oracle.so")

OG_MAX_ST = 27
SMOOTH_JACOBI, SMOOTH_RBGS = 0, 1
OP_LAPLACE_SCALED, OP_LAPLACE_UNSCALED, OP_VARCOEFF, OP_VARCOEFF27 = 0, 1, 2, 3
ORDER_BENCH, ORDER_TESTING = 0, 1


class Stencil(C.Structure):
    _fields_ = [("nst", C.c_int), ("off", (C.c_int * 3) * OG_MAX_ST), ("coef", C.c_double * OG_MAX_ST)]

    @classmethod
    def make(cls, offsets, coeffs=None):
        s = cls()
        s.nst = len(offsets)
        for k, o in enumerate(offsets):
            o = list(o) + [0] * (3 - len(o))
            for d in range(3):
                s.off[k][d] = int(o[d])
            s.coef[k] = float(coeffs[k]) if coeffs is not None else 0.0
        return s

    def offsets(self):
        return [tuple(self.off[k][d] for d in range(3)) for k in range(self.nst)]

    def coeffs(self):
        return [self.coef[k] for k in range(self.nst)]


class Config(C.Structure):
    _fields_ = [("dim", C.c_int), ("F", C.c_int * 3), ("min_level", C.c_int), ("max_level", C.c_int),
                ("op", C.c_int), ("order", C.c_int), ("kappa", C.c_double), ("smoother", C.c_int),
                ("omega", C.c_double), ("nu_pre", C.c_int), ("nu_post", C.c_int),
                ("restrict_scale", C.c_double), ("cg_max_it", C.c_int), ("cg_rel_tol", C.c_double),
                ("domain_hi", C.c_double * 3), ("use_fas", C.c_int)]


def build(force=False):
    """Compile the oracle (gcc).  Building the checker is not using it."""
    srcs = [os.path.join(_HERE, f) for f in ("examg_oracle.c", "examg_oracle_solver.c", "examg_oracle.h", "Makefile")]
    if not force and os.path.exists(_SO) and all(os.path.getmtime(_SO) >= os.path.getmtime(s) for s in srcs):
        return _SO
    subprocess.check_call(["make", "-C", _HERE, "-s"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_SO):
        build()
    L = C.CDLL(_SO)
    i3 = C.POINTER(C.c_int)
    dp = C.POINTER(C.c_double)
    stp = C.POINTER(Stencil)
    L.og_apply.argtypes = [i3, stp, dp, dp, dp]
    L.og_residual.argtypes = [i3, stp, dp, dp, dp, dp]
    L.og_smooth_jacobi.argtypes = [i3, stp, dp, C.c_double, dp, dp, dp]
    L.og_smooth_rb.argtypes = [i3, stp, dp, C.c_double, dp, dp, C.c_int]
    L.og_dot.argtypes = [i3, dp, dp]
    L.og_dot.restype = C.c_double
    L.og_max_abs_diff.argtypes = [i3, dp, dp]
    L.og_max_abs_diff.restype = C.c_double
    L.og_cr.argtypes = [i3, stp, dp, dp, dp, dp, C.c_int, C.c_double]
    L.og_cr.restype = C.c_int
    L.og_minres.argtypes = [i3, stp, dp, dp, dp, dp, C.c_int, C.c_double]
    L.og_minres.restype = C.c_int
    L.og_bicgstab.argtypes = [i3, stp, dp, dp, dp, dp, C.c_int, C.c_double]
    L.og_bicgstab.restype = C.c_int
    L.og_smoother_solve.argtypes = [i3, stp, dp, dp, dp, dp, C.c_int, C.c_double, C.c_int, C.c_double]
    L.og_smoother_solve.restype = C.c_int
    L.og_smooth_color.argtypes = [i3, stp, dp, C.c_double, dp, dp, C.c_int, C.c_int, C.c_int, C.c_int]
    L.og_smooth_color.restype = None
    L.og_cell_apply_bc.argtypes = [C.c_int, i3, dp, C.c_int, C.POINTER(dp)]
    L.og_cell_apply_bc.restype = None
    L.og_cell_restrict.argtypes = [C.c_int, i3, i3, dp, dp]
    L.og_cell_restrict.restype = None
    L.og_cell_prolong_add.argtypes = [C.c_int, i3, i3, dp, dp]
    L.og_cell_prolong_add.restype = None
    L.og_set_periodic.argtypes = [i3]
    L.og_set_periodic.restype = None
    L.og_glibc_rand_fill.argtypes = [dp, C.c_longlong]
    L.og_glibc_rand_fill.restype = None
    L.og_restrict.argtypes = [C.c_int, i3, i3, C.c_double, dp, dp]
    L.og_prolong_add.argtypes = [C.c_int, i3, i3, dp, dp]
    L.og_loop_eval.argtypes = [i3, dp, dp, C.c_int, C.POINTER(dp), dp, i3, dp, C.c_int]
    L.og_loop_eval.restype = None
    L.og_set_interior.argtypes = [i3, dp, C.c_double]
    L.og_copy_interior.argtypes = [i3, dp, dp]
    L.og_axpy.argtypes = [i3, C.c_double, dp, dp]
    L.og_xpay.argtypes = [i3, dp, C.c_double, dp]
    L.og_solver_create.argtypes = [C.POINTER(Config)]
    L.og_solver_create.restype = C.c_void_p
    L.og_solver_destroy.argtypes = [C.c_void_p]
    for f in ("og_solver_solution", "og_solver_rhs", "og_solver_residual", "og_solver_coeff"):
        getattr(L, f).argtypes = [C.c_void_p, C.c_int]
        getattr(L, f).restype = dp
    L.og_solver_dims.argtypes = [C.c_void_p, C.c_int, i3]
    L.og_solver_stencil.argtypes = [C.c_void_p, C.c_int, stp]
    L.og_solver_sync_slots.argtypes = [C.c_void_p]
    L.og_solver_update_residual_norm.argtypes = [C.c_void_p, C.c_int]
    L.og_solver_update_residual_norm.restype = C.c_double
    L.og_solver_vcycle.argtypes = [C.c_void_p, C.c_int]
    L.og_solver_last_cg_iters.argtypes = [C.c_void_p]
    L.og_solver_last_cg_iters.restype = C.c_int
    L.og_print_reduced_prec.argtypes = [C.c_double, C.c_char_p, C.c_size_t]
    L.og_set_num_threads.argtypes = [C.c_int]
    L.og_set_num_threads.restype = None
    L.og_get_max_threads.restype = C.c_int
    L.og_count_threads.restype = C.c_int
    _lib = L
    return L


def set_num_threads(n):
    """omp_numThreads of the timed CPU baseline; returns the team size an OpenMP loop really gets."""
    lib().og_set_num_threads(int(n))
    return lib().og_count_threads()


def _dp(a):
    if a is None:
        return None
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _n(a):
    """(nz, ny, nx) or (ny, nx) numpy shape -> int[3] = {nx, ny, nz}."""
    shp = a.shape if a.ndim == 3 else (1,) + tuple(a.shape)
    return (C.c_int * 3)(shp[2], shp[1], shp[0])


def _sf(sf):
    return _dp(sf) if sf is not None else None


def apply(st, u, sf=None):
    out = np.zeros_like(u)
    lib().og_apply(_n(u), C.byref(st), _sf(sf), _dp(u), _dp(out))
    return out


def residual(st, u, rhs, sf=None):
    res = np.zeros_like(u)
    lib().og_residual(_n(u), C.byref(st), _sf(sf), _dp(u), _dp(rhs), _dp(res))
    return res


def smooth_jacobi(st, omega, u, rhs, sf=None):
    un = u.copy()
    lib().og_smooth_jacobi(_n(u), C.byref(st), _sf(sf), omega, _dp(u), _dp(rhs), _dp(un))
    return un


def smooth_rb(st, omega, u, rhs, color, sf=None):
    """In place on u."""
    lib().og_smooth_rb(_n(u), C.byref(st), _sf(sf), omega, _dp(u), _dp(rhs), int(color))
    return u


def smooth_colored(st, omega, u, rhs, ncol, dim, sf=None):
    """One n^dim-way coloured Gauss-Seidel sweep in place on u: colours with the last dimension slowest."""
    for cz in range(ncol if dim == 3 else 1):
        for cy in range(ncol):
            for cx in range(ncol):
                lib().og_smooth_color(_n(u), C.byref(st), _sf(sf), omega, _dp(u), _dp(rhs), ncol, cx, cy, cz)
    return u


def smooth_color(st, omega, u, rhs, ncol, cx, cy, cz, sf=None):
    """One colour (i % ncol == cx, j % ncol == cy, k % ncol == cz by array index) of the n^dim-way colouring, in place."""
    lib().og_smooth_color(_n(u), C.byref(st), _sf(sf), omega, _dp(u), _dp(rhs), ncol, cx, cy, cz)
    return u


def smoother_solve(st, u, rhs, kind, omega, max_it, rel_tol, sf=None):
    """solver_cgs = "Smoother": in place on u; kind 0 Jacobi, 1 red-black GS; returns (outer steps, residual field)."""
    res = np.zeros_like(u)
    it = lib().og_smoother_solve(_n(u), C.byref(st), _sf(sf), _dp(u), _dp(rhs), _dp(res), int(kind), float(omega),
                                 int(max_it), float(rel_tol))
    return it, res


def conjugate_residual(st, u, rhs, max_it, rel_tol, sf=None):
    """In place on u; returns (steps, residual field)."""
    res = np.zeros_like(u)
    it = lib().og_cr(_n(u), C.byref(st), _sf(sf), _dp(u), _dp(rhs), _dp(res), int(max_it), float(rel_tol))
    return it, res


def minres(st, u, rhs, max_it, rel_tol, sf=None):
    """In place on u; returns (steps, initial residual field)."""
    res = np.zeros_like(u)
    it = lib().og_minres(_n(u), C.byref(st), _sf(sf), _dp(u), _dp(rhs), _dp(res), int(max_it), float(rel_tol))
    return it, res


def bicgstab(st, u, rhs, max_it, rel_tol, sf=None):
    """In place on u; returns (steps, residual field)."""
    res = np.zeros_like(u)
    it = lib().og_bicgstab(_n(u), C.byref(st), _sf(sf), _dp(u), _dp(rhs), _dp(res), int(max_it), float(rel_tol))
    return it, res


def dot(a, b):
    return lib().og_dot(_n(a), _dp(a), _dp(b))


# ---- cell-centred fields: padded arrays (cells + 2 per dimension); the node loops above run over exactly the cells ----
BC_CELL_DIRICHLET, BC_CELL_NEUMANN = 1, 2


def cell_apply_bc(dim, u, kind, face_values=None):
    """In place.  face_values: list of 6 arrays (face 2*d + side, the face's cells in lexicographic order) or None = 0."""
    arr = None
    keep = []
    if face_values is not None:
        arr = (C.POINTER(C.c_double) * 6)()
        for f in range(6):
            v = face_values[f] if f < len(face_values) else None
            if v is not None:
                v = np.ascontiguousarray(v, dtype=np.float64)
                keep.append(v)
                arr[f] = _dp(v)
    lib().og_cell_apply_bc(dim, _n(u), _dp(u), kind, arr)
    return u


def cell_restrict(dim, rf, coarse_shape):
    rc = np.zeros(coarse_shape, dtype=np.float64)
    lib().og_cell_restrict(dim, _n(rf), _n(rc), _dp(rf), _dp(rc))
    return rc


def cell_prolong_add(dim, uc, uf):
    lib().og_cell_prolong_add(dim, _n(uf), _n(uc), _dp(uc), _dp(uf))
    return uf


def set_periodic(per=(0, 0, 0)):
    """domain_rect_periodic_{x,y,z}: process-wide switch of the oracle's loops (examg_oracle.c)."""
    per = list(per) + [0] * (3 - len(per))
    lib().og_set_periodic((C.c_int * 3)(*[int(bool(v)) for v in per]))


def glibc_rand(n):
    """n values of (double)std::rand()/RAND_MAX from a freshly started process (Testing/Opts/base.exa4:167-169)."""
    out = np.empty(int(n), dtype=np.float64)
    lib().og_glibc_rand_fill(_dp(out), int(n))
    return out


def max_abs_diff(a, b):
    return lib().og_max_abs_diff(_n(a), _dp(a), _dp(b))


def restrict(dim, rf, coarse_shape, scale=1.0):
    rc = np.zeros(coarse_shape, dtype=np.float64)
    lib().og_restrict(dim, _n(rf), _n(rc), scale, _dp(rf), _dp(rc))
    return rc


def prolong_add(dim, uc, uf):
    """In place on uf."""
    lib().og_prolong_add(dim, _n(uf), _n(uc), _dp(uc), _dp(uf))
    return uf


def copy_interior(src, dst):
    lib().og_copy_interior(_n(src), _dp(src), _dp(dst))
    return dst


def loop_eval(dst, srcs, prog, h, x0=(0.0, 0.0, 0.0)):
    """dst = program(node position, srcs at the same node) on the loop points; prog = [(op, value), ...] with the
    EXAMG_OP_* codes of include/examg.h (15 = source field index)."""
    n = len(prog)
    ops = (C.c_int * n)(*[int(o) for o, _ in prog])
    vals = (C.c_double * n)(*[float(v) for _, v in prog])
    hh = (C.c_double * 3)(*(list(h) + [1.0] * (3 - len(h))))
    xx = (C.c_double * 3)(*(list(x0) + [0.0] * (3 - len(x0))))
    ptrs = (C.POINTER(C.c_double) * max(1, len(srcs)))(*[_dp(a) for a in srcs])
    lib().og_loop_eval(_n(dst), hh, xx, len(srcs), ptrs, _dp(dst), ops, vals, n)
    return dst


def axpy(alpha, x, y):
    lib().og_axpy(_n(x), alpha, _dp(x), _dp(y))
    return y


def xpay(x, beta, y):
    lib().og_xpay(_n(x), _dp(x), beta, _dp(y))
    return y


def print_reduced_prec(v):
    buf = C.create_string_buffer(64)
    lib().og_print_reduced_prec(v, buf, 64)
    return buf.value.decode()


class Solver:
    """Handle on og_solver; numpy views alias the oracle's own arrays."""

    def __init__(self, dim, F, min_level, max_level, op=OP_LAPLACE_SCALED, order=ORDER_BENCH, kappa=0.0,
                 smoother=SMOOTH_RBGS, omega=0.8, nu_pre=3, nu_post=3, restrict_scale=1.0,
                 cg_max_it=128, cg_rel_tol=1e-3, domain_hi=(1.0, 1.0, 1.0), use_fas=False):
        cfg = Config()
        cfg.use_fas = 1 if use_fas else 0
        for d in range(3):
            cfg.domain_hi[d] = domain_hi[d] if d < len(domain_hi) else 1.0
        cfg.dim = dim
        F = list(F) + [1] * (3 - len(F))
        for d in range(3):
            cfg.F[d] = F[d]
        cfg.min_level, cfg.max_level = min_level, max_level
        cfg.op, cfg.order, cfg.kappa = op, order, kappa
        cfg.smoother, cfg.omega, cfg.nu_pre, cfg.nu_post = smoother, omega, nu_pre, nu_post
        cfg.restrict_scale, cfg.cg_max_it, cfg.cg_rel_tol = restrict_scale, cg_max_it, cg_rel_tol
        self.cfg = cfg
        self.dim, self.min_level, self.max_level = dim, min_level, max_level
        self._h = lib().og_solver_create(C.byref(cfg))

    def close(self):
        if self._h:
            lib().og_solver_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def dims(self, lvl):
        n = (C.c_int * 3)()
        lib().og_solver_dims(self._h, lvl, n)
        return n[0], n[1], n[2]

    def _view(self, ptr, lvl, lead=None):
        nx, ny, nz = self.dims(lvl)
        shape = (nz, ny, nx) if lead is None else (lead, nz, ny, nx)
        return np.ctypeslib.as_array(ptr, shape=shape)

    def solution(self, lvl):
        return self._view(lib().og_solver_solution(self._h, lvl), lvl)

    def rhs(self, lvl):
        return self._view(lib().og_solver_rhs(self._h, lvl), lvl)

    def residual(self, lvl):
        return self._view(lib().og_solver_residual(self._h, lvl), lvl)

    def stencil(self, lvl):
        st = Stencil()
        lib().og_solver_stencil(self._h, lvl, C.byref(st))
        return st

    def coeff(self, lvl):
        p = lib().og_solver_coeff(self._h, lvl)
        if not p:
            return None
        return self._view(p, lvl, lead=self.stencil(lvl).nst)

    def node_positions(self, lvl):
        """vf_nodePosition_{x,y,z} = i*h (C/s/e/grid/ir/IR_VF_NodePosition.scala:109-111)."""
        nx, ny, nz = self.dims(lvl)
        F = self.cfg.F
        out = []
        for d, n in enumerate((nx, ny, nz)):
            hi = self.cfg.domain_hi[d] if self.cfg.domain_hi[d] > 0.0 else 1.0
            out.append(np.arange(n, dtype=np.float64) * (hi / float(n - 1)) if n > 1 else np.zeros(1))
        return out

    def sync_slots(self):
        lib().og_solver_sync_slots(self._h)

    def update_residual_norm(self, lvl=None):
        return lib().og_solver_update_residual_norm(self._h, self.max_level if lvl is None else lvl)

    def vcycle(self, lvl=None):
        lib().og_solver_vcycle(self._h, self.max_level if lvl is None else lvl)

    def last_cg_iters(self):
        return lib().og_solver_last_cg_iters(self._h)
