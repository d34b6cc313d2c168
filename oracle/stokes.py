"""ctypes wrappers of the staggered-grid Stokes part of the oracle (examg_oracle_stokes.c) -- TEST INFRASTRUCTURE ONLY.

Fields are numpy float64 views of shape (n+3,) * dim, entry (i, j[, k]) at [[k+1,] j+1, i+1] (x fastest): velocity component
d lives on the faces 0..n in dimension d and on the cells 0..n-1 in the others, p on the cells; one ghost layer around each.
Field ids: VEL + d, RHS + d, RES + d (d = 0, 1, 2), P, RHS_P, RES_P.
"""
import ctypes as C

import numpy as np

from . import oracle as _o

VEL, RHS, RES, P, RHS_P, RES_P = 0, 3, 6, 9, 10, 11
U, V, W = 0, 1, 2


class Config(C.Structure):
    _fields_ = [("dim", C.c_int), ("min_level", C.c_int), ("max_level", C.c_int), ("nu_pre", C.c_int), ("nu_post", C.c_int),
                ("nu_coarse", C.c_int), ("top_vel", C.c_double)]


_ready = False


def lib():
    global _ready
    L = _o.lib()
    if _ready:
        return L
    dp = C.POINTER(C.c_double)
    L.ogs_solver_create.argtypes = [C.POINTER(Config)]
    L.ogs_solver_create.restype = C.c_void_p
    L.ogs_solver_destroy.argtypes = [C.c_void_p]
    L.ogs_solver_field.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.ogs_solver_field.restype = dp
    L.ogs_solver_cells.argtypes = [C.c_void_p, C.c_int]
    L.ogs_solver_cells.restype = C.c_int
    for name in ("ogs_solver_apply_bc", "ogs_solver_residual", "ogs_solver_restrict", "ogs_solver_correct", "ogs_solver_vcycle"):
        getattr(L, name).argtypes = [C.c_void_p, C.c_int]
        getattr(L, name).restype = None
    L.ogs_solver_smooth.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.ogs_solver_smooth.restype = None
    L.ogs_smooth_color.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
    L.ogs_smooth_color.restype = None
    L.ogs_solver_set_solution.argtypes = [C.c_void_p, C.c_int, C.c_double]
    L.ogs_solver_set_solution.restype = None
    L.ogs_solver_norms.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double * 4)]
    L.ogs_solver_norms.restype = None
    _ready = True
    return L


class Solver:
    """Testing/Application/ExaStokes_2D (levels 2..8, lid velocity 0.02) / ExaStokes_3D (levels 1..6, 0.2): V(3,3), 10
    smoother steps on the coarsest level."""

    def __init__(self, min_level=None, max_level=None, nu_pre=3, nu_post=3, nu_coarse=10, top_vel=None, dim=2):
        min_level = (2 if dim == 2 else 1) if min_level is None else min_level
        max_level = (8 if dim == 2 else 6) if max_level is None else max_level
        top_vel = (0.02 if dim == 2 else 0.2) if top_vel is None else top_vel
        cfg = Config(dim, min_level, max_level, nu_pre, nu_post, nu_coarse, top_vel)
        self.dim, self.min_level, self.max_level = dim, min_level, max_level
        self._h = lib().ogs_solver_create(C.byref(cfg))

    def close(self):
        if self._h:
            lib().ogs_solver_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def cells(self, lvl):
        return lib().ogs_solver_cells(self._h, lvl)

    def field(self, lvl, which):
        n = self.cells(lvl) + 3
        return np.ctypeslib.as_array(lib().ogs_solver_field(self._h, lvl, which), shape=(n,) * self.dim)

    def field_ids(self):
        """all fields of this dimensionality"""
        return [k + d for k in (VEL, RHS, RES) for d in range(self.dim)] + [P, RHS_P, RES_P]

    def unknown_ids(self):
        return [VEL + d for d in range(self.dim)] + [P]

    def apply_bc(self, lvl):
        lib().ogs_solver_apply_bc(self._h, lvl)

    def smooth(self, lvl, nsweeps=1):
        lib().ogs_solver_smooth(self._h, lvl, nsweeps)

    def smooth_color(self, lvl, c0, c1, c2=0):
        lib().ogs_smooth_color(self._h, lvl, c0, c1, c2)

    def residual(self, lvl):
        lib().ogs_solver_residual(self._h, lvl)

    def norms(self, lvl):
        out = (C.c_double * 4)()
        lib().ogs_solver_norms(self._h, lvl, C.byref(out))
        return tuple(out[k] for k in range(self.dim + 1))

    def restrict(self, lvl):
        lib().ogs_solver_restrict(self._h, lvl)

    def set_solution(self, lvl, value):
        lib().ogs_solver_set_solution(self._h, lvl, value)

    def correct(self, lvl):
        lib().ogs_solver_correct(self._h, lvl)

    def vcycle(self, lvl=None):
        lib().ogs_solver_vcycle(self._h, self.max_level if lvl is None else lvl)


def solve(s, printer, max_it=10, tol=1e-6):
    """Application + Solve@finest of ExaStokes_2D.exa4: the printed tokens."""
    L = s.max_level
    out = []
    s.apply_bc(L)                       # InitFields@finest

    def norm_max():
        r = s.norms(L)
        out.extend(printer(v) for v in r)
        return max(r)

    s.residual(L)
    init = norm_max()
    out.append(printer(init))
    cur, it = init, 0
    while not (cur <= tol or it >= max_it):
        s.vcycle(L)
        s.residual(L)
        cur = norm_max()
        it += 1
    return out
