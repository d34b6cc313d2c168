"""ctypes wrapper of the 3D optical-flow part of the oracle (examg_oracle_optflow3.c) -- TEST INFRASTRUCTURE ONLY.

Dense arrays are numpy float64 of shape (nz+2, ny+2, nx+2): one ghost cell per side, x fastest.
"""
import ctypes as C

import numpy as np

from . import oracle as _o

FLOW, RHS, RES, G, IMG0, IMG1 = 0, 3, 6, 9, 15, 16
NFIELDS = 15


class Config(C.Structure):
    _fields_ = [("F", C.c_int * 3), ("min_level", C.c_int), ("max_level", C.c_int), ("alpha", C.c_double),
                ("nu_pre", C.c_int), ("nu_post", C.c_int), ("cg_max_it", C.c_int), ("cg_rel_tol", C.c_double)]


_ready = False


def lib():
    global _ready
    L = _o.lib()
    if _ready:
        return L
    L.og3_solver_create.argtypes = [C.POINTER(Config)]
    L.og3_solver_create.restype = C.c_void_p
    L.og3_solver_destroy.argtypes = [C.c_void_p]
    L.og3_solver_dims.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    L.og3_solver_field.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.og3_solver_field.restype = C.POINTER(C.c_double)
    L.og3_solver_init_images.argtypes = [C.c_void_p]
    L.og3_solver_init_fields.argtypes = [C.c_void_p]
    L.og3_solver_update_residual_norm.argtypes = [C.c_void_p, C.c_int]
    L.og3_solver_update_residual_norm.restype = C.c_double
    L.og3_solver_smooth_color.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.og3_solver_smooth.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.og3_solver_restrict.argtypes = [C.c_void_p, C.c_int]
    L.og3_solver_correct.argtypes = [C.c_void_p, C.c_int]
    L.og3_solver_vcycle.argtypes = [C.c_void_p, C.c_int]
    L.og3_solver_last_cg_iters.argtypes = [C.c_void_p]
    L.og3_solver_last_cg_iters.restype = C.c_int
    _ready = True
    return L


class Solver:
    """doSolve / mgCycle of Testing/Application/OpticalFlow3D.exa4 on one global cell grid."""

    def __init__(self, F=(1, 1, 1), min_level=2, max_level=6, alpha=1.0, nu_pre=3, nu_post=3, cg_max_it=128,
                 cg_rel_tol=1e-3):
        cfg = Config()
        cfg.F[0], cfg.F[1], cfg.F[2] = F
        cfg.min_level, cfg.max_level = min_level, max_level
        cfg.alpha = alpha
        cfg.nu_pre, cfg.nu_post = nu_pre, nu_post
        cfg.cg_max_it, cfg.cg_rel_tol = cg_max_it, cg_rel_tol
        self.cfg = cfg
        self.min_level, self.max_level = min_level, max_level
        self._h = lib().og3_solver_create(C.byref(cfg))

    def close(self):
        if self._h:
            lib().og3_solver_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def dims(self, lvl):
        n = (C.c_int * 3)()
        lib().og3_solver_dims(self._h, lvl, n)
        return n[0], n[1], n[2]

    def field(self, lvl, which):
        """numpy view (nz+2, ny+2, nx+2) of one of the solver's arrays"""
        nx, ny, nz = self.dims(self.max_level if which in (IMG0, IMG1) else lvl)
        p = lib().og3_solver_field(self._h, lvl, which)
        return np.ctypeslib.as_array(p, shape=(nz + 2, ny + 2, nx + 2))

    def init_images(self):
        lib().og3_solver_init_images(self._h)

    def init_fields(self):
        lib().og3_solver_init_fields(self._h)

    def update_residual_norm(self, lvl=None):
        return lib().og3_solver_update_residual_norm(self._h, self.max_level if lvl is None else lvl)

    def smooth_color(self, lvl, color):
        lib().og3_solver_smooth_color(self._h, lvl, color)

    def smooth(self, lvl, n):
        lib().og3_solver_smooth(self._h, lvl, n)

    def restrict(self, lvl):
        lib().og3_solver_restrict(self._h, lvl)

    def correct(self, lvl):
        lib().og3_solver_correct(self._h, lvl)

    def vcycle(self, lvl=None):
        lib().og3_solver_vcycle(self._h, self.max_level if lvl is None else lvl)

    def last_cg_iters(self):
        return lib().og3_solver_last_cg_iters(self._h)

    def solve(self, printer, max_it=128, rel_tol=1e-5):
        """doSolve@finest: prints through `printer` what the program prints through printVal"""
        init = cur = self.update_residual_norm()
        printer(init)
        it = 0
        while not (it >= max_it or cur <= rel_tol * init):
            it += 1
            self.vcycle()
            cur = self.update_residual_norm()
            printer(cur)
        return it
