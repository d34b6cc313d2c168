"""ctypes wrappers of the optical-flow part of the oracle (examg_oracle_optflow.c) -- TEST INFRASTRUCTURE ONLY.

Dense arrays are numpy float64 of shape (ny+2, nx+2): one ghost cell per side, x fastest.
"""
import ctypes as C

import numpy as np

from . import oracle as _o

U, V, RHS_U, RHS_V, RES_U, RES_V, CUU, CUV, CVV, IMG0, IMG1 = range(11)


class Config(C.Structure):
    _fields_ = [("F", C.c_int * 2), ("min_level", C.c_int), ("max_level", C.c_int), ("alpha", C.c_double),
                ("relax", C.c_double), ("nu_pre", C.c_int), ("nu_post", C.c_int), ("cg_max_it", C.c_int),
                ("cg_rel_tol", C.c_double)]


_ready = False


def lib():
    global _ready
    L = _o.lib()
    if _ready:
        return L
    dp = C.POINTER(C.c_double)
    d5 = C.c_double * 5
    L.ogf_neumann.argtypes = [C.c_int, C.c_int, dp]
    L.ogf_apply.argtypes = [C.c_int, C.c_int, d5, C.c_double] + [dp] * 7
    L.ogf_residual.argtypes = [C.c_int, C.c_int, d5, C.c_double] + [dp] * 9
    L.ogf_smooth_color.argtypes = [C.c_int, C.c_int, d5, C.c_double] + [dp] * 7 + [C.c_double, C.c_int]
    L.ogf_restrict.argtypes = [C.c_int, C.c_int, dp, dp]
    L.ogf_prolong_add.argtypes = [C.c_int, C.c_int, dp, dp]
    L.ogf_dot2.argtypes = [C.c_int, C.c_int, dp, dp, dp, dp]
    L.ogf_dot2.restype = C.c_double
    L.ogf_solver_create.argtypes = [C.POINTER(Config)]
    L.ogf_solver_create.restype = C.c_void_p
    L.ogf_solver_destroy.argtypes = [C.c_void_p]
    L.ogf_solver_dims.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    L.ogf_solver_field.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.ogf_solver_field.restype = dp
    L.ogf_solver_init_images.argtypes = [C.c_void_p]
    L.ogf_solver_init_fields.argtypes = [C.c_void_p]
    L.ogf_solver_update_residual_norm.argtypes = [C.c_void_p, C.c_int]
    L.ogf_solver_update_residual_norm.restype = C.c_double
    L.ogf_solver_vcycle.argtypes = [C.c_void_p, C.c_int]
    L.ogf_solver_last_cg_iters.argtypes = [C.c_void_p]
    L.ogf_solver_last_cg_iters.restype = C.c_int
    _ready = True
    return L


def _dp(a):
    if a is None:
        return None
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _nxy(a):
    return a.shape[1] - 2, a.shape[0] - 2


def laplace(h):
    """Stencil Laplace of the optical-flow programs, declaration order centre,-x,+x,-y,+y."""
    hx2, hy2 = h[0] * h[0], h[1] * h[1]
    return (C.c_double * 5)(2.0 / hx2 + 2.0 / hy2, -1.0 / hx2, -1.0 / hx2, -1.0 / hy2, -1.0 / hy2)


def neumann(a):
    nx, ny = _nxy(a)
    lib().ogf_neumann(nx, ny, _dp(a))
    return a


def apply(L5, a2, cuu, cuv, cvv, u, v):
    nx, ny = _nxy(u)
    ou, ov = np.zeros_like(u), np.zeros_like(v)
    lib().ogf_apply(nx, ny, L5, a2, _dp(cuu), _dp(cuv), _dp(cvv), _dp(u), _dp(v), _dp(ou), _dp(ov))
    return ou, ov


def residual(L5, a2, cuu, cuv, cvv, u, v, fu, fv):
    nx, ny = _nxy(u)
    ru, rv = np.zeros_like(u), np.zeros_like(v)
    lib().ogf_residual(nx, ny, L5, a2, _dp(cuu), _dp(cuv), _dp(cvv), _dp(u), _dp(v), _dp(fu), _dp(fv), _dp(ru), _dp(rv))
    return ru, rv


def smooth_color(L5, a2, cuu, cuv, cvv, u, v, fu, fv, relax, color):
    """in place on u, v"""
    nx, ny = _nxy(u)
    lib().ogf_smooth_color(nx, ny, L5, a2, _dp(cuu), _dp(cuv), _dp(cvv), _dp(u), _dp(v), _dp(fu), _dp(fv), relax, color)


def restrict(fine):
    nxf, nyf = _nxy(fine)
    coarse = np.zeros((nyf // 2 + 2, nxf // 2 + 2))
    lib().ogf_restrict(nxf // 2, nyf // 2, _dp(fine), _dp(coarse))
    return coarse


def prolong_add(coarse, fine):
    nxc, nyc = _nxy(coarse)
    lib().ogf_prolong_add(nxc, nyc, _dp(coarse), _dp(fine))
    return fine


def dot2(a, b, c=None, d=None):
    nx, ny = _nxy(a)
    return lib().ogf_dot2(nx, ny, _dp(a), _dp(b), _dp(c), _dp(d))


class Solver:
    """doSolve / mgCycle of the optical-flow programs on one global cell grid."""

    def __init__(self, F, min_level, max_level, alpha=1.0, relax=1.0, nu_pre=3, nu_post=3, cg_max_it=128,
                 cg_rel_tol=1e-3):
        cfg = Config()
        cfg.F[0], cfg.F[1] = F
        cfg.min_level, cfg.max_level = min_level, max_level
        cfg.alpha, cfg.relax = alpha, relax
        cfg.nu_pre, cfg.nu_post = nu_pre, nu_post
        cfg.cg_max_it, cfg.cg_rel_tol = cg_max_it, cg_rel_tol
        self.cfg = cfg
        self.min_level, self.max_level = min_level, max_level
        self._h = lib().ogf_solver_create(C.byref(cfg))

    def close(self):
        if self._h:
            lib().ogf_solver_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def dims(self, lvl):
        n = (C.c_int * 2)()
        lib().ogf_solver_dims(self._h, lvl, n)
        return n[0], n[1]

    def field(self, lvl, which):
        """numpy view (ny+2, nx+2) of one of the solver's arrays"""
        nx, ny = self.dims(self.max_level if which in (IMG0, IMG1) else lvl)
        p = lib().ogf_solver_field(self._h, lvl, which)
        return np.ctypeslib.as_array(p, shape=(ny + 2, nx + 2))

    def init_images(self):
        lib().ogf_solver_init_images(self._h)

    def init_fields(self):
        lib().ogf_solver_init_fields(self._h)

    def update_residual_norm(self, lvl=None):
        return lib().ogf_solver_update_residual_norm(self._h, self.max_level if lvl is None else lvl)

    def vcycle(self, lvl=None):
        lib().ogf_solver_vcycle(self._h, self.max_level if lvl is None else lvl)

    def last_cg_iters(self):
        return lib().ogf_solver_last_cg_iters(self._h)
