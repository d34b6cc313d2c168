"""The staggered-grid Stokes programs (Testing/Application/ExaStokes_2D.exa4 / ExaStokes_3D.exa4: lid-driven cavity) as host
code over libexamg: `Application` / `InitFields` / `Solve@finest` / `VCycle@all`.

Velocity component d lives on the faces in dimension d, p in the cells; the smoother solves the (2 dim + 1)^2 system of a
cell's faces and its pressure, 3^dim colours per sweep.  All arithmetic runs in the CUDA library; there is no CPU path here.
A field of level l is returned as a numpy array of shape (n+3,) * dim, entry (i, j[, k]) at [[k+1,] j+1, i+1], n = cells per
dimension.  Field ids: VEL + d, RHS + d, RES + d (d = 0, 1, 2), P, RHS_P, RES_P.
"""
import ctypes as C

import numpy as np

from . import runtime as rt

VEL, RHS, RES, P, RHS_P, RES_P = 0, 3, 6, 9, 10, 11


class StokesProgram:
    def __init__(self, min_level=None, max_level=None, nu_pre=3, nu_post=3, nu_coarse=10, top_vel=None, block_cells=1,
                 device=0, stream=None, dim=2):
        min_level = (2 if dim == 2 else 1) if min_level is None else min_level
        max_level = (8 if dim == 2 else 6) if max_level is None else max_level
        top_vel = (0.02 if dim == 2 else 0.2) if top_vel is None else top_vel
        self.dim = dim
        self.ctx = rt.Context(dim, min_level, max_level, block_cells=(block_cells,) * dim + (1,) * (3 - dim), device=device)
        if stream is not None:
            self.ctx.set_stream(stream)
        self.min_level, self.max_level = min_level, max_level
        d = rt.StokesDesc(nu_pre, nu_post, nu_coarse, top_vel)
        self._h = self.ctx.L.examg_stokes_create(self.ctx.h, C.byref(d))
        if not self._h:
            raise rt.ExaMGError("examg_stokes_create: " + self.ctx.L.examg_last_error().decode())

    def close(self):
        if getattr(self, "_h", None):
            self.ctx.L.examg_stokes_destroy(self._h)
            self._h = None
        if getattr(self, "ctx", None):
            self.ctx.close()
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def cells(self, lvl):
        return rt.check(self.ctx.L.examg_stokes_cells(self._h, lvl))

    def field(self, lvl, which):
        n = self.cells(lvl) + 3
        out = np.empty((n,) * self.dim, dtype=np.float64)
        rt.check(self.ctx.L.examg_stokes_download(self._h, lvl, which, out.ctypes.data_as(C.c_void_p)))
        return out

    def set_field(self, lvl, which, arr):
        n = self.cells(lvl) + 3
        a = np.ascontiguousarray(arr, dtype=np.float64)
        assert a.shape == (n,) * self.dim
        rt.check(self.ctx.L.examg_stokes_upload(self._h, lvl, which, a.ctypes.data_as(C.c_void_p)))

    def apply_bc(self, lvl):
        rt.check(self.ctx.L.examg_stokes_apply_bc(self._h, lvl))

    def smooth(self, lvl, nsweeps=1):
        rt.check(self.ctx.L.examg_stokes_smooth(self._h, lvl, nsweeps))

    def smooth_color(self, lvl, c0, c1, c2=0):
        rt.check(self.ctx.L.examg_stokes_smooth_color(self._h, lvl, c0, c1, c2))

    def residual(self, lvl):
        rt.check(self.ctx.L.examg_stokes_residual(self._h, lvl, None))

    def norms(self, lvl):
        """UpdateResidual@lvl + NormResidual_{u,v[,w],p}"""
        out = (C.c_double * 4)()
        rt.check(self.ctx.L.examg_stokes_residual(self._h, lvl, out))
        return tuple(out[k] for k in range(self.dim + 1))

    def restrict(self, lvl):
        rt.check(self.ctx.L.examg_stokes_restrict(self._h, lvl))

    def set_solution(self, lvl, value):
        rt.check(self.ctx.L.examg_stokes_set_solution(self._h, lvl, value))

    def correct(self, lvl):
        rt.check(self.ctx.L.examg_stokes_correct(self._h, lvl))

    def vcycle(self, lvl=None):
        rt.check(self.ctx.L.examg_stokes_vcycle(self._h, self.max_level if lvl is None else lvl))

    def solve(self, printer, max_it=10, tol=1e-6):
        """Application + Solve@finest: the printed tokens (dim + 1 residual norms before the first and after every cycle)"""
        L = self.max_level
        out = []
        self.apply_bc(L)  # InitFields@finest

        def norm_max():
            r = self.norms(L)
            out.extend(printer(v) for v in r)
            return max(r)

        init = norm_max()
        out.append(printer(init))
        cur, it = init, 0
        while not (cur <= tol or it >= max_it):
            self.vcycle(L)
            cur = norm_max()
            it += 1
        return out
