"""The 3D optical-flow program (Testing/Application/OpticalFlow3D.exa4) as host code over libexamg:
`Application` / `InitFields` / `doSolve@finest` / `mgCycle@all`.

A Vec3 unknown `flow` on cell centres, combinedOp = alpha**2 * Laplace + gradContribution, red-black sweeps with the
3 x 3 diagonal block inverted per cell, cell restriction / piecewise-constant correction, CG on the coarsest level.  All
arithmetic runs in the CUDA library; there is no CPU path here.  An array of level l is a numpy array of shape
(nz+2, ny+2, nx+2) (one ghost cell per side, x fastest).  Array ids: FLOW + c, RHS + c, RES + c (c = 0, 1, 2),
G + 0..5 (xx, xy, xz, yy, yz, zz), IMG0, IMG1 (finest level).
"""
import ctypes as C

import numpy as np

from . import runtime as rt

FLOW, RHS, RES, G, IMG0, IMG1 = 0, 3, 6, 9, 15, 16


class OpticalFlow3DProgram:
    def __init__(self, min_level=2, max_level=6, alpha=1.0, nu_pre=3, nu_post=3, cg_max_it=128, cg_rel_tol=1e-3,
                 block_cells=(1, 1, 1), device=0, stream=None):
        self.ctx = rt.Context(3, min_level, max_level, block_cells=tuple(block_cells), device=device)
        if stream is not None:
            self.ctx.set_stream(stream)
        self.min_level, self.max_level = min_level, max_level
        d = rt.OptFlow3Desc(alpha, nu_pre, nu_post, cg_max_it, cg_rel_tol)
        self._h = self.ctx.L.examg_optflow3_create(self.ctx.h, C.byref(d))
        if not self._h:
            raise rt.ExaMGError("examg_optflow3_create: " + self.ctx.L.examg_last_error().decode())

    def close(self):
        if getattr(self, "_h", None):
            self.ctx.L.examg_optflow3_destroy(self._h)
            self._h = None
        if getattr(self, "ctx", None):
            self.ctx.close()
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def cells(self, lvl=-1):
        n = (C.c_int * 3)()
        rt.check(self.ctx.L.examg_optflow3_cells(self._h, lvl, n))
        return n[0], n[1], n[2]

    def _shape(self, lvl, which):
        nx, ny, nz = self.cells(-1 if which in (IMG0, IMG1) else lvl)
        return (nz + 2, ny + 2, nx + 2)

    def field(self, lvl, which):
        out = np.empty(self._shape(lvl, which), dtype=np.float64)
        rt.check(self.ctx.L.examg_optflow3_download(self._h, lvl, which, out.ctypes.data_as(C.c_void_p)))
        return out

    def set_field(self, lvl, which, arr):
        a = np.ascontiguousarray(arr, dtype=np.float64)
        assert a.shape == self._shape(lvl, which)
        rt.check(self.ctx.L.examg_optflow3_upload(self._h, lvl, which, a.ctypes.data_as(C.c_void_p)))

    def cell_centres(self):
        """vf_cellCenter_{x,y,z} of the finest level as broadcastable arrays over the inner cells [k, j, i]"""
        nx, ny, nz = self.cells()
        x = (np.arange(nx, dtype=np.float64) + 0.5) * (1.0 / nx)
        y = (np.arange(ny, dtype=np.float64) + 0.5) * (1.0 / ny)
        z = (np.arange(nz, dtype=np.float64) + 0.5) * (1.0 / nz)
        return x[None, None, :], y[None, :, None], z[:, None, None]

    def set_images(self, img0, img1):
        """initImgs@finest with the inner cells of both images, shape (nz, ny, nx); apply bc is part of init_fields"""
        nx, ny, nz = self.cells()
        for which, img in ((IMG0, img0), (IMG1, img1)):
            a = np.zeros((nz + 2, ny + 2, nx + 2))
            a[1:-1, 1:-1, 1:-1] = img
            self.set_field(self.max_level, which, a)

    def init_images(self):
        """the program's analytic images (OpticalFlow3D.exa4:334-346)"""
        x, y, z = self.cell_centres()
        self.set_images(np.sin(x) + np.cos(y) + np.sin(z), np.sin(x - 0.01) + np.cos(y + 0.01) + np.sin(z - 0.005))

    def init_fields(self):
        rt.check(self.ctx.L.examg_optflow3_init_fields(self._h))

    def smooth(self, lvl, nsweeps=1):
        rt.check(self.ctx.L.examg_optflow3_smooth(self._h, lvl, nsweeps))

    def smooth_color(self, lvl, color):
        rt.check(self.ctx.L.examg_optflow3_smooth_color(self._h, lvl, color))

    def residual(self, lvl=-1):
        rt.check(self.ctx.L.examg_optflow3_residual(self._h, lvl, None))

    def update_residual_norm(self, lvl=-1):
        out = C.c_double(0.0)
        rt.check(self.ctx.L.examg_optflow3_residual(self._h, lvl, C.byref(out)))
        return out.value

    def restrict(self, lvl):
        rt.check(self.ctx.L.examg_optflow3_restrict(self._h, lvl))

    def correct(self, lvl):
        rt.check(self.ctx.L.examg_optflow3_correct(self._h, lvl))

    def cg(self):
        it = C.c_int(0)
        rt.check(self.ctx.L.examg_optflow3_cg(self._h, C.byref(it)))
        return it.value

    def last_cg_iters(self):
        it = C.c_int(0)
        rt.check(self.ctx.L.examg_optflow3_last_cg_iters(self._h, C.byref(it)))
        return it.value

    def vcycle(self, lvl=-1):
        rt.check(self.ctx.L.examg_optflow3_vcycle(self._h, lvl))

    def cycles(self, n):
        out = C.c_double(0.0)
        rt.check(self.ctx.L.examg_optflow3_cycles(self._h, n, C.byref(out)))
        return out.value

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
