"""The optical-flow programs (Benchmark/OptFlow2D/2D_FD_OptFlow.exa4, Testing/Application/OpticalFlow2D.exa4)
as host code over libexamg: `Application` / `InitFields` / `doSolve@finest` / `mgCycle@all`.

Fields are cell-localised with Neumann boundaries; the smoother is the red-black 2x2 block solve of
`solve locally`.  All arithmetic runs in the CUDA library; there is no CPU path here.
"""
import ctypes as C

import numpy as np

from . import runtime as rt

U, V, RHS_U, RHS_V, RES_U, RES_V, IXIX, IXIY, IYIY, IMG0, IMG1 = range(11)


def analytic_images(nx, ny):
    """initImgs@finest: img0 = sin(x) + cos(y), img1 = sin(x - 0.01) + cos(y + 0.01) at the cell centres of
    [0,1]^2 (2D_FD_OptFlow.exa4 `initImgs`)."""
    x = (np.arange(nx, dtype=np.float64) + 0.5) * (1.0 / nx)
    y = (np.arange(ny, dtype=np.float64) + 0.5) * (1.0 / ny)
    img0 = np.sin(x)[None, :] + np.cos(y)[:, None]
    img1 = np.sin(x - 0.01)[None, :] + np.cos(y + 0.01)[:, None]
    return np.ascontiguousarray(img0), np.ascontiguousarray(img1)


class OpticalFlowProgram:
    def __init__(self, block_cells, min_level, max_level, alpha=1.0, relax=1.0, nu_pre=3, nu_post=3, cg_max_it=128,
                 cg_rel_tol=1e-3, device=0, stream=None, num_blocks=(1, 1), rank=0, dist=None):
        """num_blocks / rank / dist: this process owns block `rank` (x fastest) of a num_blocks grid of blocks with
        block_cells cells each on level 0; the peers are wired through the process group `dist` (one process per GPU)."""
        self.num_blocks, self.rank = tuple(num_blocks), rank
        self.ctx = rt.Context(2, min_level, max_level, block_cells=tuple(block_cells) + (1,), device=device,
                              num_blocks=tuple(num_blocks) + (1,), rank=rank)
        if stream is not None:
            self.ctx.set_stream(stream)
        self.min_level, self.max_level = min_level, max_level
        d = rt.OptFlowDesc(alpha, relax, nu_pre, nu_post, cg_max_it, cg_rel_tol)
        self._h = self.ctx.L.examg_optflow_create(self.ctx.h, C.byref(d))
        if not self._h:
            raise rt.ExaMGError("examg_optflow_create: " + self.ctx.L.examg_last_error().decode())
        if dist is not None:
            from .distributed import wire_peers
            wire_peers(self.ctx, dist, device="cuda:%d" % device)

    def block_slice(self, level=None):
        """this block's part of a global (ny, nx) cell array of the level"""
        nx, ny = self.cells(level)
        bx, by = self.rank % self.num_blocks[0], self.rank // self.num_blocks[0]
        return (slice(by * ny, (by + 1) * ny), slice(bx * nx, (bx + 1) * nx))

    def close(self):
        if getattr(self, "_h", None):
            self.ctx.L.examg_optflow_destroy(self._h)
            self._h = None
        self.ctx.close()

    def field(self, which, level=None):
        level = self.max_level if level is None else level
        return rt.check(self.ctx.L.examg_optflow_field(self._h, level, which))

    def cells(self, level=None):
        nx, ny, _ = self.ctx.cells(self.max_level if level is None else level)
        return nx, ny

    def set_images(self, img0, img1):
        """upload the two frames (dense (ny, nx) arrays of this block's cells) and run InitFields"""
        self.ctx.upload(self.field(IMG0), np.ascontiguousarray(img0, dtype=np.float64))
        self.ctx.upload(self.field(IMG1), np.ascontiguousarray(img1, dtype=np.float64))
        rt.check(self.ctx.L.examg_optflow_init_fields(self._h))

    def download(self, which, level=None):
        level = self.max_level if level is None else level
        return self.ctx.download_cells(self.field(which, level), level)[0]

    def vcycle(self, level=-1):
        rt.check(self.ctx.L.examg_optflow_vcycle(self._h, level))

    def residual_norm(self):
        out = C.c_double()
        rt.check(self.ctx.L.examg_optflow_residual_norm(self._h, C.byref(out)))
        return out.value

    def cycles(self, n):
        out = C.c_double()
        rt.check(self.ctx.L.examg_optflow_cycles(self._h, n, C.byref(out)))
        return out.value

    def last_cg_iters(self):
        out = C.c_int()
        rt.check(self.ctx.L.examg_optflow_last_cg_iters(self._h, C.byref(out)))
        return out.value

    def solve(self, max_it=128, reduction=1e-5):
        """doSolve@finest: residual norms printed by the program, initial one first"""
        r0 = self.residual_norm()
        hist, r, it = [r0], r0, 0
        while it < max_it and r > reduction * r0:
            it += 1
            r = self.cycles(1)
            hist.append(r)
        return hist
