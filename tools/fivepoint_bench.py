#!/usr/bin/env python
"""Benchmark/FivePointStencil/5pt_Jac_Cell.exa4 on one GPU: single-level 2D cell-centred 5-point omega = 0.8 Jacobi on
[0,2] x [0,1] (level 11 on 2 x 2 fragments = 4096^2 cells, anisotropic widths), Dirichlet data through the user-written
ghost-cell function applyBC_solution (value 0 on three faces, sin(2 pi x) sinh(2 pi) on the top face), one sweep +
residual + norm per iteration -- the third benchmark app of the reference, the pure-smoother bandwidth test.

  python tools/fivepoint_bench.py [--level 11] [--its 100]

Prints the Solve loop's time per iteration and the Jacobi sweep alone (24 B per cell).  Development tool.
"""
import argparse
import json
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def setup(level, stream=None):
    import numpy as np
    from exastencils_b200 import runtime as rt
    ctx = rt.Context(2, level, level, block_cells=(2, 2), domain_hi=(2.0, 1.0, 1.0))
    if stream is not None:
        ctx.set_stream(stream)
    hx, hy, _ = ctx.width(level)
    offs = [(0, 0, 0), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0)]
    A = ctx.stencil_const(level, offs, [(2.0 / (hx * hx)) + (2.0 / (hy * hy)), -1.0 / (hx * hx), -1.0 / (hx * hx),
                                        -1.0 / (hy * hy), -1.0 / (hy * hy)])
    u, f, r = ctx.field_create_cell("Solution", level, 2), ctx.field_create_cell("RHS", level), ctx.field_create_cell("Residual", level)
    # applyBC_solution: ghost = value - inner = 2 * (value / 2) - inner; value = 0 except on the top face, where it is
    # sin(2 pi vf_nodePosition_x) * sinh(2 pi) with the NODE position of the cell (= cell centre - hx / 2)
    ctx.field_set_bc(u, rt.BC_DIRICHLET_CONST, 0.0)
    top = [(rt.OP_CONST, 0.5), (rt.OP_CONST, 2.0 * math.pi), (rt.OP_X, 0), (rt.OP_CONST, 0.5 * hx), (rt.OP_SUB, 0),
           (rt.OP_MUL, 0), (rt.OP_SIN, 0), (rt.OP_CONST, math.sinh(2.0 * math.pi)), (rt.OP_MUL, 0), (rt.OP_MUL, 0)]
    ctx.field_set_bc_face(u, 3, rt.BC_DIRICHLET_EXPR, prog=top)
    ctx.field_set_bc(r, rt.BC_DIRICHLET_CONST, 0.0)
    nx, ny, _ = ctx.cells(level)
    xc = (np.arange(nx) * hx) + 0.5 * hx
    yc = (np.arange(ny) * hy) + 0.5 * hy
    rhs = (4.0 * math.pi * math.pi) * np.sin(2.0 * math.pi * xc)[None, None, :] * np.sinh(2.0 * math.pi * yc)[None, :, None]
    ctx.upload(f, np.ascontiguousarray(rhs))
    ctx.apply_bc(u)
    return ctx, A, u, f, r, (nx, ny, hx, hy)


def iteration(ctx, A, u, f, r):
    """Smoother@finest; UpResidual@finest; NormResidual@finest"""
    ctx.apply_bc(u)
    ctx.smooth_jacobi(A, u, f, 0.8, 1)
    ctx.apply_bc(u)
    ctx.residual(A, u, f, r)
    return math.sqrt(ctx.dot(r, r))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--level", type=int, default=11)
    ap.add_argument("--its", type=int, default=100)
    args = ap.parse_args()
    import torch
    peak = 6650.0
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = float(json.load(open(p))["hbm_gbs"])
    stream = torch.cuda.Stream()
    ctx, A, u, f, r, (nx, ny, hx, hy) = setup(args.level, stream.cuda_stream)
    with torch.cuda.stream(stream):
        ctx.residual(A, u, f, r)
        res0 = math.sqrt(ctx.dot(r, r))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.its):
            res = iteration(ctx, A, u, f, r)
        e1.record(stream)
        torch.cuda.synchronize()
        ms_it = e0.elapsed_time(e1) / args.its
        e0.record(stream)
        for _ in range(args.its):
            ctx.smooth_jacobi(A, u, f, 0.8, 1)
        e1.record(stream)
        torch.cuda.synchronize()
        ms_k = e0.elapsed_time(e1) / args.its
    cells = float(nx) * ny
    gbs = 24.0 * cells / (ms_k * 1e-3) / 1e9
    print("5pt_Jac_Cell %d x %d cells: residual %.4e -> %.4e after %d iterations; %.3f ms per iteration (sweep + residual "
          "+ norm); Jacobi sweep alone %.4f ms = %.1f GB/s algorithmic (24 B per cell) = %.1f %% of %.0f"
          % (nx, ny, res0, res, args.its, ms_it, ms_k, gbs, 100.0 * gbs / peak, peak))
    ctx.close()


if __name__ == "__main__":
    main()
