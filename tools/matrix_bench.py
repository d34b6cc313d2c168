#!/usr/bin/env python
"""north_star (a) as a table: one sweep of every smoother x stencil x coefficient kind on one level that exceeds L2,
CUDA events, algorithmic GB/s (24 B per point and sweep + 8 B per stencil-field entry) against the measured
copy bandwidth.  3D: 256^3 (stencil fields) / 512^3 (constant coefficients); 2D: 8192^2 / 4096^2.  Development tool."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def offsets(dim, npts):
    if npts in (5, 7):
        offs = [(0, 0, 0)]
        for d in range(dim):
            for s in (-1, 1):
                o = [0, 0, 0]
                o[d] = s
                offs.append(tuple(o))
        return offs
    r = range(-1, 2)
    return [(dx, dy, 0) for dy in r for dx in r] if dim == 2 else [(dx, dy, dz) for dz in r for dy in r for dx in r]


def main():
    import torch
    from exastencils_b200 import runtime as rt
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    stream = torch.cuda.Stream()
    rows = []
    cases = [(3, 7, False), (3, 27, False), (3, 7, True), (3, 27, True), (2, 5, False), (2, 9, False), (2, 5, True),
             (2, 9, True)]
    if len(sys.argv) > 1:  # e.g. `matrix_bench.py 4 6`: only those rows of the list above
        cases = [cases[int(v)] for v in sys.argv[1:]]
    for dim, npts, var in cases:
        lvl = (8 if var else 9) if dim == 3 else (12 if var else 13)
        ctx = rt.Context(dim, lvl, lvl, block_cells=(1,) * dim)
        ctx.set_stream(stream.cuda_stream)
        offs = offsets(dim, npts)
        centre = offs.index((0, 0, 0))
        coef = [-1.0 / (1 + sum(abs(v) for v in o)) for o in offs]
        coef[centre] = -(sum(coef) - coef[centre]) + 0.5
        if var:
            cf = ctx.field_create("coef", lvl, 1, npts)
            for k in range(npts):
                ctx.field_eval(cf, [(rt.OP_CONST, coef[k])], comp=k)
            A = ctx.stencil_field(lvl, offs, cf)
        else:
            A = ctx.stencil_const(lvl, offs, coef)
        u, f, r = ctx.field_create("u", lvl, 2), ctx.field_create("f", lvl), ctx.field_create("r", lvl)
        n = ctx.dims(lvl)
        pts = float(np.prod([v - 2 for v in n[:dim]]))
        bytes_sweep = (24.0 + (8.0 * npts if var else 0.0)) * pts

        def timed(fn, reps=5):
            with torch.cuda.stream(stream):
                for _ in range(2):
                    fn()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(reps):
                    fn()
                e1.record(stream)
                torch.cuda.synchronize()
            return e0.elapsed_time(e1) / reps

        res = {"jacobi": timed(lambda: ctx.smooth_jacobi(A, u, f, 0.8, 1)),
               "jacobi x2": timed(lambda: ctx.smooth_jacobi(A, u, f, 0.8, 2)) / 2,
               "residual": timed(lambda: ctx.residual(A, u, f, r))}
        if npts in (5, 7):
            res["rbgs"] = timed(lambda: ctx.smooth_rbgs(A, u, f, 0.8, 1))
            res["rbgs x2"] = timed(lambda: ctx.smooth_rbgs(A, u, f, 0.8, 2)) / 2
            res["rbgs x3"] = timed(lambda: ctx.smooth_rbgs(A, u, f, 0.8, 3)) / 3
        else:
            res["gs %d-way" % (2 ** dim)] = timed(lambda: ctx.smooth_colored(A, u, f, 0.8, 2, 1))
        for name, ms in res.items():
            gbs = bytes_sweep / (ms * 1e-3) / 1e9
            rows.append((dim, npts, "stencil field" if var else "constant", "x".join(str(v) for v in n[:dim]), name, ms, gbs,
                         gbs / peak))
        ctx.close()
    print("%-3s %-4s %-14s %-16s %-12s %9s %10s %6s" % ("dim", "pts", "coefficients", "nodes", "loop", "ms/sweep", "alg GB/s", "frac"))
    for row in rows:
        print("%-3d %-4d %-14s %-16s %-12s %9.3f %10.0f %6.2f" % row)


if __name__ == "__main__":
    main()
