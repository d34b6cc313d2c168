#!/usr/bin/env python
"""Per-kernel timings of the optical-flow path (config 4 of BASELINE.json: 2D, cell-centred, 2x2 block smoother).

  python tools/optflow_bench.py --level 13 [--reps 10]

Algorithmic bytes per cell: block RB sweep 80 (u,v,rhs_u,rhs_v + 3 coefficients read, u,v written),
residual of both unknowns 72, restriction of two fields 20, correction of two fields 36.
Development tool; bench.py is the contract.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--level", type=int, default=13)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--nu", type=int, default=3)
    args = ap.parse_args()
    import torch
    from exastencils_b200 import OpticalFlowProgram
    from exastencils_b200 import optflow as gof
    peak = 6650.0
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = float(json.load(open(p))["hbm_gbs"])
    stream = torch.cuda.Stream()
    L = args.level
    prog = OpticalFlowProgram((1, 1), 2, L, nu_pre=args.nu, nu_post=args.nu, stream=stream.cuda_stream)
    ctx = prog.ctx
    nx, ny = prog.cells()
    prog.set_images(*gof.analytic_images(nx, ny))
    cells = float(nx) * ny
    f = {w: prog.field(w) for w in range(9)}
    fc = {w: prog.field(w, L - 1) for w in range(9)}
    sys_h = 0 + (L - 2)  # systems are created per level in order

    def timeit(name, fn, bytes_per_cell, n=cells):
        with torch.cuda.stream(stream):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            l0 = ctx.launch_count()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(args.reps):
                fn()
            e1.record(stream)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.reps
            nl = (ctx.launch_count() - l0) / args.reps
        gbs = bytes_per_cell * n / (ms * 1e-3) / 1e9
        print("%-40s %8.3f ms  %6.1f launches  %8.1f GB/s alg  %5.1f%% of %.0f" % (name, ms, nl, gbs, 100 * gbs / peak, peak),
              flush=True)

    U, V, FU, FV, RU, RV = (f[k] for k in range(6))
    timeit("block RB sweep (80 B/cell)", lambda: ctx.system2_smooth_rb(sys_h, U, V, FU, FV, 1.0, 1), 80.0)
    timeit("block RB %d sweeps" % args.nu, lambda: ctx.system2_smooth_rb(sys_h, U, V, FU, FV, 1.0, args.nu), 80.0 * args.nu)
    timeit("residual u,v (72 B/cell)", lambda: ctx.system2_residual(sys_h, U, V, FU, FV, RU, RV), 72.0)
    timeit("residual + norm", lambda: ctx.system2_residual(sys_h, U, V, FU, FV, RU, RV, norm=True), 72.0)
    timeit("restrict 2 fields (20 B/cell)", lambda: ctx.restrict_cell([RU, RV], [fc[2], fc[3]]), 20.0)
    timeit("correct 2 fields (36 B/cell)", lambda: ctx.prolong_add_cell([fc[0], fc[1]], [U, V]), 36.0)
    cyc = ((2 * args.nu) * 80.0 + 72 + 20 + 36) * 4.0 / 3.0
    prog.set_images(*gof.analytic_images(nx, ny))
    timeit("V(%d,%d) cycle (%.1f B/cell)" % (args.nu, args.nu, cyc), lambda: prog.vcycle(), cyc)
    timeit("V cycle + residual norm", lambda: prog.cycles(1), cyc + 72.0)
    print("unknowns: %d, device memory %.2f GB" % (2 * nx * ny, ctx.device_bytes() / 1e9))
    prog.close()


if __name__ == "__main__":
    main()
