#!/usr/bin/env python
"""Testing/Application/ExaStokes_{2D,3D} through libexamg: time per V(3,3) cycle (CUDA events) on the reference's grid
(levels 2..8, 256^2 cells) and on larger ones, next to the CPU oracle on the host cores.  Development tool."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from exastencils_b200 import StokesProgram
    from oracle import stokes
    for dim, lo, hi in ((2, 2, 8), (2, 2, 10), (2, 2, 12), (3, 1, 6), (3, 1, 8)):
        g = StokesProgram(lo, hi, dim=dim)
        g.apply_bc(hi)
        for _ in range(2):
            g.vcycle()
        torch.cuda.synchronize()
        l0 = g.ctx.launch_count()
        t0 = time.perf_counter()
        reps = 5
        for _ in range(reps):
            g.vcycle()
        g.ctx.sync()
        ms = (time.perf_counter() - t0) * 1e3 / reps
        launches = (g.ctx.launch_count() - l0) // reps
        n = g.cells(hi)
        cpu = float("nan")
        if hi <= (10 if dim == 2 else 6):
            r = stokes.Solver(lo, hi, dim=dim)
            r.apply_bc(hi)
            r.vcycle()
            t0 = time.perf_counter()
            r.vcycle()
            cpu = (time.perf_counter() - t0) * 1e3
            r.close()
        print("%dD levels %d..%d (%d^%d cells, %.1f M unknowns): %.3f ms per V(3,3) cycle, %d launches; CPU oracle %.1f ms" %
              (dim, lo, hi, n, dim, (dim + 1.0) * n ** dim / 1e6, ms, launches, cpu), flush=True)
        g.close()


if __name__ == "__main__":
    main()
