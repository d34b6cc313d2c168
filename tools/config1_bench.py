#!/usr/bin/env python
"""BASELINE config 1 on one GPU: Examples/Poisson 2D 5-point FD, V(3,3) red-black GS (omega 0.8), CG coarse solve,
double, max level 10 -- the 1025^2-node grid of Testing/Poisson/2D_FD_Poisson_fromL4 (4x4 fragments, levels 0..8).

  python tools/config1_bench.py [--reps 50]

Prints the time of one mgCycle@finest + residual norm (CUDA events on the launching stream) and DoF-cycles/s.
The grid is 8.4 MB per field: everything lives in L2 and the cycle is launch/latency bound (one CUDA graph).
Development tool; bench.py is the contract (this configuration is a parity-test case there).
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=50)
    args = ap.parse_args()
    import numpy as np
    import torch
    from exastencils_b200 import MultigridProgram
    stream = torch.cuda.Stream()
    prog = MultigridProgram(2, (4, 4), 0, 8, op="scaled", order="bench", smoother="rbgs", omega=0.8, nu_pre=3, nu_post=3,
                            restrict_scale=1.0, cg_max_it=128, cg_rel_tol=1e-3, stream=stream.cuda_stream)
    L = 8
    x, y, _ = prog.node_positions(L)
    X, Y = x[None, None, :], y[None, :, None]
    pi = np.pi
    u = prog.solution(L)
    u[...] = np.broadcast_to(np.cos(pi * X) - np.sin(2.0 * pi * Y), u.shape)
    u[:, 1:-1, 1:-1] = 0.0
    prog.rhs(L)[...] = np.broadcast_to(pi ** 2 * np.cos(pi * X) - 4.0 * pi ** 2 * np.sin(2.0 * pi * Y), u.shape)
    prog.commit()
    n = prog.dims(L)
    dof = float(n[0] - 2) * (n[1] - 2)
    with torch.cuda.stream(stream):
        r0 = prog.cycles(1)
        for _ in range(3):
            prog.cycles(1)
        torch.cuda.synchronize()
        l0 = prog.ctx.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.reps):
            r = prog.cycles(1)
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.reps
        nl = (prog.ctx.launch_count() - l0) / args.reps
    print("config 1 (2D 5-point, V(3,3) RBGS, %d^2 nodes): %.3f ms per cycle + residual norm, %.0f launches, "
          "%.2f GDoF-cycles/s (residual %.3e -> %.3e)" % (n[0], ms, nl, dof / (ms * 1e-3) / 1e9, r0, r))
    prog.close()


if __name__ == "__main__":
    main()
