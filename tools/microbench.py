#!/usr/bin/env python
"""Per-kernel timings of libexamg on one GPU (CUDA events on the launching stream).

  python tools/microbench.py --level 9 [--reps 10]

Prints one line per loop shape: ms, points/s, algorithmic GB/s and the fraction of the
measured HBM peak (MEASURED_PEAKS.json).  Development tool; bench.py is the contract.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--level", type=int, default=9)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--nu", type=int, default=2)
    args = ap.parse_args()
    import torch
    from exastencils_b200 import MultigridProgram
    peak = 6650.0
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = float(json.load(open(p))["hbm_gbs"])
    stream = torch.cuda.Stream()
    L = args.level
    prog = MultigridProgram(3, (1, 1, 1), 2, L, smoother="rbgs", omega=0.8, nu_pre=args.nu, nu_post=args.nu,
                            stream=stream.cuda_stream)
    ctx = prog.ctx
    # the benchmark problem (Dirichlet data, RHS = 0): with an all-zero problem the coarse CG divides 0 by 0 and
    # runs to its iteration limit, which is not what a real cycle costs
    sys.path.insert(0, ROOT)
    from bench import set_benchmark_problem
    set_benchmark_problem(prog)
    n = ctx.dims(L)
    pts = float(n[0] - 2) * (n[1] - 2) * (n[2] - 2)
    A, u, f, r = prog.stencil[L], prog.f_solution[L], prog.f_rhs[L], prog.f_residual[L]
    uc, fc = prog.f_solution[L - 1], prog.f_rhs[L - 1]

    def timeit(name, fn, bytes_per_pt, npts=pts):
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
        gbs = bytes_per_pt * npts / (ms * 1e-3) / 1e9
        print("%-34s %8.3f ms  %6.1f launches  %8.1f GB/s alg  %5.1f%% of %.0f" % (name, ms, nl, gbs, 100 * gbs / peak,
                                                                              peak), flush=True)

    timeit("rbgs 1 sweep (24 B/pt)", lambda: ctx.smooth_rbgs(A, u, f, 0.8, 1), 24.0)
    timeit("rbgs %d sweeps (24 B/pt/sweep)" % args.nu, lambda: ctx.smooth_rbgs(A, u, f, 0.8, args.nu), 24.0 * args.nu)
    timeit("rb one colour (12 B/pt)", lambda: ctx.smooth_rb_color(A, u, f, 0.8, 0), 12.0)
    timeit("residual (24 B/pt)", lambda: ctx.residual(A, u, f, r), 24.0)
    timeit("residual + norm (24 B/pt)", lambda: ctx.residual(A, u, f, r, norm=True), 24.0)
    timeit("restrict (9 B/fine pt)", lambda: ctx.restrict(r, fc), 9.0)
    timeit("prolong+correct (17 B/fine pt)", lambda: ctx.prolong_add(uc, u), 17.0)
    timeit("dot (16 B/pt)", lambda: ctx.dot(r, r), 16.0)
    timeit("set coarse (8 B/coarse pt)", lambda: ctx.set(uc, 0.0), 1.0)
    # whole cycle: (4 sweeps*24 + 24 + 9 + 17) * 8/7 per finest point
    cyc_bytes = ((2 * args.nu) * 24.0 + 24 + 9 + 17) * 8.0 / 7.0
    set_benchmark_problem(prog)
    timeit("V(%d,%d) cycle (%.1f B/pt)" % (args.nu, args.nu, cyc_bytes), lambda: prog.vcycle(), cyc_bytes)
    set_benchmark_problem(prog)
    timeit("V cycle + residual norm", lambda: prog.cycles(1), cyc_bytes + 24.0)
    # coarse part alone: mgCycle@(L-1)
    prog2 = prog
    set_benchmark_problem(prog)
    timeit("mgCycle@%d (coarse part)" % (L - 1), lambda: prog2.ctx.L.examg_solver_vcycle(prog2._solver, L - 1), cyc_bytes,
           npts=pts / 8)
    prog.close()
    if args.level > 9:
        return
    # Jacobi (two Solution slots): 1, 2 and 3 sweeps per pass over HBM
    jp = MultigridProgram(3, (1, 1, 1), 2, L, smoother="jacobi", omega=0.8, nu_pre=args.nu, nu_post=args.nu,
                          stream=stream.cuda_stream)
    set_benchmark_problem(jp)
    ctx = jp.ctx
    A, u, f = jp.stencil[L], jp.f_solution[L], jp.f_rhs[L]
    for nu in (1, 2, 3, 4):
        timeit("jacobi %d sweeps (24 B/pt/sweep)" % nu, lambda: ctx.smooth_jacobi(A, u, f, 0.8, nu), 24.0 * nu)
    set_benchmark_problem(jp)
    timeit("Jacobi V(%d,%d) cycle (%.1f B/pt)" % (args.nu, args.nu, cyc_bytes), lambda: jp.vcycle(), cyc_bytes)
    jp.close()
    # 27-point stencil field (BASELINE config 3's operator shape) one level down: Jacobi sweep and residual read
    # 27 coefficients + Solution + RHS and write one value = 240 B per point
    import numpy as np
    from exastencils_b200 import runtime as rt
    L3 = L - 1
    c3 = rt.Context(3, L3, L3, block_cells=(1, 1, 1))
    c3.set_stream(stream.cuda_stream)
    ctx = c3
    n3 = c3.dims(L3)
    pts3 = float(n3[0] - 2) * (n3[1] - 2) * (n3[2] - 2)
    offs = [(dx, dy, dz) for dz in (-1, 0, 1) for dy in (-1, 0, 1) for dx in (-1, 0, 1)]
    cf = c3.field_create("coef", L3, 1, 27)
    rng = np.random.default_rng(1)
    coef = rng.uniform(-1.0, 1.0, size=(27,) + c3.shape(L3))
    coef[13] = 30.0 + np.abs(coef[13])
    c3.upload(cf, coef)
    del coef
    A27 = c3.stencil_field(L3, offs, cf)
    u3, f3, r3 = c3.field_create("u", L3, 2, 1), c3.field_create("f", L3), c3.field_create("r", L3)
    timeit("27-pt stencil-field Jacobi (240 B/pt) @%d" % L3, lambda: c3.smooth_jacobi(A27, u3, f3, 0.8, 1), 240.0, npts=pts3)
    timeit("27-pt stencil-field residual (240 B/pt)", lambda: c3.residual(A27, u3, f3, r3), 240.0, npts=pts3)
    c3.close()


if __name__ == "__main__":
    main()
