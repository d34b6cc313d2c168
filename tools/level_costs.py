#!/usr/bin/env python
"""Cost of mgCycle@k for every level k of the benchmark hierarchy (CUDA events): shows where the coarse part of
the cycle spends its time.  Development tool."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--level", type=int, default=9)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--smoother", default="rbgs")
    args = ap.parse_args()
    import torch
    from exastencils_b200 import MultigridProgram
    from bench import set_benchmark_problem
    stream = torch.cuda.Stream()
    prog = MultigridProgram(3, (1, 1, 1), 2, args.level, smoother=args.smoother, omega=0.8, nu_pre=2, nu_post=2,
                            stream=stream.cuda_stream)
    set_benchmark_problem(prog)
    ctx = prog.ctx
    prev = 0.0
    for k in range(2, args.level + 1):
        with torch.cuda.stream(stream):
            for _ in range(3):
                ctx.L.examg_solver_vcycle(prog._solver, k)
            torch.cuda.synchronize()
            l0 = ctx.launch_count()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(args.reps):
                ctx.L.examg_solver_vcycle(prog._solver, k)
            e1.record(stream)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.reps
            nl = (ctx.launch_count() - l0) / args.reps
        print("mgCycle@%d: %8.3f ms (%5.1f launches)  this level alone: %8.3f ms" % (k, ms, nl, ms - prev), flush=True)
        prev = ms
    prog.close()


if __name__ == "__main__":
    main()
