#!/usr/bin/env python
"""The finest-level kernels of the V(2,2) cycle at 512^3, one launch each after a warm-up launch, for
`ncu --set full -k regex:'k_rbw|k_resrestrict3|k_stencil'`: pre-smoother k_rbw<4,0,false,true>, post-smoother with the
correction folded in k_rbw<4,0,true>, k_resrestrict3, the norm-only residual pass.  Prints CUDA-event timings when run
without ncu.  Development tool (round 2)."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from exastencils_b200 import MultigridProgram
    from exastencils_b200 import runtime as rt
    from bench import set_benchmark_problem
    stream = torch.cuda.Stream()
    L = int(os.environ.get("PROF_LEVEL", "9"))
    prog = MultigridProgram(3, (1, 1, 1), 2, L, smoother="rbgs", omega=0.8, nu_pre=2, nu_post=2, stream=stream.cuda_stream)
    set_benchmark_problem(prog)
    ctx = prog.ctx
    A, u, f, r = prog.stencil[L], prog.f_solution[L], prog.f_rhs[L], prog.f_residual[L]
    uc, fc = prog.f_solution[L - 1], prog.f_rhs[L - 1]

    def timed(name, fn, reps=int(os.environ.get("PROF_REPS", "2"))):
        with torch.cuda.stream(stream):
            fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(reps):
                fn()
            e1.record(stream)
            torch.cuda.synchronize()
        print("%-52s %8.3f ms" % (name, e0.elapsed_time(e1) / reps), flush=True)

    def norm_only():
        out = C.c_double()
        rt.check(ctx.L.examg_solver_cycles(prog._solver, 0, C.byref(out)))

    timed("pre-smoother, 2 sweeps (k_rbw<4,0,false,true>)", lambda: ctx.smooth_rbgs(A, u, f, 0.8, 2))
    timed("correction + post-smoother (k_rbw<4,0,true>)", lambda: ctx.prolong_add_smooth_rbgs(A, uc, u, f, 0.8, 2))
    timed("residual + restriction (k_resrestrict3)", lambda: ctx.residual_restrict(A, u, f, r, fc))
    timed("residual norm, Residual not written (k_stencil<2,7,0>)", norm_only)
    timed("residual + norm, Residual written", lambda: ctx.residual(A, u, f, r, norm=True))
    timed("residual (k_stencil<1,7,0>)", lambda: ctx.residual(A, u, f, r))
    prog.close()


if __name__ == "__main__":
    main()
