#!/usr/bin/env python
"""A short program for `ncu --set full -k regex:...`: a few launches of the fused kernels at benchmark size
(3D 512^3: k_rbw, k_jacw, k_resrestrict3, k_prolong3_fast; 2D 8192^2: k_of_rbw, k_of_residual).  Prints CUDA-event
timings when run without ncu.  Development tool."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from exastencils_b200 import MultigridProgram, OpticalFlowProgram
    from exastencils_b200 import optflow as gof
    from bench import set_benchmark_problem
    stream = torch.cuda.Stream()
    L = int(os.environ.get("PROF_LEVEL", "9"))

    def timed(name, fn, reps=3):
        with torch.cuda.stream(stream):
            fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(reps):
                fn()
            e1.record(stream)
            torch.cuda.synchronize()
        print("%-36s %8.3f ms" % (name, e0.elapsed_time(e1) / reps), flush=True)

    for smoother in ("rbgs", "jacobi"):
        prog = MultigridProgram(3, (1, 1, 1), 2, L, smoother=smoother, omega=0.8, nu_pre=2, nu_post=2,
                                stream=stream.cuda_stream)
        set_benchmark_problem(prog)
        ctx = prog.ctx
        A, u, f, r = prog.stencil[L], prog.f_solution[L], prog.f_rhs[L], prog.f_residual[L]
        uc, fc = prog.f_solution[L - 1], prog.f_rhs[L - 1]
        if smoother == "rbgs":
            timed("rbgs 2 sweeps (k_rbw<4>)", lambda: ctx.smooth_rbgs(A, u, f, 0.8, 2))
            timed("rbgs 1 sweep (k_rbw<2>)", lambda: ctx.smooth_rbgs(A, u, f, 0.8, 1))
            timed("residual+restrict (k_resrestrict3)", lambda: ctx.residual_restrict(A, u, f, r, fc))
            timed("prolong+correct (k_prolong3_fast)", lambda: ctx.prolong_add(uc, u))
        else:
            timed("jacobi 2 sweeps (k_jacw<2>)", lambda: ctx.smooth_jacobi(A, u, f, 0.8, 2))
        prog.close()
    L2 = int(os.environ.get("PROF_LEVEL_2D", "13"))
    of = OpticalFlowProgram((1, 1), 2, L2, stream=stream.cuda_stream)
    nx, ny = of.cells()
    of.set_images(*gof.analytic_images(nx, ny))
    ctx = of.ctx
    F = [of.field(k) for k in range(6)]
    sys_h = L2 - 2
    timed("optflow 3 sweeps (k_of_rbw<6>)", lambda: ctx.system2_smooth_rb(sys_h, F[0], F[1], F[2], F[3], 1.0, 3))
    timed("optflow 2 sweeps (k_of_rbw<4>)", lambda: ctx.system2_smooth_rb(sys_h, F[0], F[1], F[2], F[3], 1.0, 2))
    timed("optflow residual (k_of_residual)", lambda: ctx.system2_residual(sys_h, F[0], F[1], F[2], F[3], F[4], F[5]))
    of.close()


if __name__ == "__main__":
    main()
