"""One run of a benchmark-size program on cuda:0, printing the residual history and a SHA-256 of the final Solution.

tests/test_gpu_fullsize.py runs it twice -- with the fused sm_100a kernels and with every fast path switched off
(EXAMG_NO_FUSED, EXAMG_NO_FUSED_RESTRICT, EXAMG_NO_FAST_PROLONG, EXAMG_NO_GRAPH: the generic kernels that the small
oracle parity tests pin bit for bit) -- and requires identical output: parity at BASELINE.json's full sizes, where the
oracle is too slow to serve as the checker.

  python tests/fullsize_worker.py poisson7|jacobi7|varcoef27 LEVEL CYCLES
"""
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    kind, level, cycles = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    from exastencils_b200 import MultigridProgram
    import bench
    if kind == "varcoef27":
        prog = MultigridProgram(3, (1, 1, 1), 2, level, op="varcoeff27", order="testing", kappa=bench.KAPPA,
                                smoother="jacobi", omega=0.85, nu_pre=2, nu_post=2, cg_max_it=128, cg_rel_tol=1e-3)
        bench.set_varcoef_problem(prog)
    else:
        prog = MultigridProgram(3, (1, 1, 1), 2, level, op="scaled", order="bench",
                                smoother="rbgs" if kind == "poisson7" else "jacobi", omega=0.8, nu_pre=2,
                                nu_post=2 if kind == "poisson7" else 3, cg_max_it=128, cg_rel_tol=1e-3)
        bench.set_benchmark_problem(prog)
        if kind == "jacobi7":  # both slots carry the Dirichlet data
            f = prog.f_solution[level]
            prog.ctx.apply_bc(f, 0)
            prog.ctx.apply_bc(f, 1)
    hist = [prog.update_residual_norm()]
    for _ in range(cycles):
        prog.vcycle()
        hist.append(prog.update_residual_norm())
    u = prog.fetch_solution()
    print("HISTORY " + " ".join(v.hex() for v in hist))
    print("CG " + str(prog.last_cg_iters()))
    print("SHA256 " + hashlib.sha256(u.tobytes()).hexdigest())
    print("LAUNCHES %d" % prog.ctx.launch_count())
    prog.close()


if __name__ == "__main__":
    main()
