"""One rank of Testing/Large/Hybrid.{exa4,knowledge} on 4 GPUs (launched by torch.distributed.run).

The reference runs this program on 4 MPI ranks x 10 OpenMP fragments: domain_rect_numBlocks 1x1x4, fragments per block
5x2x1, levels 0..9 => 2560 x 1024 x 2048 cells (5.4e9 unknowns), the Smoothers/Jac cycle (unscaled 7-point stencil,
omega = 0.8 Jacobi, V(3,3), 4.0 * restriction, CG on the coarsest level), Dirichlet data x^2 - y^2/2 - z^2/2, Solution = 0,
RHS = 0.  Here: one block of 5x2x1 fragments per GPU (11 GB per array).  The oracle cannot hold this problem; the check
is the reference's own golden stdout, Testing/Large/Hybrid.results (tests/golden/Large_Hybrid.results).

  python -m torch.distributed.run --nproc-per-node 4 tests/mgpu/large_worker.py [--level 9]
"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--level", type=int, default=9)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from exastencils_b200 import MultigridProgram, runtime as rt
    from exastencils_b200.distributed import wire_peers

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    assert world == 4
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = args.level
    prog = MultigridProgram(3, (5, 2, 1), 0, L, op="unscaled", order="testing", smoother="jacobi", omega=0.8, nu_pre=3,
                            nu_post=3, restrict_scale=4.0, cg_max_it=512, cg_rel_tol=1e-3, num_blocks=(1, 1, 4),
                            rank=rank, device=local)
    wire_peers(prog.ctx, dist, device="cuda:%d" % local)
    ctx = prog.ctx
    f = prog.f_solution[L]
    bc = [(rt.OP_X, 0), (rt.OP_X, 0), (rt.OP_MUL, 0),
          (rt.OP_CONST, 0.5), (rt.OP_Y, 0), (rt.OP_MUL, 0), (rt.OP_Y, 0), (rt.OP_MUL, 0), (rt.OP_SUB, 0),
          (rt.OP_CONST, 0.5), (rt.OP_Z, 0), (rt.OP_MUL, 0), (rt.OP_Z, 0), (rt.OP_MUL, 0), (rt.OP_SUB, 0)]
    ctx.field_set_bc_expr(f, bc)
    ctx.fill(f, 0.0)
    ctx.apply_bc(f, 0)  # apply bc to Solution<0>@9; apply bc to Solution<1>@9
    ctx.apply_bc(f, 1)
    ctx.fill(prog.f_rhs[L], 0.0)

    # Solve: print the residual after every cycle until res < 1e-5 * resStart or 100 cycles, then the cycle count
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res0 = prog.update_residual_norm()
    hist, res, it = [res0], res0, 0
    while not (res < 1.0e-5 * res0 or it >= 100):
        it += 1
        prog.vcycle()
        res = prog.update_residual_norm()
        hist.append(res)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    ok = True
    if rank == 0:
        from oracle import oracle as og
        og.build()
        out = [og.print_reduced_prec(v) for v in hist] + [str(it)]
        print("large: %d cycles on %d GPUs in %.2f s, device bytes per rank %.1f GB" % (it, world, dt, ctx.device_bytes() / 1e9))
        print("large: printed  %s" % " ".join(out))
        if L == 9:
            want = open(os.path.join(ROOT, "tests", "golden", "Large_Hybrid.results")).read().split()
            print("large: golden   %s" % " ".join(want))
            ok = out == want
        print("LARGE GOLDEN %s" % ("OK" if ok else "FAILED"))
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    prog.close()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
