"""One rank of the multi-GPU parity check (launched by torch.distributed.run, one rank per GPU).

Every rank owns one block of the Benchmark/Poisson3D program (V(2,2) RBGS, CG on the agglomerated
coarse grid), fed with a seeded random global Solution/RHS.  After every cycle the residual norm
and -- at the end -- the assembled global Solution are compared with the oracle's run of the same
program on the global grid: fields bit-exact, norms within 1e-13 relative.

  python -m torch.distributed.run --nproc-per-node 2 tests/mgpu/worker.py --level 7 --cycles 3
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--level", type=int, default=7)
    ap.add_argument("--min-level", type=int, default=2)
    ap.add_argument("--cycles", type=int, default=3)
    ap.add_argument("--sweeps", type=int, default=2, help="nu_pre = nu_post")
    ap.add_argument("--smoother", default="rbgs")
    ap.add_argument("--blocks", default="", help="domain_rect_numBlocks as x,y,z (default: x first)")
    ap.add_argument("--op", default="scaled", help="scaled | varcoeff27 (27-point stencil field assembled on the device)")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from exastencils_b200 import MultigridProgram
    from exastencils_b200.distributed import BLOCK_GRIDS, block_slices, wire_peers

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nb = tuple(int(v) for v in args.blocks.split(",")) if args.blocks else BLOCK_GRIDS[world]
    assert nb[0] * nb[1] * nb[2] == world
    L = args.level
    var27 = args.op == "varcoeff27"
    omega = 0.85 if var27 else 0.8
    kw = dict(op=args.op, order="bench", kappa=10.0 if var27 else 0.0, smoother=args.smoother, omega=omega,
              nu_pre=args.sweeps, nu_post=args.sweeps, cg_max_it=128, cg_rel_tol=1e-3)
    prog = MultigridProgram(3, (1, 1, 1), args.min_level, L, num_blocks=nb, rank=rank, device=local, **kw)
    wire_peers(prog.ctx, dist, device="cuda:%d" % local)

    gshape = tuple((nb[d] << L) + 1 for d in (2, 1, 0))
    rng = np.random.default_rng(1234)
    u0 = rng.uniform(-1.0, 1.0, size=gshape)
    f0 = rng.uniform(-1.0, 1.0, size=gshape)
    sl = block_slices(rank, nb, gshape)
    prog.solution(L)[...] = u0[sl]
    prog.rhs(L)[...] = f0[sl]
    prog.commit()

    norms = [prog.update_residual_norm()]
    for _ in range(args.cycles):
        prog.vcycle()
        norms.append(prog.update_residual_norm())
    mine = torch.from_numpy(np.ascontiguousarray(prog.fetch_solution())).cuda()
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)
    ok = True
    if rank == 0 and var27:
        # the operator is assembled on the device (exp of the device's libm), so the bit-level reference is the same
        # program on ONE block covering the global grid; the oracle (host exp) is compared to 1e-9
        one = MultigridProgram(3, nb, args.min_level, L, device=local, **kw)
        one.solution(L)[...] = u0
        one.rhs(L)[...] = f0
        one.commit()
        n1 = [one.update_residual_norm()]
        for _ in range(args.cycles):
            one.vcycle()
            n1.append(one.update_residual_norm())
        got = np.empty(gshape)
        for r in range(world):
            got[block_slices(r, nb, gshape)] = parts[r].cpu().numpy()
        exp1 = one.fetch_solution()
        one.close()
        nbad1 = int(np.count_nonzero(got != exp1))
        rel1 = max(abs(a - b) / b for a, b in zip(norms, n1))
        print("mgpu: varcoeff27 vs one block on the global grid: fields differing in %d of %d nodes, max rel norm "
              "diff %.2e" % (nbad1, got.size, rel1))
        ok = ok and nbad1 == 0 and rel1 <= 1e-13
    if rank == 0:
        from oracle import oracle as og
        og.build()
        ref = og.Solver(3, nb, args.min_level, L, op=og.OP_VARCOEFF27 if var27 else og.OP_LAPLACE_SCALED,
                        order=og.ORDER_BENCH, kappa=10.0 if var27 else 0.0,
                        smoother=og.SMOOTH_RBGS if args.smoother == "rbgs" else og.SMOOTH_JACOBI, omega=omega,
                        nu_pre=args.sweeps, nu_post=args.sweeps, cg_max_it=128, cg_rel_tol=1e-3)
        ref.solution(L)[...] = u0
        ref.rhs(L)[...] = f0
        ref.sync_slots()
        rn = [ref.update_residual_norm()]
        for _ in range(args.cycles):
            ref.vcycle()
            rn.append(ref.update_residual_norm())
        got = np.empty(gshape)
        for r in range(world):
            got[block_slices(r, nb, gshape)] = parts[r].cpu().numpy()
        exp = ref.solution(L)
        nbad = int(np.count_nonzero(got != exp))
        rel = max(abs(a - b) / b for a, b in zip(norms, rn))
        print("mgpu: world %d blocks %s level %d: norms %s" % (world, nb, L, ["%.6e" % v for v in norms]))
        print("mgpu: oracle norms            %s" % ["%.6e" % v for v in rn])
        print("mgpu: fields differing in %d of %d nodes (max abs %.3e), max rel norm diff %.2e" % (
            nbad, got.size, float(np.max(np.abs(got - exp))), rel))
        if var27:
            ok = ok and float(np.max(np.abs(got - exp))) <= 1e-9 * float(np.max(np.abs(exp))) and rel <= 1e-9
        else:
            ok = nbad == 0 and rel <= 1e-13
        print("MGPU PARITY %s" % ("OK" if ok else "FAILED"))
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    prog.close()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
