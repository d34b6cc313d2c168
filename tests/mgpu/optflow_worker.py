"""One rank of the multi-GPU check of the optical-flow program (launched by torch.distributed.run, one rank per GPU).

Every rank owns one block of Benchmark/OptFlow2D (two coupled unknowns on cell fields, red-black 2 x 2 block smoother
with `communicate` between the colours, CG on the coarsest level with global sums).  After every cycle the residual norm,
and at the end the assembled flow field, are compared with (a) the same program on ONE block covering the global grid
on rank 0's GPU and (b) the CPU oracle: fields bit-identical, norms within 1e-12.

  python -m torch.distributed.run --nproc-per-node 2 tests/mgpu/optflow_worker.py --level 8 --blocks 2,1
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--level", type=int, default=8)
    ap.add_argument("--min-level", type=int, default=2)
    ap.add_argument("--cycles", type=int, default=3)
    ap.add_argument("--blocks", default="2,1")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from exastencils_b200 import OpticalFlowProgram
    from exastencils_b200 import optflow as gof

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nb = tuple(int(v) for v in args.blocks.split(","))
    assert nb[0] * nb[1] == world
    L, lo = args.level, args.min_level
    prog = OpticalFlowProgram((1, 1), lo, L, num_blocks=nb, rank=rank, device=local, dist=dist)
    nx, ny = prog.cells()
    gnx, gny = nx * nb[0], ny * nb[1]
    img0, img1 = gof.analytic_images(gnx, gny)
    sl = prog.block_slice()
    prog.set_images(img0[sl], img1[sl])
    norms = [prog.residual_norm()]
    iters = []
    for _ in range(args.cycles):
        norms.append(prog.cycles(1))
        iters.append(prog.last_cg_iters())
    parts = {}
    for which in (gof.U, gof.V):
        mine = torch.from_numpy(np.ascontiguousarray(prog.download(which))).cuda()
        got = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(got, mine)
        parts[which] = got
    ok = True
    if rank == 0:
        def assemble(which):
            out = np.empty((gny, gnx))
            for r in range(world):
                bx, by = r % nb[0], r // nb[0]
                out[by * ny:(by + 1) * ny, bx * nx:(bx + 1) * nx] = parts[which][r].cpu().numpy()
            return out

        one = OpticalFlowProgram(nb, lo, L, device=local)
        one.set_images(img0, img1)
        n1 = [one.residual_norm()]
        it1 = []
        for _ in range(args.cycles):
            n1.append(one.cycles(1))
            it1.append(one.last_cg_iters())
        for a, b in zip(norms, n1):
            if abs(a - b) > 1e-12 * abs(b):
                ok = False
                print("norm differs from the one-block run:", a, b)
        if iters != it1:
            ok = False
            print("CG iterations differ:", iters, it1)
        for which in (gof.U, gof.V):
            got, want = assemble(which), one.download(which)
            if not np.array_equal(got, want):
                ok = False
                print("field %d differs from the one-block run: max |d| = %.3e, %d cells" %
                      (which, float(np.max(np.abs(got - want))), int(np.sum(got != want))))
        one.close()
        if gnx * gny <= 1 << 22:
            from oracle import optflow as oof
            ref = oof.Solver(nb, lo, L, alpha=1.0, relax=1.0, nu_pre=3, nu_post=3, cg_max_it=128, cg_rel_tol=1e-3)
            ref.field(L, oof.IMG0)[1:-1, 1:-1] = img0
            ref.field(L, oof.IMG1)[1:-1, 1:-1] = img1
            ref.init_fields()
            nr = [ref.update_residual_norm()]
            for _ in range(args.cycles):
                ref.vcycle()
                nr.append(ref.update_residual_norm())
            for a, b in zip(norms, nr):
                if abs(a - b) > 1e-12 * abs(b):
                    ok = False
                    print("norm differs from the oracle:", a, b)
            for which, ow in ((gof.U, oof.U), (gof.V, oof.V)):
                if not np.array_equal(assemble(which), ref.field(L, ow)[1:-1, 1:-1]):
                    ok = False
                    print("field %d differs from the oracle" % which)
            ref.close()
        if norms[-1] >= 0.1 * norms[0]:
            ok = False
            print("no convergence:", norms)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    prog.close()
    if rank == 0 and int(flag.item()) == 1:
        print("OPTFLOW MGPU OK", args.blocks, "level", L, "norms", ["%.4e" % v for v in norms], "CG", iters)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
