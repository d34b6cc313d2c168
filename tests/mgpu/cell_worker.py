"""One rank of the multi-GPU check of cell-localised fields (launched by torch.distributed.run, one rank per GPU).

Every rank owns one block of a Testing/CellBased program (cell-centred finite-volume Poisson: ghost cells at block faces
come from `communicate`, at the physical boundary from `apply bc`; CG on the coarsest level with global dot products).
The run must print the reference's golden stdout, and the assembled global Solution must equal the CPU oracle's run on
the global grid (bit for bit for the Neumann programs, 1e-12 where the Dirichlet data goes through the device's libm).

  python -m torch.distributed.run --nproc-per-node 2 tests/mgpu/cell_worker.py --name 3D_Neumann --blocks 1,1,2
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--name", default="2D_Neumann")
    ap.add_argument("--blocks", default="2,1")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import cell_programs as cp
    from oracle import oracle as og

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nb = tuple(int(v) for v in args.blocks.split(","))
    assert int(np.prod(nb)) == world
    prog = [p for p in cp.PROGRAMS if p.name == args.name][0]
    assert len(nb) == prog.dim
    g = cp.GpuCells(prog, num_blocks=nb, rank=rank, device=local, dist=dist)
    out = cp.solve(g, og.print_reduced_prec)
    L = prog.max_level
    mine = torch.from_numpy(np.ascontiguousarray(g.interior("u", L))).cuda()
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)
    ok = out == prog.expected()
    if rank == 0:
        if not ok:
            print("stdout differs:", out, prog.expected())
        ref = cp.OracleCells(prog)
        cp.solve(ref, og.print_reduced_prec)
        want = ref.interior("u", L)
        shape = tuple(mine.shape)
        got = np.empty((shape[0] * g.nb[2], shape[1] * g.nb[1], shape[2] * g.nb[0]))
        for r in range(world):
            b = [r % g.nb[0], (r // g.nb[0]) % g.nb[1], r // (g.nb[0] * g.nb[1])]
            sl = tuple(slice(b[2 - k] * shape[k], (b[2 - k] + 1) * shape[k]) for k in range(3))
            got[sl] = parts[r].cpu().numpy()
        want = want.reshape(got.shape)
        if prog.neumann:
            same = np.array_equal(got, want)
        else:
            same = np.allclose(got, want, rtol=1e-12, atol=1e-12 * np.max(np.abs(want)))
        if not same:
            print("fields differ: max |d| = %.3e, %d entries" % (float(np.max(np.abs(got - want))), int(np.sum(got != want))))
        ok = ok and same
        ref.close()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    g.close() if hasattr(g, "close") else None
    if rank == 0 and int(flag.item()) == 1:
        print("CELL MGPU OK", args.name, args.blocks, out[-1], "cycles")
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
