#!/usr/bin/env python
"""Timings of the multi-GPU building blocks (run under torch.distributed.run, one rank per GPU):
halo exchange per level and depth, global sum, whole cycle.  Development tool.

  python -m torch.distributed.run --nproc-per-node 2 tools/mgpu_microbench.py --level 9 --blocks 1,1,2
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--level", type=int, default=9)
    ap.add_argument("--blocks", default="")
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from exastencils_b200 import MultigridProgram
    from exastencils_b200.distributed import wire_peers
    from bench import BLOCKS, set_benchmark_problem
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nb = tuple(int(v) for v in args.blocks.split(",")) if args.blocks else BLOCKS[world]
    stream = torch.cuda.Stream()
    L = args.level
    prog = MultigridProgram(3, (1, 1, 1), 2, L, smoother="rbgs", omega=0.8, nu_pre=2, nu_post=2, num_blocks=nb, rank=rank,
                            device=local, stream=stream.cuda_stream, domain_hi=tuple(float(b) for b in nb))
    wire_peers(prog.ctx, dist, device="cuda:%d" % local)
    set_benchmark_problem(prog)
    ctx = prog.ctx

    def timeit(name, fn):
        with torch.cuda.stream(stream):
            for _ in range(3):
                fn()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(args.reps):
                fn()
            e1.record(stream)
            torch.cuda.synchronize()
            us = 1e3 * e0.elapsed_time(e1) / args.reps
        t = torch.tensor([us], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print("%-44s %9.1f us" % (name, float(t.item())), flush=True)

    import ctypes as C
    lib = ctx.L
    lib.examg_exchange_depth.restype = C.c_int
    lib.examg_exchange_depth.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
    for lvl in range(L, max(L - 4, prog.min_level), -1):
        u = prog.f_solution[lvl]
        for depth in (1, 2, 4):
            timeit("exchange Solution@%d depth %d" % (lvl, depth),
                   lambda: lib.examg_exchange_depth(ctx.h, u, -1, depth))
    A, u, f, r = prog.stencil[L], prog.f_solution[L], prog.f_rhs[L], prog.f_residual[L]
    timeit("rbgs 2 sweeps @%d (exchange + fused kernel)" % L, lambda: ctx.smooth_rbgs(A, u, f, 0.8, 2))
    timeit("residual @%d (exchange + kernel)" % L, lambda: ctx.residual(A, u, f, r))
    timeit("residual + global norm @%d" % L, lambda: ctx.residual(A, u, f, r, norm=True))
    timeit("dot (global sum) @%d" % (L - 3), lambda: ctx.dot(prog.f_residual[L - 3], prog.f_residual[L - 3]))
    set_benchmark_problem(prog)
    timeit("V(2,2) cycle", lambda: prog.vcycle())
    timeit("V(2,2) cycle + residual norm", lambda: prog.cycles(1))
    prog.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
