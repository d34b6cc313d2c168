#!/usr/bin/env python
"""Stencil-field (variable-coefficient) damped Jacobi on one GPU: two generic sweeps against one fused pass
(`k_jacv2`, csrc/examg_jacv.cuh), CUDA events on the launching stream, and a bit-for-bit comparison of the two
results at the benchmark size (BASELINE config 3's operator shape: 27-point, omega = 0.85).

  python tools/varcoef_bench.py [--level 8] [--reps 10] [--npts 27|7]

EXAMG_JACV_FUSE=2 also routes the 7-point stencil field through the fused kernel.  Development tool; bench.py is
the contract.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--level", type=int, default=8)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--npts", type=int, default=27)
    args = ap.parse_args()
    import numpy as np
    import torch
    from exastencils_b200 import runtime as rt
    peak = 6650.0
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = float(json.load(open(p))["hbm_gbs"])
    stream = torch.cuda.Stream()
    L, npts = args.level, args.npts
    ctx = rt.Context(3, L, L, block_cells=(1, 1, 1))
    ctx.set_stream(stream.cuda_stream)
    n = ctx.dims(L)
    shape = ctx.shape(L)
    pts = float(n[0] - 2) * (n[1] - 2) * (n[2] - 2)
    if npts == 27:
        offs = [(dx, dy, dz) for dz in (-1, 0, 1) for dy in (-1, 0, 1) for dx in (-1, 0, 1)]
    else:
        offs = [(0, 0, 0), (-1, 0, 0), (1, 0, 0), (0, -1, 0), (0, 1, 0), (0, 0, -1), (0, 0, 1)]
    centre = offs.index((0, 0, 0))
    rng = np.random.default_rng(1)
    base = rng.uniform(-1.0, 1.0, size=shape)
    scale = rng.uniform(0.5, 1.0, size=npts)
    coef = np.empty((npts,) + shape)
    for k in range(npts):
        np.multiply(np.roll(base, 7 * k, axis=2), scale[k], out=coef[k])
    coef[centre] = (30.0 if npts == 27 else 8.0) + np.abs(coef[centre])
    cf = ctx.field_create("coef", L, 1, npts)
    ctx.upload(cf, coef)
    del coef
    A = ctx.stencil_field(L, offs, cf)
    fu, ff = ctx.field_create("u", L, 2, 1), ctx.field_create("f", L)
    u0 = rng.uniform(-1.0, 1.0, size=shape)
    ctx.upload(ff, np.roll(base, 3, axis=1))
    bpp = 8.0 * npts + 24.0

    def reset():
        ctx.upload(fu, u0, 0)
        ctx.upload(fu, u0, 1)

    def timeit(name, fn, sweeps):
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
        gbs = sweeps * bpp * pts / (ms * 1e-3) / 1e9
        print("%-44s %8.3f ms  %4.1f launches  %8.1f GB/s alg  %5.1f%% of %.0f" % (name, ms, nl, gbs, 100 * gbs / peak, peak),
              flush=True)
        return ms

    with torch.cuda.stream(stream):
        reset()
        ctx.smooth_jacobi(A, fu, ff, 0.85, 1)
        ctx.smooth_jacobi(A, fu, ff, 0.85, 1)
        want = ctx.download(fu, L)
        reset()
        ctx.smooth_jacobi(A, fu, ff, 0.85, 2)
        got = ctx.download(fu, L)
    same = bool(np.array_equal(want, got))
    print("%d-point stencil field @ level %d (%d^3 nodes): one pass of two sweeps == two single sweeps, bit for bit: %s"
          % (npts, L, n[0], same), flush=True)
    reset()
    t1 = timeit("%d-pt Jacobi, 2 x 1 sweep (%.0f B/pt/sweep)" % (npts, bpp), lambda: (ctx.smooth_jacobi(A, fu, ff, 0.85, 1), ctx.smooth_jacobi(A, fu, ff, 0.85, 1)), 2)
    t2 = timeit("%d-pt Jacobi, 2 sweeps in one call" % npts, lambda: ctx.smooth_jacobi(A, fu, ff, 0.85, 2), 2)
    print("speed-up of the pair: %.2fx" % (t1 / t2))
    ctx.close()
    if not same:
        sys.exit(1)


if __name__ == "__main__":
    main()
