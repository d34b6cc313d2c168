#!/usr/bin/env python
"""bench.py -- 3D Poisson V-cycle throughput (DoF-cycles/s) on N B200s.

Workload: Benchmark/Poisson3D 7-point constant-coefficient Poisson, V(2,2) red-black
Gauss-Seidel (omega 0.8), CG coarse solve, double precision, Dirichlet data
x^2 - y^2/2 - z^2/2, RHS = 0.
  N = 1   BASELINE.json configs[1]: 512^3 cells (levels 2..9 as in
          Benchmark/Poisson3D/3D_FD_Poisson_fromL4_CUDA.knowledge).  The line also carries a
          "config5_n1" record: the same step at 1024^3 cells, the N = 1 point of configs[4].
  N > 1   BASELINE.json configs[4]: weak scaling, one block of 1024^3 cells per GPU (levels 2..10),
          domain_rect_numBlocks from BLOCKS, domain [0,nbx]x[0,nby]x[0,nbz] (the mesh width and with it
          the operator stay those of one block).  --scaling strong: 1024^3 cells in all.
  --level / --blocks override both.

A step = one mgCycle@finest + the outer-loop residual and its L2 norm
(Solve@finest's loop body, ...fromL4.exa4:135-149).  Prints ONE JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# domain_rect_numBlocks_{x,y,z} (rank -> block x fastest, C/s/e/domain/ir/IR_InitGeneratedDomain.scala:41-47); the
# domain grows with the block grid ([0,nbx]x[0,nby]x[0,nbz]), so the mesh width and with it the operator stay those
# of the 1-GPU problem (weak scaling of the same isotropic problem).  --blocks x,y,z overrides; BLOCKS_X_FIRST is the
# reference's own choice for 2/4/8 ranks and what --scaling strong uses.
BLOCKS = {1: (1, 1, 1), 2: (1, 1, 2), 4: (1, 2, 2), 8: (1, 2, 4)}
BLOCKS_X_FIRST = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}
# --scaling strong: 1024^3 cells in all; every block has 512 cells in a split dimension, 1024 in the others
# (block_cells at level 0 x 2^9), and the level range 1..9 gives the hierarchy of the one-block run (2..10)
STRONG_LEVEL, STRONG_MIN_LEVEL = 9, 1
PARITY_LEVEL = 7  # per-block finest level of the cross-N parity check (outside the timed region)
NU_PRE = NU_POST = 2
OMEGA = 0.8
MIN_LEVEL = 2
# algorithmic bytes per fine-grid point (DESIGN.md "Byte model")
BYTES_RBGS_SWEEP = 24.0
BYTES_RESIDUAL = 24.0


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            c = [x.strip() for x in r.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0]))
                mx = float(c[1])
            except ValueError:
                continue
            for nm, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


KAPPA = 10.0  # getCoefficient of Testing/SISC/3D_VarCoeff.exa4

# --workload: poisson7 = BASELINE configs[1] (the bench contract), varcoef27 = configs[2] (27-point stencil field,
# damped Jacobi with pairs of sweeps fused into one pass over the coefficients; 1 GPU)
VBLOCKS = {1: (1, 1, 1), 8: (2, 2, 2)}  # varcoef27: the unit cube, cubic cells
WORKLOADS = {
    "poisson7": dict(op="scaled", order="bench", smoother="rbgs", omega=OMEGA, kappa=0.0, bytes_sweep=24.0,
                     bytes_residual=24.0),
    "varcoef27": dict(op="varcoeff27", order="testing", smoother="jacobi", omega=0.85, kappa=KAPPA,
                      bytes_sweep=240.0, bytes_residual=240.0),
}


def _bump(rt, op):  # X - X * X
    return [(op, 0), (op, 0), (op, 0), (rt.OP_MUL, 0), (rt.OP_SUB, 0)]


def set_varcoef_problem(prog):
    """Testing/SISC/3D_VarCoeff.exa4 on the 27-point operator: RHS = 2 kappa ((x-x^2)(y-y^2) + (x-x^2)(z-z^2) +
    (y-y^2)(z-z^2)), Dirichlet data 1 - exp(-kappa (x-x^2)(y-y^2)(z-z^2)), Solution = 0 inside; all on the device."""
    from exastencils_b200 import runtime as rt
    ctx, lvl = prog.ctx, prog.max_level
    bx, by, bz = _bump(rt, rt.OP_X), _bump(rt, rt.OP_Y), _bump(rt, rt.OP_Z)
    mul, add = [(rt.OP_MUL, 0)], [(rt.OP_ADD, 0)]
    rhs = [(rt.OP_CONST, 2.0 * KAPPA)] + bx + by + mul + bx + bz + mul + add + by + bz + mul + add + mul
    ctx.field_eval(prog.f_rhs[lvl], rhs)
    bc = ([(rt.OP_CONST, 1.0), (rt.OP_CONST, -1.0 * KAPPA)] + bx + by + mul + bz + mul + mul +
          [(rt.OP_EXP, 0), (rt.OP_SUB, 0)])
    f = prog.f_solution[lvl]
    ctx.field_set_bc_expr(f, bc)
    ctx.fill(f, 0.0)
    ctx.apply_bc(f)


def set_problem(prog, workload):
    if workload == "varcoef27":
        set_varcoef_problem(prog)
    else:
        set_benchmark_problem(prog)


def host_threads():
    """CPUs this process may run on (cgroup / affinity aware): the thread count of the CPU arms."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def prepare_openmp():
    """Thread count and placement of the CPU arms, set BEFORE libgomp is loaded: a launcher such as torchrun exports
    OMP_NUM_THREADS=1, which round 1's reference arm inherited at N > 1.  BASELINE.md section 4: one thread per
    available core, OMP_PROC_BIND=close, OMP_PLACES=cores."""
    n = host_threads()
    os.environ["OMP_NUM_THREADS"] = str(n)
    os.environ.setdefault("OMP_PROC_BIND", "close")
    os.environ.setdefault("OMP_PLACES", "cores")
    os.environ.setdefault("OMP_DYNAMIC", "false")
    return n


def load_oracle():
    """The CPU oracle with its OpenMP team sized explicitly; returns (module, threads OpenMP really uses)."""
    want = prepare_openmp()
    from oracle import oracle as og
    og.build()
    used = og.set_num_threads(want)  # omp_set_num_threads + the size of an actual parallel region
    return og, used


def _oracle_solver(og, workload, level, blocks=(1, 1, 1), block_cells=(1, 1, 1), min_level=MIN_LEVEL,
                   domain_hi=(1.0, 1.0, 1.0)):
    """The oracle's solver on the GLOBAL grid of the block decomposition (its arithmetic does not depend on the
    decomposition: shared nodes are computed redundantly from identical inputs)."""
    w = WORKLOADS[workload]
    F = tuple(blocks[d] * block_cells[d] for d in range(3))
    if workload == "varcoef27":
        s = og.Solver(3, F, min_level, level, op=og.OP_VARCOEFF27, order=og.ORDER_TESTING, kappa=KAPPA,
                      smoother=og.SMOOTH_JACOBI, omega=w["omega"], nu_pre=NU_PRE, nu_post=NU_POST, cg_max_it=128,
                      cg_rel_tol=1e-3, domain_hi=domain_hi)
        import numpy as np
        x, y, z = s.node_positions(level)
        X, Y, Z = x[None, None, :], y[None, :, None], z[:, None, None]
        bx, by, bz = X - X * X, Y - Y * Y, Z - Z * Z
        s.rhs(level)[...] = (2.0 * KAPPA) * (((bx * by) + (bx * bz)) + (by * bz))
        u = s.solution(level)
        u[...] = 1.0 - np.exp((-1.0 * KAPPA) * ((bx * by) * bz))
        u[1:-1, 1:-1, 1:-1] = 0.0
        s.sync_slots()
        return s
    s = og.Solver(3, F, min_level, level, op=og.OP_LAPLACE_SCALED, order=og.ORDER_BENCH,
                  smoother=og.SMOOTH_RBGS, omega=OMEGA, nu_pre=NU_PRE, nu_post=NU_POST, cg_max_it=128, cg_rel_tol=1e-3,
                  domain_hi=domain_hi)
    _oracle_problem(s, level)
    return s


def set_benchmark_problem(prog):
    """Solution@finest = Dirichlet data on the boundary, 0 inside; RHS = 0 (Benchmark/Poisson3D)."""
    from exastencils_b200 import runtime as rt
    f = prog.f_solution[prog.max_level]
    bc = [(rt.OP_X, 0), (rt.OP_X, 0), (rt.OP_MUL, 0),
          (rt.OP_CONST, 0.5), (rt.OP_Y, 0), (rt.OP_MUL, 0), (rt.OP_Y, 0), (rt.OP_MUL, 0), (rt.OP_SUB, 0),
          (rt.OP_CONST, 0.5), (rt.OP_Z, 0), (rt.OP_MUL, 0), (rt.OP_Z, 0), (rt.OP_MUL, 0), (rt.OP_SUB, 0)]
    prog.ctx.field_set_bc_expr(f, bc)
    prog.ctx.fill(f, 0.0)
    prog.ctx.apply_bc(f)
    prog.ctx.fill(prog.f_rhs[prog.max_level], 0.0)


def host_mem_available():
    try:
        with open("/proc/meminfo") as fh:
            for line in fh:
                if line.startswith("MemAvailable:"):
                    return int(line.split()[1]) * 1024
    except OSError:
        pass
    return 1 << 62


def cpu_sample_level(decomp, workload, want=None):
    """Per-block finest level of the CPU arms: the GPU arm's block grid at the largest level <= 9 (512^3 cells per
    block) whose global grid fits comfortably into host memory -- a bounded sample of the
    workload; DoF-cycles/s of this bandwidth-bound code does not depend on the level."""
    level = min(decomp["level"], 9 if workload != "varcoef27" else 8)
    if want is not None:
        level = want
    per_node = 8.0 * (3.3 + (27.0 if workload == "varcoef27" else 0.0)) * 8.0 / 7.0
    while level > decomp["min_level"] + 2:
        nodes = 1.0
        for d in range(3):
            nodes *= (decomp["blocks"][d] * decomp["block_cells"][d] << level) + 1
        if nodes * per_node < 0.5 * host_mem_available():
            break
        level -= 1
    return level


def time_oracle(og, s, warmup, steps):
    """(total seconds of `steps` steps, best single step): a step = mgCycle@finest + residual + norm."""
    for _ in range(warmup):
        s.vcycle()
        s.update_residual_norm()
    best, t_all = float("inf"), time.perf_counter()
    for _ in range(steps):
        t0 = time.perf_counter()
        s.vcycle()
        s.update_residual_norm()
        best = min(best, time.perf_counter() - t0)
    return time.perf_counter() - t_all, best


def run_reference(args, decomp):
    """--impl reference: the CPU restatement of the reference's generated C++/OpenMP solver
    (oracle/, kind "port": the generator needs a JVM, absent here) on the host cores, on the global grid of the
    GPU arm's block decomposition at a bounded level."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    og, threads = load_oracle()
    level = cpu_sample_level(decomp, args.workload, args.cpu_level)
    s = _oracle_solver(og, args.workload, level, decomp["blocks"], decomp["block_cells"], decomp["min_level"],
                       decomp["domain_hi"])
    dt, best = time_oracle(og, s, args.warmup, args.steps)
    res = s.update_residual_norm()
    dof = 1.0
    for d in range(3):
        dof *= (decomp["blocks"][d] * decomp["block_cells"][d] << level) - 1
    val = dof * args.steps / dt
    cells = "x".join(str(decomp["blocks"][d] * decomp["block_cells"][d] << level) for d in range(3))
    sample = ("%d V(2,2) cycles + residual norm on the global grid of the %dx%dx%d block decomposition with %d^3 cells "
              "(level %d) per block (%s cells in all), one host, OpenMP with %d threads (OMP_PROC_BIND=%s, OMP_PLACES=%s), "
              "%.1f s; best single step %.1f ms" % (
                  args.steps, decomp["blocks"][0], decomp["blocks"][1], decomp["blocks"][2], 1 << level, level, cells,
                  threads, os.environ.get("OMP_PROC_BIND"), os.environ.get("OMP_PLACES"), dt, 1e3 * best))
    line = {"impl": "reference", "metric": "3D Poisson V-cycle DoF-cycles/s", "value": val, "unit": "DoF-cycles/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(decomp, args.workload),
            "cpu_baseline": {"value": val, "unit": "DoF-cycles/s", "cores": threads, "kind": "port", "sample": sample,
                             "best_of_steps": dof / best, "level_per_block": level, "dof": dof},
            "e2e": {"value": val, "unit": "DoF-cycles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "final_residual": res}
    print(json.dumps(line))


def _oracle_problem(s, level):
    import numpy as np
    x, y, z = s.node_positions(level)
    X, Y, Z = x[None, None, :], y[None, :, None], z[:, None, None]
    u = s.solution(level)
    u[...] = ((X * X) - ((0.5 * Y) * Y)) - ((0.5 * Z) * Z)
    u[1:-1, 1:-1, 1:-1] = 0.0


def decomposition(args):
    """Block grid, cells per block at level 0, level range and domain of the run."""
    n = args.gpus
    if args.workload == "varcoef27":
        blocks = VBLOCKS[n]
        return {"blocks": blocks, "block_cells": (1, 1, 1), "level": args.level if args.level is not None else 8,
                "min_level": MIN_LEVEL, "domain_hi": (1.0, 1.0, 1.0), "scaling": "weak"}
    if args.scaling == "strong":
        blocks = args.blocks or BLOCKS_X_FIRST[n]
        if n == 1:
            return {"blocks": (1, 1, 1), "block_cells": (1, 1, 1), "level": args.level if args.level is not None else 10,
                    "min_level": MIN_LEVEL, "domain_hi": (1.0, 1.0, 1.0), "scaling": "strong"}
        level = args.level if args.level is not None else STRONG_LEVEL
        bc = tuple(2 // blocks[d] if blocks[d] <= 2 else 0 for d in range(3))
        if 0 in bc:
            raise SystemExit("--scaling strong: at most two blocks per dimension")
        return {"blocks": blocks, "block_cells": bc, "level": level, "min_level": STRONG_MIN_LEVEL,
                "domain_hi": (1.0, 1.0, 1.0), "scaling": "strong"}
    blocks = args.blocks or BLOCKS[n]
    level = args.level if args.level is not None else (9 if n == 1 else 10)
    return {"blocks": blocks, "block_cells": (1, 1, 1), "level": level, "min_level": MIN_LEVEL,
            "domain_hi": tuple(float(b) for b in blocks), "scaling": "weak"}


def workload_config(decomp, workload="poisson7"):
    level, blocks, bc = decomp["level"], decomp["blocks"], decomp["block_cells"]
    per_gpu = [bc[d] << level for d in range(3)]
    total = [blocks[d] * per_gpu[d] for d in range(3)]
    cells_per_gpu = per_gpu[0] * per_gpu[1] * per_gpu[2]
    nodes_per_gpu = (per_gpu[0] + 1) * (per_gpu[1] + 1) * (per_gpu[2] + 1)
    if workload == "varcoef27":
        return {"workload": "3D 27-point variable-coefficient Poisson (stencil field, coefficient exp(10 (x-x^2)(y-y^2)"
                            "(z-z^2))), V(2,2) damped Jacobi omega=0.85 with pairs of sweeps fused into one pass, CG "
                            "coarse solve, %d^3 cells per GPU (%d^3 in all), levels %d..%d" % (
                                1 << level, blocks[0] << level, MIN_LEVEL, level),
                "blocks": list(blocks), "cells_per_gpu": cells_per_gpu, "smoother": "jacobi", "nu_pre": NU_PRE,
                "nu_post": NU_POST, "l2_policy": "inputs larger than L2 (27 coefficient components x %.2f GB on the "
                "finest level)" % (8.0 * nodes_per_gpu / 1e9)}
    name = ("BASELINE configs[4], strong scaling" if decomp["scaling"] == "strong" else
            "BASELINE configs[1]" if blocks == (1, 1, 1) and level == 9 else
            "BASELINE configs[4], weak scaling" if level == 10 else "weak scaling")
    return {"workload": "Benchmark/Poisson3D 7-point constant-coefficient, V(2,2) RBGS omega=0.8, CG coarse solve, "
                        "%dx%dx%d cells per GPU, %dx%dx%d in all, levels %d..%d (%s)" % (
                            per_gpu[0], per_gpu[1], per_gpu[2], total[0], total[1], total[2], decomp["min_level"], level,
                            name),
            "blocks": list(blocks), "block_cells_level0": list(bc), "level": level, "cells_per_gpu": cells_per_gpu,
            "smoother": "rbgs", "nu_pre": NU_PRE, "nu_post": NU_POST,
            "l2_policy": "inputs larger than L2 (3 fields x %.2f GB on the finest level)" % (8.0 * nodes_per_gpu / 1e9)}


def cpu_baseline(args, decomp):
    """Bounded sample of the same workload on the host cores with the oracle (kind "port"): best of `cpu_cycles`
    single steps after one warm-up step."""
    og, threads = load_oracle()
    level = cpu_sample_level(decomp, args.workload, args.cpu_level)
    s = _oracle_solver(og, args.workload, level, decomp["blocks"], decomp["block_cells"], decomp["min_level"],
                       decomp["domain_hi"])
    reps = max(3, args.cpu_cycles)
    dt, best = time_oracle(og, s, 1, reps)
    s.close()
    dof = 1.0
    for d in range(3):
        dof *= (decomp["blocks"][d] * decomp["block_cells"][d] << level) - 1
    return {"value": dof / best, "unit": "DoF-cycles/s", "cores": threads, "kind": "port",
            "sample": "best of %d single steps (V(2,2) cycle + residual norm) at %d^3 cells per block, %d block(s) "
                      "(OpenMP, %d threads really used, OMP_PROC_BIND=%s, OMP_PLACES=%s; mean %.1f ms, best %.1f ms)" % (
                          reps, 1 << level, decomp["blocks"][0] * decomp["blocks"][1] * decomp["blocks"][2], threads,
                          os.environ.get("OMP_PROC_BIND"), os.environ.get("OMP_PLACES"), 1e3 * dt / reps, 1e3 * best),
            "mean_value": dof * reps / dt}



# ---- BASELINE configs[0] and configs[3] (parity-test cases of the bench contract; measured the same way) -----------
AUX = ("poisson2d", "optflow")


def _aux_config(workload, level):
    if workload == "poisson2d":
        return {"workload": "Examples/Poisson 2D 5-point FD, V(3,3) RBGS omega=0.8, CG coarse solve, max level 10 "
                            "(1025^2 nodes = 4x4 fragments of level 8, Testing/Poisson/2D_FD_Poisson_fromL4)",
                "cells_per_gpu": 1024 ** 2, "smoother": "rbgs", "nu_pre": 3, "nu_post": 3,
                "l2_policy": "none: 8.4 MB per field, the whole problem lives in L2 (latency-bound cycle)"}
    n = 1 << level
    return {"workload": "Examples/OpticalFlow 2D coupled two-component system, V(3,3) red-black 2x2 block smoother, "
                        "coupled CG, %d^2 cells, levels 2..%d" % (n, level),
            "cells_per_gpu": n * n, "smoother": "block2x2-rb", "nu_pre": 3, "nu_post": 3,
            "l2_policy": "inputs larger than L2 (9 fields x %.2f GB on the finest level)" % (8.0 * n * n / 1e9)}


def _aux_oracle(workload, level):
    """(solver, step function, unknowns) of the CPU oracle for the bounded sample."""
    import numpy as np
    og, _ = load_oracle()
    if workload == "poisson2d":
        s = og.Solver(2, (4, 4), 0, 8, op=og.OP_LAPLACE_SCALED, order=og.ORDER_BENCH, smoother=og.SMOOTH_RBGS, omega=0.8,
                      nu_pre=3, nu_post=3, cg_max_it=128, cg_rel_tol=1e-3)
        x, y, _ = s.node_positions(8)
        X, Y = x[None, None, :], y[None, :, None]
        u = s.solution(8)
        u[...] = np.broadcast_to(np.cos(np.pi * X) - np.sin(2.0 * np.pi * Y), u.shape)
        u[:, 1:-1, 1:-1] = 0.0
        s.rhs(8)[...] = np.broadcast_to(np.pi ** 2 * np.cos(np.pi * X) - 4.0 * np.pi ** 2 * np.sin(2.0 * np.pi * Y), u.shape)
        return s, 1023.0 ** 2
    from oracle import optflow as oof
    s = oof.Solver((1, 1), 2, level)
    s.init_images()
    s.init_fields()
    return s, 2.0 * float(1 << level) ** 2


def _aux_cpu(workload, level, reps, warm=1):
    s, unknowns = _aux_oracle(workload, level)
    _, ncores = load_oracle()
    for _ in range(warm):
        s.vcycle()
        s.update_residual_norm()
    t0 = time.perf_counter()
    for _ in range(reps):
        s.vcycle()
        s.update_residual_norm()
    dt = time.perf_counter() - t0
    s.close()
    what = "1025^2 nodes" if workload == "poisson2d" else "%d^2 cells" % (1 << level)
    return {"value": unknowns * reps / dt, "unit": "DoF-cycles/s", "cores": ncores, "kind": "port",
            "sample": "%d V(3,3) cycles + residual norm at %s (OpenMP, %d threads, %.1f s)" % (reps, what, ncores, dt)}, dt


def run_aux(args):
    """--workload poisson2d | optflow on one GPU: same JSON line as the main workload."""
    level = args.level if args.level is not None else (8 if args.workload == "poisson2d" else 13)
    cpu_level = args.cpu_level if args.cpu_level is not None else (8 if args.workload == "poisson2d" else min(level, 12))
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    multi = args.workload == "optflow" and args.gpus > 1 and args.impl != "reference"
    if rank != 0 and not multi:
        return
    if args.gpus != 1 and not multi and args.impl != "reference":
        raise SystemExit("--workload %s runs on one GPU" % args.workload)
    if multi and world != args.gpus:
        raise SystemExit("--gpus %d needs torch.distributed.run with %d ranks" % (args.gpus, args.gpus))
    if args.impl == "reference":
        cb, dt = _aux_cpu(args.workload, cpu_level, args.steps, args.warmup)
        line = {"impl": "reference", "metric": "V-cycle DoF-cycles/s", "value": cb["value"], "unit": "DoF-cycles/s",
                "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": _aux_config(args.workload, level), "cpu_baseline": cb,
                "e2e": {"value": cb["value"], "unit": "DoF-cycles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return
    import numpy as np
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libexamg has no CPU path")
    torch.cuda.set_device(local_rank)
    dist = None
    if multi:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    of_blocks = {1: (1, 1), 2: (2, 1), 4: (2, 2), 8: (4, 2)}.get(args.gpus, (args.gpus, 1))
    stream = torch.cuda.Stream()
    peak, peak_kind = peaks()
    if args.workload == "poisson2d":
        from exastencils_b200 import MultigridProgram
        prog = MultigridProgram(2, (4, 4), 0, 8, op="scaled", order="bench", smoother="rbgs", omega=0.8, nu_pre=3,
                                nu_post=3, restrict_scale=1.0, cg_max_it=128, cg_rel_tol=1e-3, stream=stream.cuda_stream)
        ctx = prog.ctx
        x, y, _ = prog.node_positions(8)
        X, Y = x[None, None, :], y[None, :, None]
        u = prog.solution(8)
        u[...] = np.broadcast_to(np.cos(np.pi * X) - np.sin(2.0 * np.pi * Y), u.shape)
        u[:, 1:-1, 1:-1] = 0.0
        rhs = np.ascontiguousarray(np.broadcast_to(np.pi ** 2 * np.cos(np.pi * X) - 4.0 * np.pi ** 2 * np.sin(2.0 * np.pi * Y), u.shape))
        prog.rhs(8)[...] = rhs
        prog.commit()
        unknowns = 1023.0 ** 2
        inputs = [(prog.f_rhs[8], torch.from_numpy(rhs.copy()).pin_memory())]
        prepare = lambda: None
        A, fu, ff = prog.stencil[8], prog.f_solution[8], prog.f_rhs[8]
        kernel = lambda: ctx.smooth_rbgs(A, fu, ff, 0.8, 3)
        kname, kbytes = "RBGS smoother, 3 sweeps at 1025^2 (generic colour kernels, L2-resident)", 3 * 24.0 * unknowns
    else:
        from exastencils_b200 import OpticalFlowProgram
        from exastencils_b200 import optflow as gof
        # several GPUs: one block of 2^level x 2^level cells per GPU (weak scaling), blocks 2x1 / 2x2 / 4x2
        prog = OpticalFlowProgram((1, 1), 2, level, nu_pre=3, nu_post=3, stream=stream.cuda_stream, num_blocks=of_blocks,
                                  rank=rank, device=local_rank, dist=dist)
        ctx = prog.ctx
        nx, ny = prog.cells()
        x = (np.arange(nx, dtype=np.float64) + (rank % of_blocks[0]) * nx + 0.5) * (1.0 / (nx * of_blocks[0]))
        y = (np.arange(ny, dtype=np.float64) + (rank // of_blocks[0]) * ny + 0.5) * (1.0 / (ny * of_blocks[1]))
        img0 = np.ascontiguousarray(np.sin(x)[None, :] + np.cos(y)[:, None])      # initImgs@finest, this block's part
        img1 = np.ascontiguousarray(np.sin(x - 0.01)[None, :] + np.cos(y + 0.01)[:, None])
        prog.set_images(img0, img1)
        unknowns = 2.0 * nx * ny * args.gpus
        inputs = [(prog.field(gof.IMG0), torch.from_numpy(img0).pin_memory()),
                  (prog.field(gof.IMG1), torch.from_numpy(img1).pin_memory())]
        from exastencils_b200.runtime import check as rt_check
        prepare = lambda: rt_check(ctx.L.examg_optflow_init_fields(prog._h))
        sys_h = level - 2  # systems are created per level in order
        f = {w: prog.field(w) for w in range(9)}
        kernel = lambda: ctx.system2_smooth_rb(sys_h, f[gof.U], f[gof.V], f[gof.RHS_U], f[gof.RHS_V], 1.0, 3)
        kname, kbytes = "red-black 2x2 block smoother, 3 sweeps on the finest level", 3 * 80.0 * nx * ny

    def sync_all():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_ms(v):
        if dist is None:
            return v
        t = torch.tensor([v], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    with torch.cuda.stream(stream):
        sampler = ClockSampler(0)
        sampler.start()
        for _ in range(args.warmup):
            prog.cycles(1)
        sync_all()
        l0 = ctx.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            res = prog.cycles(1)
        e1.record(stream)
        sync_all()
        ms = max_ms(e0.elapsed_time(e1))
        launches = ctx.launch_count() - l0
        # end to end: every step uploads its input from pinned host memory, (re)initialises, cycles, reads the norm back
        h2d = sum(t.numel() * 8 for _, t in inputs)
        for _ in range(args.warmup):
            for fld, t in inputs:
                ctx.upload_async(fld, t.data_ptr())
            prepare()
            prog.cycles(1)
        sync_all()
        e0.record(stream)
        for _ in range(args.steps):
            for fld, t in inputs:
                ctx.upload_async(fld, t.data_ptr())
            prepare()
            res_e2e = prog.cycles(1)
        e1.record(stream)
        sync_all()
        ms_e2e = max_ms(e0.elapsed_time(e1))
        clocks = sampler.stop()
        reps = max(5, args.steps)
        for _ in range(3):
            kernel()
        torch.cuda.synchronize()
        l1 = ctx.launch_count()
        e0.record(stream)
        for _ in range(reps):
            kernel()
        e1.record(stream)
        torch.cuda.synchronize()
        ms_k = e0.elapsed_time(e1) / reps
        k_launches = (ctx.launch_count() - l1) // reps
    achieved = kbytes / (ms_k * 1e-3) / 1e9
    if rank != 0:
        prog.close()
        dist.destroy_process_group()
        return
    cb = None
    if not args.no_cpu_baseline and not multi:
        cb, _ = _aux_cpu(args.workload, cpu_level, args.cpu_cycles)
    line = {"metric": "V-cycle DoF-cycles/s", "value": unknowns * args.steps / (ms * 1e-3), "unit": "DoF-cycles/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": _aux_config(args.workload, level),
            "e2e": {"value": unknowns * args.steps / (ms_e2e * 1e-3), "unit": "DoF-cycles/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": 8, "ms_per_step": ms_e2e / args.steps,
                    "what": "per step: input H2D from pinned memory, (re)initialisation, one cycle + residual, norm D2H"},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": "%s (%d launch(es))" % (kname, k_launches), "achieved": achieved,
                         "peak": peak, "peak_kind": peak_kind + " (MEASURED_PEAKS.json hbm_gbs, burst)", "unit": "GB/s",
                         "frac": achieved / peak, "traffic": None, "algorithmic_bytes_per_launch_group": kbytes, "ms": ms_k},
            "cpu_baseline": cb, "final_residual": res, "device_bytes": ctx.device_bytes()}
    if multi:
        line["config"]["blocks"] = list(of_blocks)
        line["config"]["workload"] += "; one such block per GPU, blocks %d x %d" % of_blocks
        line["e2e"]["h2d_bytes_per_step"] = h2d * args.gpus
    print(json.dumps(line))
    prog.close()
    if dist is not None:
        dist.destroy_process_group()


def parity_vs_one_block(args, decomp, wl, dist, rank, local_rank, world, stream):
    """Cross-N parity, outside every timed region: the N-block program at a reduced level against the SAME program on
    one block that covers the global grid, run by rank 0 on its GPU (seeded random Solution and RHS, 2 cycles).  The
    arithmetic contract makes shared nodes and sums independent of the decomposition, so the residual
    norms must agree to 1e-12 relative (they do to the last bit) and the fields bit for bit."""
    import numpy as np
    import torch
    from exastencils_b200 import MultigridProgram
    from exastencils_b200.distributed import block_slices, wire_peers
    L = min(decomp["level"], PARITY_LEVEL)
    blocks, bc, lo = decomp["blocks"], decomp["block_cells"], decomp["min_level"]
    kw = dict(op=wl["op"], order=wl["order"], smoother=wl["smoother"], omega=wl["omega"], kappa=wl["kappa"],
              nu_pre=NU_PRE, nu_post=NU_POST, cg_max_it=128, cg_rel_tol=1e-3, domain_hi=decomp["domain_hi"])
    prog = MultigridProgram(3, bc, lo, L, num_blocks=blocks, rank=rank, device=local_rank,
                            stream=stream.cuda_stream, **kw)
    wire_peers(prog.ctx, dist, device="cuda:%d" % local_rank)
    gshape = tuple((blocks[d] * bc[d] << L) + 1 for d in (2, 1, 0))
    rng = np.random.default_rng(20240607)
    u0 = rng.uniform(-1.0, 1.0, size=gshape)
    f0 = rng.uniform(-1.0, 1.0, size=gshape)
    sl = block_slices(rank, blocks, gshape)
    prog.solution(L)[...] = u0[sl]
    prog.rhs(L)[...] = f0[sl]
    prog.commit()
    cycles = 2
    with torch.cuda.stream(stream):
        norms = [prog.update_residual_norm()]
        for _ in range(cycles):
            prog.vcycle()
            norms.append(prog.update_residual_norm())
        mine = torch.from_numpy(np.ascontiguousarray(prog.fetch_solution())).cuda()
    torch.cuda.synchronize()
    # blocks may differ in shape only through --blocks with ragged cells, which the bench does not offer
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)
    prog.close()
    out = None
    if rank == 0:
        one = MultigridProgram(3, tuple(blocks[d] * bc[d] for d in range(3)), lo, L, device=local_rank,
                               stream=stream.cuda_stream, **kw)
        one.solution(L)[...] = u0
        one.rhs(L)[...] = f0
        one.commit()
        with torch.cuda.stream(stream):
            n1 = [one.update_residual_norm()]
            for _ in range(cycles):
                one.vcycle()
                n1.append(one.update_residual_norm())
            exp = one.fetch_solution()
        one.close()
        got = np.empty(gshape)
        for r in range(world):
            got[block_slices(r, blocks, gshape)] = parts[r].cpu().numpy()
        rel = max(abs(a - b) / abs(b) for a, b in zip(norms, n1))
        nbad = int(np.count_nonzero(got != exp))
        out = {"ok": bool(rel <= 1e-12 and nbad == 0), "level_per_block": L, "cycles": cycles,
               "max_rel_norm_diff": rel, "nodes_differing": nbad, "nodes": int(got.size),
               "norms": norms, "norms_one_block": n1,
               "what": "N-block program vs the same program on one block covering the global grid (rank 0's GPU), "
                       "seeded random Solution/RHS; tolerance 1e-12 relative on norms, fields bit-identical"}
    return out


def measure_steps(prog, ctx, stream, steps, warmup, barrier):
    """warmup untimed steps, then `steps` timed ones (CUDA events on the library's stream)."""
    import torch
    for _ in range(warmup):
        prog.cycles(1)
    barrier()
    l0 = ctx.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    res = None
    for _ in range(steps):
        res = prog.cycles(1)
    e1.record(stream)
    barrier()
    return e0.elapsed_time(e1), ctx.launch_count() - l0, res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="examg", choices=["examg", "reference"])
    ap.add_argument("--workload", default="poisson7", choices=sorted(WORKLOADS) + list(AUX))
    ap.add_argument("--level", type=int, default=None, help="finest level per GPU (9 = 512^3 cells; default: 9 on one "
                    "GPU, 10 on several = BASELINE configs[4]; varcoef27: 8)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"], help="strong: 1024^3 cells in all "
                    "(x-first block grid, 512 cells per split dimension)")
    ap.add_argument("--blocks", default=None, help="domain_rect_numBlocks as x,y,z (default: BLOCKS / BLOCKS_X_FIRST)")
    ap.add_argument("--cpu-level", type=int, default=None)
    ap.add_argument("--cpu-cycles", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the cross-N parity check (N > 1)")
    ap.add_argument("--no-config5", action="store_true", help="N = 1: skip the 1024^3 sub-record")
    args = ap.parse_args()
    if args.gpus not in BLOCKS:
        raise SystemExit("--gpus must be 1, 2, 4 or 8")
    if args.workload in AUX:
        return run_aux(args)
    if args.blocks:
        args.blocks = tuple(int(v) for v in args.blocks.split(","))
        if len(args.blocks) != 3 or args.blocks[0] * args.blocks[1] * args.blocks[2] != args.gpus:
            raise SystemExit("--blocks x,y,z must multiply to --gpus")
    wl = WORKLOADS[args.workload]
    if args.workload == "varcoef27" and args.gpus not in VBLOCKS:
        raise SystemExit("--workload varcoef27 runs on 1 GPU or on 8 (2x2x2 blocks of the unit cube: 1024^3 cells at "
                         "level 9 per GPU)")
    decomp = decomposition(args)
    args.scaling = decomp["scaling"]
    blocks, level = decomp["blocks"], decomp["level"]
    if args.impl == "reference":
        return run_reference(args, decomp)

    import torch
    from exastencils_b200 import MultigridProgram
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE %d (launch with torch.distributed.run)" % (args.gpus, world))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libexamg has no CPU path")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    stream = torch.cuda.Stream()

    def make_program(lvl, dc):
        p = MultigridProgram(3, dc["block_cells"], dc["min_level"], lvl, op=wl["op"], order=wl["order"],
                             smoother=wl["smoother"], omega=wl["omega"], kappa=wl["kappa"], nu_pre=NU_PRE, nu_post=NU_POST,
                             cg_max_it=128, cg_rel_tol=1e-3, num_blocks=dc["blocks"], rank=rank, device=local_rank,
                             stream=stream.cuda_stream, domain_hi=dc["domain_hi"])
        if world > 1:
            from exastencils_b200.distributed import wire_peers
            wire_peers(p.ctx, dist, device="cuda:%d" % local_rank)
        set_problem(p, args.workload)
        return p

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if dist is None:
            return ms
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- cross-N parity (outside the timed regions) ---------------------------------
    parity = None
    if world > 1 and not args.no_parity and args.workload == "poisson7":
        parity = parity_vs_one_block(args, decomp, wl, dist, rank, local_rank, world, stream)
        barrier()

    prog = make_program(level, decomp)
    ctx = prog.ctx
    dev_bytes = ctx.device_bytes()
    dof = 1.0
    for d in range(3):
        dof *= (blocks[d] * decomp["block_cells"][d] << level) - 1

    # ---- device-resident throughput ------------------------------------------------
    with torch.cuda.stream(stream):
        sampler = ClockSampler(local_rank)
        if rank == 0:
            sampler.start()  # samples cover the warm-up and both timed regions (nvidia-smi needs ~100 ms per sample)
        ms, launches, res = measure_steps(prog, ctx, stream, args.steps, args.warmup, barrier)
    ms = max_over_ranks(ms)
    value = dof * args.steps / (ms * 1e-3)

    # ---- end to end through the C ABI with host buffers -----------------------------
    # every step: H2D of the step's input (RHS@finest, pinned), one cycle + residual norm, D2H of the norm
    shape = ctx.shape(level)
    rhs_host = torch.zeros(shape, dtype=torch.float64).pin_memory()
    h2d = rhs_host.numel() * 8
    set_problem(prog, args.workload)
    if args.workload == "varcoef27":  # the step's input: the RHS the device just assembled, staged in pinned memory
        rhs_host.copy_(torch.from_numpy(ctx.download(prog.f_rhs[level], level)))
    with torch.cuda.stream(stream):
        # software pipeline through the ABI's upload_begin/_finish: the PCIe transfer of step i+1's input runs on
        # the library's copy stream while step i computes; every step still moves its own input
        for _ in range(args.warmup):
            ctx.upload_begin(prog.f_rhs[level], rhs_host.data_ptr())
            ctx.upload_finish()
            prog.cycles(1)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        ctx.upload_begin(prog.f_rhs[level], rhs_host.data_ptr())
        for i in range(args.steps):
            ctx.upload_finish()
            if i + 1 < args.steps:
                ctx.upload_begin(prog.f_rhs[level], rhs_host.data_ptr())
            res_e2e = prog.cycles(1)  # synchronises: the norm is read back on the host
        e1.record(stream)
        barrier()
        ms_e2e = e0.elapsed_time(e1)
    ms_e2e = max_over_ranks(ms_e2e)
    clocks = sampler.stop() if rank == 0 else None
    e2e = {"value": dof * args.steps / (ms_e2e * 1e-3), "unit": "DoF-cycles/s", "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": 16, "ms_per_step": ms_e2e / args.steps,
           "h2d_gbs_per_gpu": h2d / (ms_e2e / args.steps * 1e-3) / 1e9,
           "what": "per step: RHS@finest H2D from pinned memory (upload_begin/_finish: the transfer of the next "
                   "step's input overlaps this step's kernels), mgCycle + residual, norm D2H",
           "bound": "host-to-device transfer: every rank uploads its block's RHS every step; the step time is the "
                    "PCIe / host-memory time of that upload (h2d_gbs_per_gpu), all GPUs of the box share one host"}
    del rhs_host

    # ---- roofline of the dominant kernel (RBGS sweeps on the finest level), timed alone ----
    peak, peak_kind = peaks()
    A, u, f = prog.stencil[level], prog.f_solution[level], prog.f_rhs[level]
    n_loc = ctx.dims(level)
    pts = float(n_loc[0] - 2) * (n_loc[1] - 2) * (n_loc[2] - 2) if world == 1 else float(n_loc[0]) * n_loc[1] * n_loc[2]
    reps = max(5, args.steps)
    if args.workload == "varcoef27":
        smooth = lambda: ctx.smooth_jacobi(A, u, f, wl["omega"], NU_PRE)
    else:
        smooth = lambda: ctx.smooth_rbgs(A, u, f, OMEGA, NU_PRE)
    with torch.cuda.stream(stream):
        for _ in range(3):
            smooth()
        barrier()
        l1 = ctx.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            smooth()
        e1.record(stream)
        barrier()
        ms_k = e0.elapsed_time(e1) / reps
        k_launches = (ctx.launch_count() - l1) // reps
        e0.record(stream)
        for _ in range(reps):
            ctx.residual(A, u, f, prog.f_residual[level])
        e1.record(stream)
        barrier()
        ms_r = e0.elapsed_time(e1) / reps
    alg_bytes = wl["bytes_sweep"] * NU_PRE * pts
    achieved = alg_bytes / (ms_k * 1e-3) / 1e9
    kname = "damped Jacobi on the 27-point stencil field" if args.workload == "varcoef27" else "RBGS smoother"
    roofline = {"bound": "hbm", "kernel": "%s, %d sweeps on the finest level (%d launch(es)%s)" % (
                    kname, NU_PRE, k_launches, ", halo exchange included" if world > 1 else ""),
                "achieved": achieved, "peak": peak, "peak_kind": peak_kind + " (MEASURED_PEAKS.json hbm_gbs, burst)",
                "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "algorithmic_bytes_per_launch_group": alg_bytes, "ms": ms_k,
                "fused_minimum_bytes": alg_bytes / NU_PRE,
                "frac_of_fused_minimum": alg_bytes / NU_PRE / (ms_k * 1e-3) / 1e9 / peak,
                "residual": {"achieved": wl["bytes_residual"] * pts / (ms_r * 1e-3) / 1e9, "ms": ms_r,
                             "frac": wl["bytes_residual"] * pts / (ms_r * 1e-3) / 1e9 / peak}}
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(traffic_file) and level == (8 if args.workload == "varcoef27" else 9) and world == 1:
        with open(traffic_file) as fh:
            tj = json.load(fh)
        roofline["traffic"] = tj.get("jacv27_bytes_per_launch_group" if args.workload == "varcoef27"
                                     else "rbgs_bytes_per_launch_group")
        roofline["traffic_source"] = tj.get("source", "profiles/traffic.json (ncu --set full capture committed with the "
                                                      "kernel; not measured in this run)")

    # ---- N = 1: the same step at 1024^3 cells, the N = 1 point of BASELINE configs[4] ----
    config5 = None
    if (world == 1 and args.workload == "poisson7" and level == 9 and args.scaling == "weak" and not args.no_config5):
        prog.close()
        prog = None
        torch.cuda.empty_cache()
        dc5 = dict(decomp, level=10)
        p5 = make_program(10, dc5)
        with torch.cuda.stream(stream):
            k5 = max(5, min(args.steps, 10))
            ms5, _, res5 = measure_steps(p5, p5.ctx, stream, k5, max(3, min(args.warmup, 5)), barrier)
        dof5 = float((1 << 10) - 1) ** 3
        config5 = {"value": dof5 * k5 / (ms5 * 1e-3), "unit": "DoF-cycles/s", "ms_per_step": ms5 / k5, "steps": k5,
                   "config": workload_config(dc5, args.workload), "final_residual": res5,
                   "what": "weak-scaling efficiency of the N > 1 lines (1024^3 cells per GPU) = value(N) / (N x this value)"}
        dev_bytes5 = p5.ctx.device_bytes()
        config5["device_bytes"] = dev_bytes5
        p5.close()

    cb = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cb = cpu_baseline(args, decomp)

    if rank == 0:
        line = {"metric": "3D Poisson V-cycle DoF-cycles/s", "value": value, "unit": "DoF-cycles/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": workload_config(decomp, args.workload), "e2e": e2e,
                "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cb,
                "final_residual": res, "device_bytes": dev_bytes,
                "parity_vs_n1": parity["ok"] if parity else None, "parity": parity, "config5_n1": config5}
        print(json.dumps(line))
    if prog is not None:
        prog.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
