"""Whole-solver parity on the GPU: the reference's regression programs run through
libexamg must (a) print the reference's golden stdout (4 s.f., tests/golden/) and
(b) match the CPU oracle -- Solution fields bit-for-bit after every V-cycle,
residual norms within 1e-12 relative (the tolerance BASELINE.json's north_star states).
Both host modes are covered: the library cycle (CUDA graph) and the call-by-call
sequence the generator back-end emits.
"""
import numpy as np
import pytest

import programs
from oracle_driver import make_oracle_solver

pytestmark = pytest.mark.gpu

RTOL_NORM = 1e-12


def make_gpu_solver(cfg, mode="library", **override):
    from exastencils_b200 import MultigridProgram
    c = dict(cfg)
    c.update(override)
    return MultigridProgram(c["dim"], tuple(c["F"]), c["min_level"], c["max_level"], op=c["op"], order=c["order"],
                            kappa=c.get("kappa", 0.0), smoother=c["smoother"], omega=c["omega"], nu_pre=c["nu_pre"],
                            nu_post=c["nu_post"], restrict_scale=c["restrict_scale"], cg_max_it=c["cg_max_it"],
                            cg_rel_tol=c["cg_rel_tol"], mode=mode, periodic=c.get("periodic", (0, 0, 0)))


def _setup_pair(prog, mode, **override):
    ref = make_oracle_solver(prog.cfg, **override)
    gpu = make_gpu_solver(prog.cfg, mode, **override)
    prog.setup(ref)
    ref.sync_slots()
    prog.setup(gpu)
    if prog.cfg["op"] in ("varcoeff", "varcoeff27"):  # identical operator data on both sides
        for lvl in range(ref.min_level, ref.max_level + 1):
            gpu.coeff(lvl)[...] = ref.coeff(lvl).reshape(gpu.coeff(lvl).shape)
    gpu.sync_slots()
    return ref, gpu


@pytest.mark.parametrize("mode", ["library", "calls"])
@pytest.mark.parametrize("prog", [programs.POISSON_2D, programs.POISSON_2D_FV, programs.SISC_2D_CONST, programs.SISC_2D_VAR,
                                  programs.SMOOTHERS_RBGS, programs.SMOOTHERS_JAC, programs.COMMBASIC_PURE,
                                  programs.BC_2D_POLY,
                                  programs.BC_2D_TRIG, programs.COMMBASIC_2D, programs.FMG_2D_CONST,
                                  programs.FMG_2D_POLY, programs.BC_2D_PERIODIC], ids=lambda p: p.name)
def test_golden_stdout(prog, mode, oracle):
    """Same printed history as the reference's .results file."""
    if mode == "calls" and prog.dim == 3:
        pytest.skip("call-by-call mode is covered in 2D and by test_oracle_parity")
    gpu = make_gpu_solver(prog.cfg, mode)
    prog.setup(gpu)
    if prog.cfg["op"] in ("varcoeff", "varcoeff27"):
        ref = make_oracle_solver(prog.cfg)
        for lvl in range(ref.min_level, ref.max_level + 1):
            gpu.coeff(lvl)[...] = ref.coeff(lvl).reshape(gpu.coeff(lvl).shape)
        ref.close()
    gpu.sync_slots()
    out = programs.run_solve(prog, gpu, oracle.print_reduced_prec, oracle.max_abs_diff)
    gpu.close()
    assert out == prog.expected()


@pytest.mark.parametrize("prog", [programs.SISC_3D_CONST, programs.SISC_3D_VAR, programs.BC_3D_POLY,
                                  programs.BC_3D_TRIG, programs.OPTS_SEQ, programs.OPTS_PAR, programs.FMG_3D_TRIG,
                                  programs.FMG_3D_VAR, programs.BC_3D_PERIODIC], ids=lambda p: p.name)
def test_golden_stdout_3d_sisc(prog, oracle):
    """256^3-cell programs: Testing/SISC (the reference's own CUDA target is tested on these) and Testing/BC
    (polynomial and sin*sin*sinh Dirichlet data); Testing/Opts seq (65^3) and par (513^3 nodes: three Jacobi sweeps
    per pass through k_jacw<3>) -- the history every optimisation variant and the temporally blocked program of the
    reference must print."""
    gpu = make_gpu_solver(prog.cfg, "library")
    prog.setup(gpu)
    if prog.cfg["op"] in ("varcoeff", "varcoeff27"):
        ref = make_oracle_solver(prog.cfg)
        for lvl in range(ref.min_level, ref.max_level + 1):
            gpu.coeff(lvl)[...] = ref.coeff(lvl).reshape(gpu.coeff(lvl).shape)
        ref.close()
    gpu.sync_slots()
    out = programs.run_solve(prog, gpu, oracle.print_reduced_prec, oracle.max_abs_diff)
    gpu.close()
    assert out == prog.expected()


SMALL = dict(max_level=4)


@pytest.mark.parametrize("mode", ["library", "calls"])
@pytest.mark.parametrize("prog,override", [
    (programs.POISSON_2D, dict(max_level=6)),
    (programs.SISC_2D_VAR, dict(max_level=6)),
    (programs.SMOOTHERS_RBGS, dict(max_level=3)),
    (programs.SMOOTHERS_JAC, dict(max_level=3)),
    (programs.SISC_3D_CONST, dict(max_level=3)),
    (programs.SISC_3D_VAR, dict(max_level=3)),
    (programs.SMOOTHERS_RBGS, dict(max_level=3, nu_pre=2, nu_post=2, omega=0.8)),   # V(2,2), north-star smoother
    (programs.SMOOTHERS_JAC, dict(max_level=3, nu_pre=2, nu_post=1)),                # odd sweep count: slot parity flips
    (programs.BC_2D_PERIODIC, dict(max_level=5)),      # domain_rect_periodic_y: self-exchange, wrap inside the device CG
    (programs.BC_3D_PERIODIC, dict(max_level=3)),      # periodic in y and z
    (programs.CONFIG3_VARCOEFF27, dict(max_level=6)),   # 27-point stencil field, V(2,2): pairs of sweeps fused on 33^3, 65^3
    (programs.CONFIG3_VARCOEFF27, dict(max_level=5, nu_pre=3, nu_post=3)),           # 2 fused + 1 generic sweep
], ids=lambda v: v.name if hasattr(v, "name") else "-".join("%s%s" % kv for kv in v.items()))
def test_oracle_parity(prog, override, mode, oracle):
    ref, gpu = _setup_pair(prog, mode, **override)
    lvl = ref.max_level
    r_ref, r_gpu = ref.update_residual_norm(), gpu.update_residual_norm()
    assert abs(r_gpu - r_ref) <= RTOL_NORM * r_ref
    assert np.array_equal(gpu.fetch_residual(), ref.residual(lvl).reshape(gpu.ctx.shape(lvl)))
    for cyc in range(4):
        ref.vcycle()
        gpu.vcycle()
        assert gpu.last_cg_iters() == ref.last_cg_iters(), "CG iteration count, cycle %d" % cyc
        assert np.array_equal(gpu.fetch_solution(), ref.solution(lvl).reshape(gpu.ctx.shape(lvl))), "cycle %d" % cyc
        r_ref, r_gpu = ref.update_residual_norm(), gpu.update_residual_norm()
        assert abs(r_gpu - r_ref) <= RTOL_NORM * r_ref
    # coarse levels too (restriction / coarse solution of the last cycle)
    for l in range(ref.min_level, lvl):
        assert np.array_equal(gpu.fetch_rhs(l), ref.rhs(l).reshape(gpu.ctx.shape(l))), "RHS@%d" % l
        assert np.array_equal(gpu.fetch_solution(l), ref.solution(l).reshape(gpu.ctx.shape(l))), "Solution@%d" % l
    gpu.close()
    ref.close()


@pytest.mark.parametrize("dim,ncol", [(2, 2), (2, 3), (3, 2), (3, 3)])
def test_coloured_smoother_in_the_library_cycle(dim, ncol, oracle):
    """solver_smoother_coloring "4-way" / "9-way" / "8-way" / "27-way" inside the library cycle (EXAMG_SMOOTH_COLORED_n,
    reached through examg_cycle_desc_from_options): the CUDA-graph cycle equals the same statements issued one by one
    (examg_smooth_colored, residual, restrict, set, recursion, prolong_add; CG at the bottom) bit for bit, and the
    pieces are pinned against the oracle by tests/test_gpu_kernels.py."""
    from exastencils_b200 import MultigridProgram
    from exastencils_b200 import runtime as rt
    text = "solver_smoother_coloring = %s; solver_smoother_numPre = 2; solver_smoother_numPost = 1; solver_smoother_damping = 0.9; " \
           "solver_cgs_maxNumIts = 128" % {(2, 2): "4-way", (2, 3): "9-way", (3, 2): "8-way", (3, 3): "27-way"}[(dim, ncol)]
    d, nc = rt.cycle_desc_from_options(text, dim)
    assert nc == ncol and d.smoother == (rt.SMOOTH_COLORED_2 if ncol == 2 else rt.SMOOTH_COLORED_3)
    lo, hi = 1, 5 if dim == 2 else 4
    kw = dict(op="scaled", order="bench", smoother="colored%d" % ncol, omega=d.omega, nu_pre=d.nu_pre, nu_post=d.nu_post,
              cg_max_it=d.cg_max_it, cg_rel_tol=d.cg_rel_tol)
    progs = [MultigridProgram(dim, (1,) * dim, lo, hi, **kw) for _ in range(2)]
    rng = np.random.default_rng(31 * dim + ncol)
    u0, f0 = rng.uniform(-1.0, 1.0, size=progs[0].ctx.shape(hi)), rng.uniform(-1.0, 1.0, size=progs[0].ctx.shape(hi))
    for p in progs:
        p.solution(hi)[...] = u0
        p.rhs(hi)[...] = f0
        p.commit()
    lib, man = progs
    ctx = man.ctx

    def cycle(lvl):
        A, u, f, r = man.stencil[lvl], man.f_solution[lvl], man.f_rhs[lvl], man.f_residual[lvl]
        if lvl == lo:
            ctx.cg_coarse(A, u, f, r, man.f_cgtmp0, man.f_cgtmp1, d.cg_max_it, d.cg_rel_tol, want_iters=False)
            return
        ctx.smooth_colored(A, u, f, d.omega, ncol, d.nu_pre)
        ctx.residual(A, u, f, r)
        ctx.restrict(r, man.f_rhs[lvl - 1], 1.0)
        ctx.set(man.f_solution[lvl - 1], 0.0)
        cycle(lvl - 1)
        ctx.prolong_add(man.f_solution[lvl - 1], u)
        ctx.smooth_colored(A, u, f, d.omega, ncol, d.nu_post)

    h = [lib.update_residual_norm()]
    for _ in range(3):
        lib.vcycle()
        cycle(hi)
        h.append(lib.update_residual_norm())
        assert np.array_equal(lib.fetch_solution(), man.fetch_solution())
    assert h[-1] < 0.05 * h[0]
    for p in progs:
        p.close()


def test_binary_field_io_and_benchmark_markers(tmp_path):
    """writeField / readField in binary mode (raw doubles, [comp][z][y][x], no ghost layers; rank-ordered append for
    several blocks) and benchmarkStart / benchmarkStop (NVTX range + device timer of the same name)."""
    from exastencils_b200 import runtime as rt
    ctx = rt.Context(3, 3, 3, block_cells=(2, 1, 3))
    shape = ctx.shape(3)
    rng = np.random.default_rng(5)
    a = rng.uniform(-1.0, 1.0, size=(3,) + shape)
    f, g = ctx.field_create("f", 3, 1, 3), ctx.field_create("g", 3, 2, 3)
    ctx.upload(f, a)
    path = tmp_path / "f.bin"
    ctx.benchmark_start("io")
    ctx.field_write(f, path)
    ctx.field_write(f, path, append=True)   # what a second rank would do
    ctx.benchmark_stop("io")
    raw = np.fromfile(path, dtype=np.float64)
    assert raw.size == 2 * a.size and np.array_equal(raw[:a.size].reshape(a.shape), a) and np.array_equal(raw[a.size:], raw[:a.size])
    ctx.field_read(g, path, slot=1, offset_doubles=a.size)
    assert np.array_equal(ctx.download(g, 3, slot=1, ncomp=3), a)
    with pytest.raises(rt.ExaMGError, match="holds"):
        ctx.field_read(g, path, offset_doubles=a.size + 1)   # short file
    with pytest.raises(rt.ExaMGError, match="without benchmarkStart"):
        ctx.benchmark_stop("io")
    total, count = ctx.timer_total_ms("io")
    assert count == 1 and total > 0.0
    cf = ctx.field_create_cell("c", 3)
    c = rng.uniform(-1.0, 1.0, size=ctx.cell_shape(3))
    ctx.upload(cf, c)
    ctx.field_write(cf, tmp_path / "c.bin")
    assert np.array_equal(np.fromfile(tmp_path / "c.bin").reshape(c.shape), c)
    ctx.close()


@pytest.mark.parametrize("prog,override", [(programs.SMOOTHERS_RBGS, dict(max_level=3, nu_pre=2, nu_post=2, omega=0.8)),
                                           (programs.SMOOTHERS_JAC, dict(max_level=3))],
                         ids=["rbgs", "jacobi"])
def test_solver_cycles_norm_only(prog, override, oracle):
    """examg_solver_cycles(n, &norm): n cycles and the residual norm of Solve@finest's loop body.  The norm equals the
    oracle's (the same loop and sums as examg_solver_residual_norm), and Residual@finest is left alone -- the pass that
    computes the norm does not write it (16 instead of 24 bytes per point; include/examg.h)."""
    ref, gpu = _setup_pair(prog, "library", **override)
    lvl = ref.max_level
    r0 = gpu.update_residual_norm()           # writes Residual@finest
    before = gpu.fetch_residual().copy()
    assert abs(r0 - ref.update_residual_norm()) <= RTOL_NORM * r0
    for n in (1, 2):
        for _ in range(n):
            ref.vcycle()
        r_ref = ref.update_residual_norm()
        r_gpu = gpu.cycles(n)
        assert abs(r_gpu - r_ref) <= RTOL_NORM * r_ref
        assert np.array_equal(gpu.fetch_solution(), ref.solution(lvl).reshape(gpu.ctx.shape(lvl)))
    assert np.array_equal(gpu.fetch_residual(), before), "examg_solver_cycles wrote Residual@finest"
    assert gpu.update_residual_norm() == r_gpu  # the explicit statement pair gives the same number, and writes the field
    assert np.array_equal(gpu.fetch_residual(), ref.residual(lvl).reshape(gpu.ctx.shape(lvl)))
    gpu.close()
    ref.close()


def test_varcoeff27_device_assembly(oracle):
    """InitLaplace of BASELINE config 3 on the device (examg_field_eval: postfix programs with exp) against the
    oracle's host assembly: equal up to the last bits of exp (device libm vs glibc), rows sum to zero."""
    prog = programs.CONFIG3_VARCOEFF27
    ref = make_oracle_solver(prog.cfg, max_level=4)
    gpu = make_gpu_solver(prog.cfg, "library", max_level=4)
    for lvl in range(ref.min_level, ref.max_level + 1):
        got = gpu.ctx.download(gpu.coeff_field[lvl], lvl, ncomp=27)
        want = ref.coeff(lvl).reshape(got.shape)
        assert np.allclose(got, want, rtol=4e-15, atol=0.0), "level %d" % lvl
        off = np.delete(got, 13, axis=0)
        assert np.all(np.abs(off.sum(axis=0) + got[13]) <= 1e-12 * np.abs(got[13]))
    # the cycle runs on the device-assembled operator: same contraction as with the oracle's coefficients
    prog.setup(ref)
    ref.sync_slots()
    prog.setup(gpu)
    gpu.sync_slots()
    r_ref, r_gpu = ref.update_residual_norm(), gpu.update_residual_norm()
    for _ in range(3):
        ref.vcycle()
        gpu.vcycle()
    a, b = ref.update_residual_norm() / r_ref, gpu.update_residual_norm() / r_gpu
    assert abs(a - b) <= 1e-9 * a and b < 0.01
    gpu.close()
    ref.close()


def test_field_eval_ops():
    """examg_field_eval: every postfix op against numpy on a ragged 2D grid (ghost layer included in the loop)."""
    from exastencils_b200 import runtime as rt
    ctx = rt.Context(2, 3, 3, block_cells=(3, 5))
    f = ctx.field_create("f", 3, 1, 2)
    n = ctx.dims(3)
    h = ctx.width(3)
    x = np.arange(n[0]) * h[0]
    y = np.arange(n[1]) * h[1]
    X, Y = x[None, None, :], y[None, :, None]
    ctx.field_eval(f, [(rt.OP_X, 0), (rt.OP_SIN, 0), (rt.OP_Y, 0), (rt.OP_COS, 0), (rt.OP_MUL, 0), (rt.OP_CONST, 2.0),
                       (rt.OP_SQRT, 0), (rt.OP_X, 0), (rt.OP_MUL, 0), (rt.OP_SINH, 0), (rt.OP_ADD, 0)], comp=0)
    ctx.field_eval(f, [(rt.OP_COMP, 0.0), (rt.OP_NEG, 0), (rt.OP_Y, 0), (rt.OP_EXP, 0), (rt.OP_DIV, 0),
                       (rt.OP_CONST, 1.5), (rt.OP_SUB, 0)], comp=1)
    got = ctx.download(f, 3, ncomp=2)
    want0 = np.sin(X) * np.cos(Y) + np.sinh(np.sqrt(2.0) * X)
    assert np.allclose(got[0], np.broadcast_to(want0, got[0].shape), rtol=1e-14, atol=1e-15)
    assert np.allclose(got[1], -got[0] / np.exp(np.broadcast_to(Y, got[0].shape)) - 1.5, rtol=1e-14, atol=1e-15)
    with pytest.raises(rt.ExaMGError):
        ctx.field_eval(f, [(rt.OP_COMP, 2.0)], comp=0)   # component out of range
    with pytest.raises(rt.ExaMGError):
        ctx.field_eval(f, [(rt.OP_X, 0), (rt.OP_ADD, 0)], comp=0)   # stack underflow
    ctx.close()


def test_benchmark_program_properties(oracle):
    """Benchmark/Poisson3D (config 2) at a size the oracle cannot check quickly:
    size-independent properties.  The Dirichlet data x^2 - y^2/2 - z^2/2 is harmonic and
    quadratic, so the 7-point discrete solution equals it: the error must fall to rounding
    level and the residual must contract by > 4x per V(2,2) cycle (omega = 0.8 under-relaxes RBGS;
    the oracle shows the same 5.5x, see test_vcycle_matches_oracle for the bit-level comparison)."""
    from exastencils_b200 import MultigridProgram, runtime as rt
    lvl = 7  # 129^3
    gpu = MultigridProgram(3, (1, 1, 1), 2, lvl, op="scaled", order="bench", smoother="rbgs", omega=0.8, nu_pre=2,
                           nu_post=2, cg_max_it=128, cg_rel_tol=1e-3)
    f = gpu.f_solution[lvl]
    prog = [(rt.OP_X, 0), (rt.OP_X, 0), (rt.OP_MUL, 0),
            (rt.OP_CONST, 0.5), (rt.OP_Y, 0), (rt.OP_MUL, 0), (rt.OP_Y, 0), (rt.OP_MUL, 0), (rt.OP_SUB, 0),
            (rt.OP_CONST, 0.5), (rt.OP_Z, 0), (rt.OP_MUL, 0), (rt.OP_Z, 0), (rt.OP_MUL, 0), (rt.OP_SUB, 0)]
    gpu.ctx.field_set_bc_expr(f, prog)
    gpu.ctx.apply_bc(f)
    hist = gpu.Solve(max_it=20, target_reduction=1e-10)
    assert len(hist) <= 16
    for a, b in zip(hist[:-2], hist[1:-1]):
        assert b < 0.25 * a
    x, y, z = gpu.node_positions(lvl)
    exact = (x[None, None, :] ** 2 - 0.5 * y[None, :, None] ** 2) - 0.5 * z[:, None, None] ** 2
    assert np.max(np.abs(gpu.fetch_solution() - exact)) < 1e-9
    gpu.close()


def test_print_field(tmp_path):
    """printField ( "file", F ) in ascii (C/s/e/field/ir/IR_PrintField.scala:49-100): position and value per point, x fastest,
    std::defaultfloat -- node field in 3D, cell field with two components in 2D and a custom separator."""
    from exastencils_b200 import runtime as rt
    ctx = rt.Context(3, 2, 2, block_cells=(1, 2, 1))
    f = ctx.field_create("u", 2)
    shape = ctx.shape(2)
    u = np.random.default_rng(3).uniform(-1.0, 1.0, size=shape)
    ctx.upload(f, u)
    path = str(tmp_path / "u.txt")
    ctx.field_print(f, path)
    h = ctx.width(2)
    lines = open(path).read().splitlines()
    assert len(lines) == u.size
    n = ctx.dims(2)
    for t in (0, 1, n[0], n[0] * n[1] + 3, u.size - 1):
        i, j, k = t % n[0], (t // n[0]) % n[1], t // (n[0] * n[1])
        assert lines[t] == "%g %g %g %g" % (i * h[0], j * h[1], k * h[2], u[k, j, i])
    ctx.close()
    ctx = rt.Context(2, 3, 3, block_cells=(1, 1))
    f = ctx.field_create_cell("c", 3, 1, 2)
    cs = ctx.cell_shape(3)
    v = np.random.default_rng(4).uniform(-1.0, 1.0, size=(2,) + cs)
    ctx.upload(f, np.ascontiguousarray(v))
    path = str(tmp_path / "c.csv")
    ctx.field_print(f, path, separator=",")
    lines = open(path).read().splitlines()
    hh = ctx.width(3)
    assert len(lines) == cs[1] * cs[2]
    assert lines[9] == "%g,%g,%g,%g" % (1 * hh[0] + 0.5 * hh[0], 1 * hh[1] + 0.5 * hh[1], v[0, 0, 1, 1], v[1, 0, 1, 1])
    ctx.close()


@pytest.mark.parametrize("cgs", ["bicgstab", "cr", "minres", "smoother"])
def test_coarse_solver_choice(cgs, oracle):
    """solver_cgs of the layer-3 generator inside the library cycle (examg_cycle_desc.coarse_solver): on the SPD Poisson
    operator every coarse solver reaches the same 1e-3 coarse reduction, so the V-cycle history stays that of the CG
    program to a few per cent and the Solve loop takes the same number of cycles."""
    prog = programs.SMOOTHERS_JAC
    hist = {}
    for name in ("cg", cgs):
        from exastencils_b200 import MultigridProgram
        c = dict(prog.cfg, max_level=3)
        gpu = MultigridProgram(c["dim"], tuple(c["F"]), c["min_level"], c["max_level"], op=c["op"], order=c["order"],
                               smoother=c["smoother"], omega=c["omega"], nu_pre=c["nu_pre"], nu_post=c["nu_post"],
                               restrict_scale=c["restrict_scale"], cg_max_it=c["cg_max_it"], cg_rel_tol=c["cg_rel_tol"],
                               coarse_solver=name)
        prog.setup(gpu)
        gpu.sync_slots()
        h = [gpu.update_residual_norm()]
        for _ in range(6):
            gpu.vcycle()
            h.append(gpu.update_residual_norm())
        assert gpu.last_cg_iters() >= 1
        gpu.close()
        hist[name] = h
    assert all(abs(a - b) <= 0.05 * b for a, b in zip(hist[cgs], hist["cg"])), (hist[cgs], hist["cg"])
    assert hist[cgs][-1] < 1e-3 * hist[cgs][0]


def test_fas_2d_basic_golden_and_oracle(oracle):
    """Testing/NonLinear/FAS_2D_Basic: the nonlinear full approximation scheme program through libexamg --
    stencil convolutions by examg_apply, the pointwise nonlinear parts by examg_loop_eval, transfers by examg_restrict /
    examg_prolong_add -- prints the reference's golden stdout, and follows the CPU oracle to 1e-9 (exp() is the
    device's libm on one side and the host's on the other)."""
    import fas_program as fp
    g = fp.GpuFas(fp.FasProgram())
    assert fp.solve(g, oracle.print_reduced_prec) == fp.expected()
    g.close()
    prog = fp.FasProgram(max_level=5)
    ref, gpu = fp.OracleFas(prog, oracle), fp.GpuFas(prog)
    for b in (ref, gpu):
        b.init_finest()
        fp.up_residual(b, 5)
    assert abs(gpu.norm(5) - ref.norm(5)) <= 1e-12 * ref.norm(5)
    for cyc in range(3):
        for b in (ref, gpu):
            fp.vcycle(b, 5)
            fp.up_residual(b, 5)
        ug, ur = gpu.solution(5), ref.solution(5)
        assert np.max(np.abs(ug.reshape(ur.shape) - ur)) <= 1e-9 * np.max(np.abs(ur)), cyc
        assert abs(gpu.norm(5) - ref.norm(5)) <= 1e-8 * ref.norm(5)
        assert abs(gpu.max_error(5) - ref.max_error(5)) <= 1e-9 * ref.max_error(5)
    gpu.close()
    ref.close()


def test_loop_eval_is_bit_exact_without_libm(oracle):
    """examg_loop_eval against og_loop_eval on a program without transcendental functions (+ - * / neg sqrt, node
    positions, three source fields, one of them the destination): bit-identical; and its argument checks."""
    import fas_program as fp
    from exastencils_b200 import runtime as rt
    ctx = rt.Context(3, 2, 2, block_cells=(3, 2, 1))
    shape = ctx.shape(2)
    rng = np.random.default_rng(8)
    a, b, d = (rng.uniform(0.5, 2.0, size=shape) for _ in range(3))
    fa, fb, fd = ctx.field_create("a", 2), ctx.field_create("b", 2), ctx.field_create("d", 2, 2)
    ctx.upload(fa, a)
    ctx.upload(fb, b)
    ctx.upload(fd, d)
    F, K, O = fp.fld, fp.c, fp.op
    prog = F(2) + F(0) + O(fp.X) + O(fp.MUL) + F(1) + O(fp.Y) + K(0.25) + O(fp.ADD) + O(fp.DIV) + O(fp.SUB) + O(fp.ADD) \
        + F(0) + O(fp.SQRT) + O(fp.NEG) + O(fp.Z) + O(fp.MUL) + O(fp.SUB)
    ctx.loop_eval(fd, [fa, fb, fd], prog)
    want = d.copy()
    h = [1.0 / (c << 2) for c in (3, 2, 1)]
    oracle.set_periodic((0, 0, 0))
    oracle.loop_eval(want, [a, b, d.copy()], prog, h)
    assert np.array_equal(ctx.download(fd, 2), want)
    assert not np.array_equal(want, d)
    with pytest.raises(rt.ExaMGError, match="source field"):
        ctx.loop_eval(fd, [fa], F(1))
    ctx.close()
    two = rt.Context(3, 1, 2, block_cells=(3, 2, 1))
    with pytest.raises(rt.ExaMGError, match="level"):
        two.loop_eval(two.field_create("x", 2), [two.field_create("y", 1)], F(0))
    two.close()


@pytest.mark.parametrize("dim,smoother", [(2, "jacobi"), (3, "rbgs")])
def test_fas_in_the_library_cycle(dim, smoother, oracle):
    """solver_useFAS = true inside the library cycle (examg_cycle_desc.use_fas, reached through
    examg_cycle_desc_from_options): bit-identical to the oracle's FAS cycle (og_config.use_fas) from a random first
    guess, and the same iterates as the correction scheme up to rounding."""
    from exastencils_b200 import MultigridProgram
    from exastencils_b200 import runtime as rt
    d, _ = rt.cycle_desc_from_options("solver_useFAS = true; solver_smoother_coloring = rb", dim)
    assert d.use_fas == 1
    lo, hi = 1, 5 if dim == 2 else 4
    cells = (3, 2) if dim == 2 else (1, 2, 1)
    kw = dict(op="scaled", order="bench", smoother=smoother, omega=0.8, nu_pre=2, nu_post=2, cg_max_it=128, cg_rel_tol=1e-3)
    fas, cs = MultigridProgram(dim, cells, lo, hi, use_fas=True, **kw), MultigridProgram(dim, cells, lo, hi, **kw)
    ref = oracle.Solver(dim, cells, lo, hi, smoother=oracle.SMOOTH_JACOBI if smoother == "jacobi" else oracle.SMOOTH_RBGS,
                        omega=0.8, nu_pre=2, nu_post=2, cg_max_it=128, cg_rel_tol=1e-3, use_fas=True)
    rng = np.random.default_rng(17 + dim)
    u0, f0 = rng.uniform(-1.0, 1.0, size=fas.ctx.shape(hi)), rng.uniform(-1.0, 1.0, size=fas.ctx.shape(hi))
    for p in (fas, cs):
        p.solution(hi)[...] = u0
        p.rhs(hi)[...] = f0
        p.commit()
    ref.solution(hi)[...] = u0.reshape(ref.solution(hi).shape)
    ref.rhs(hi)[...] = f0.reshape(ref.rhs(hi).shape)
    ref.sync_slots()
    h = [fas.update_residual_norm()]
    assert h[0] == ref.update_residual_norm()
    for _ in range(3):
        fas.vcycle()
        cs.vcycle()
        ref.vcycle()
        h.append(fas.update_residual_norm())
        assert h[-1] == ref.update_residual_norm()
        assert fas.last_cg_iters() == ref.last_cg_iters()
        assert np.array_equal(fas.fetch_solution(), ref.solution(hi).reshape(fas.fetch_solution().shape))
        assert np.max(np.abs(fas.fetch_solution() - cs.fetch_solution())) <= 1e-10 * np.max(np.abs(u0))
    assert h[-1] < 0.02 * h[0]
    for p in (fas, cs, ref):
        p.close()
