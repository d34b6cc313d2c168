"""Cell-centred finite-volume Poisson on the GPU (Testing/CellBased/{2D,3D}_{Basic,Neumann}.exa4): the
programs' text (tests/cell_programs.py) on libexamg's cell-localised fields must print the reference's golden stdout and
match the CPU oracle -- bit for bit where the boundary data is a constant or Neumann, to 1e-12 where the Dirichlet
expression goes through the device's sin / sinh."""
import numpy as np
import pytest

import cell_programs as cp

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("prog", cp.PROGRAMS, ids=lambda p: p.name)
def test_golden_stdout(prog, oracle):
    b = cp.GpuCells(prog)
    out = cp.solve(b, oracle.print_reduced_prec)
    b.close()
    assert out == prog.expected()


def test_golden_stdout_hybrid_cell(oracle):
    """Testing/CommBasic/{HybridCell,Strategy26Cell}: 576^3 cells (1.9e8 unknowns) on one GPU; the oracle pins the same
    program at a lower finest level (test_oracle_parity_hybrid_cell)."""
    b = cp.GpuCells(cp.HYBRID_CELL)
    out = cp.solve(b, oracle.print_reduced_prec)
    b.close()
    assert out == cp.HYBRID_CELL.expected()
    import programs
    assert programs.golden("CommBasic_Strategy26Cell") == cp.HYBRID_CELL.expected()


def test_oracle_parity_hybrid_cell(oracle):
    prog = cp.CellProgram("HybridCell", 3, False, max_level=2, F=9, kind="hybridcell", golden="CommBasic_HybridCell")
    ref, gpu = cp.OracleCells(prog), cp.GpuCells(prog)
    out_ref, out_gpu = cp.solve(ref, float.hex), cp.solve(gpu, float.hex)
    assert out_gpu == out_ref  # polynomial boundary data: no libm involved, residual history bit-identical
    assert np.array_equal(gpu.interior("u", 2), ref.interior("u", 2).reshape(gpu.interior("u", 2).shape))
    gpu.close()
    ref.close()


@pytest.mark.parametrize("name,dim,neumann,lvl", [("2D_Basic", 2, False, 5), ("2D_Neumann", 2, True, 5),
                                                  ("3D_Basic", 3, False, 3), ("3D_Neumann", 3, True, 3)])
def test_oracle_parity(name, dim, neumann, lvl, oracle):
    prog = cp.CellProgram(name, dim, neumann, max_level=lvl)
    ref, gpu = cp.OracleCells(prog), cp.GpuCells(prog)
    L = prog.max_level
    for b in (ref, gpu):
        cp.up_residual(b, L)
    assert abs(gpu.norm(L) - ref.norm(L)) <= 1e-12 * ref.norm(L)
    for cyc in range(3):
        for b in (ref, gpu):
            cp.vcycle(b, L)
            if neumann:
                b.shift("u", L, -(b.sum("u", L) / float(prog.cells(L) ** dim)))
            cp.up_residual(b, L)
        ug = gpu.interior("u", L)
        ur = ref.interior("u", L).reshape(ug.shape)
        if neumann:
            assert np.array_equal(ug, ur), "cycle %d" % cyc
        else:
            assert np.allclose(ug, ur, rtol=1e-12, atol=1e-12 * np.max(np.abs(ur))), "cycle %d" % cyc
        assert abs(gpu.norm(L) - ref.norm(L)) <= 1e-10 * ref.norm(L) + 1e-300
        assert abs(gpu.max_error(L) - ref.max_error(L)) <= 1e-10 * ref.max_error(L)
    gpu.close()
    ref.close()


def test_node_only_entry_points_reject_cell_fields():
    """Entry points whose loops exist for node fields only fail loudly on cell-localised handles instead of running
    with the wrong iteration space."""
    from exastencils_b200 import runtime as rt
    ctx = rt.Context(3, 1, 2, block_cells=(2, 2, 2))
    cf, cc = ctx.field_create_cell("f", 2), ctx.field_create_cell("c", 1)
    offs = [(0, 0, 0), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]
    A = ctx.stencil_const(1, offs, [6.0] + [-1.0] * 6)
    tmp = [ctx.field_create_cell("t%d" % k, 1) for k in range(4)]
    with pytest.raises(rt.ExaMGError):
        ctx.restrict(cf, cc)
    with pytest.raises(rt.ExaMGError):
        ctx.prolong_add(cc, cf)
    with pytest.raises(rt.ExaMGError):
        ctx.cg_coarse(A, cc, tmp[0], tmp[1], tmp[2], tmp[3], 10, 1e-3)
    nf = ctx.field_create("node", 1)
    with pytest.raises(rt.ExaMGError):
        ctx.residual(A, cc, nf, tmp[0])  # mixed localisations
    ctx.close()


def test_five_point_benchmark_app(oracle):
    """Benchmark/FivePointStencil/5pt_Jac_Cell.exa4 (tools/fivepoint_bench.py) at level 5: anisotropic cell widths, the
    user-written ghost-cell function (a condition of its own on the top face, examg_field_set_bc_face) -- 20 iterations of
    Smoother + UpResidual against the oracle's cell loops."""
    import math
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import fivepoint_bench as fb
    og = oracle
    og.set_periodic((0, 0, 0))
    lvl = 5
    ctx, A, u, f, r, (nx, ny, hx, hy) = fb.setup(lvl)
    offs = [(0, 0, 0), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0)]
    st = og.Stencil.make(offs, [(2.0 / (hx * hx)) + (2.0 / (hy * hy)), -1.0 / (hx * hx), -1.0 / (hx * hx),
                                -1.0 / (hy * hy), -1.0 / (hy * hy)])
    U = np.zeros((1, ny + 2, nx + 2))
    F = np.zeros_like(U)
    F[:, 1:-1, 1:-1] = ctx.download_cells(f, lvl)
    top = 0.5 * (np.sin((2.0 * math.pi) * (np.arange(nx) * hx)) * math.sinh(2.0 * math.pi))
    faces = [None, None, None, np.ascontiguousarray(top.reshape(1, nx)), None, None]
    for it in range(20):
        got = fb.iteration(ctx, A, u, f, r)
        og.cell_apply_bc(2, U, og.BC_CELL_DIRICHLET, faces)
        U = og.smooth_jacobi(st, 0.8, U, F)
        og.cell_apply_bc(2, U, og.BC_CELL_DIRICHLET, faces)
        R = og.residual(st, U, F)
        want = math.sqrt(og.dot(R, R))
        assert abs(got - want) <= 1e-11 * want, it
    ug = ctx.download_cells(u, lvl)
    assert np.allclose(ug, U[:, 1:-1, 1:-1], rtol=1e-11, atol=1e-11 * np.max(np.abs(U)))
    ctx.close()


@pytest.mark.parametrize("dim,neumann", [(2, False), (2, True), (3, False), (3, True)])
def test_cell_smoothers_refresh_ghost_cells(dim, neumann, oracle):
    """Round-1 advisor finding: on cell-localised fields every colour of examg_smooth_rbgs / examg_smooth_rb_color /
    examg_smooth_colored and every sweep of a multi-sweep examg_smooth_jacobi is { update; apply bc } -- the ghost cells
    (Dirichlet 2g - u, Neumann u) follow the inner cells -- as the smoother bodies of Testing/CellBased/*.exa4 spell it
    out.  Against the oracle's loops with og_cell_apply_bc after every colour / sweep: bit-identical."""
    from exastencils_b200 import runtime as rt
    og = oracle
    og.set_periodic((0, 0, 0))
    lvl = 4 if dim == 2 else 3
    cells = (3, 5) if dim == 2 else (2, 3, 1)
    ctx = rt.Context(dim, lvl, lvl, block_cells=cells)
    n = ctx.cell_shape(lvl)  # (nz, ny, nx) of the cells
    rng = np.random.default_rng(11 + dim + 2 * int(neumann))
    offs = [(0, 0, 0), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0)] + ([(0, 0, 1), (0, 0, -1)] if dim == 3 else [])
    coef = [2.0 * dim + 0.5] + [-1.0] * (2 * dim)
    A = ctx.stencil_const(lvl, offs, coef)
    st = og.Stencil.make(offs, coef)
    u0, f0 = rng.uniform(-1.0, 1.0, size=n), rng.uniform(-1.0, 1.0, size=n)
    kind_gpu, kind_og = (rt.BC_NEUMANN, og.BC_CELL_NEUMANN) if neumann else (rt.BC_DIRICHLET_CONST, og.BC_CELL_DIRICHLET)
    g = 0.0 if neumann else 0.75
    pad = [(1, 1)] * 3 if dim == 3 else [(0, 0), (1, 1), (1, 1)]
    faces = None
    if not neumann:
        faces = []
        for fc in range(6):
            d = fc // 2
            if d >= dim:
                faces.append(None)
                continue
            shp = [n[2 - w] for w in range(3) if w != d and w < dim]  # the face's cells, x fastest
            faces.append(np.full(int(np.prod(shp)), g))

    def bc(U):
        og.cell_apply_bc(dim, U, kind_og, faces)

    def fresh():
        fu, ff = ctx.field_create_cell("u", lvl, 2), ctx.field_create_cell("f", lvl, 1)
        ctx.field_set_bc(fu, kind_gpu, g)
        for s in (0, 1):
            ctx.upload(fu, u0, s)
            ctx.apply_bc(fu, s)
        ctx.upload(ff, f0)
        U = np.pad(u0, pad)
        F = np.pad(f0, pad)
        bc(U)
        return fu, ff, U, F

    # red-black: colour c of the cells = colour (c + dim) % 2 of the padded array the oracle loops over
    fu, ff, U, F = fresh()
    ctx.smooth_rbgs(A, fu, ff, 0.8, 2)
    for _ in range(2):
        for colour in (0, 1):
            og.smooth_rb(st, 0.8, U, F, (colour + dim) % 2)
            bc(U)
    sl = (slice(1, -1),) * 3 if dim == 3 else (slice(None), slice(1, -1), slice(1, -1))
    assert np.array_equal(ctx.download_cells(fu, lvl), U[sl])
    # three Jacobi sweeps in one call
    fu, ff, U, F = fresh()
    ctx.smooth_jacobi(A, fu, ff, 0.8, 3)
    for _ in range(3):
        U = og.smooth_jacobi(st, 0.8, U, F)
        bc(U)
    assert np.array_equal(ctx.download_cells(fu, lvl), U[sl])
    # 2^dim-way coloured Gauss-Seidel
    fu, ff, U, F = fresh()
    ctx.smooth_colored(A, fu, ff, 0.8, 2, 1)
    for cz in range(2 if dim == 3 else 1):
        for cy in range(2):
            for cx in range(2):
                # cell colour c_d <-> padded-array index colour (c_d + 1) % 2 in the padded dimensions
                og.smooth_color(st, 0.8, U, F, 2, (cx + 1) % 2, (cy + 1) % 2, (cz + 1) % 2 if dim == 3 else 0)
                bc(U)
    assert np.array_equal(ctx.download_cells(fu, lvl), U[sl])
    ctx.close()
