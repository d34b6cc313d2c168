"""The C++ host programs under host/ -- the code the generator back-end prints, calling the C ABI directly.

CPU: they compile and link against libexamg.so.  GPU: their stdout equals the reference's golden .results files
(Testing/Smoothers/RBGS.results, Testing/Application/OpticalFlow2D.results and OpticalFlow3D.results,
Testing/BC/3D_Periodic.results)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "host")


def _build():
    import exastencils_b200
    exastencils_b200.build()
    subprocess.check_call(["make", "-C", HOST, "-s"])


def test_host_programs_link():
    _build()
    for name in ("Testing_Smoothers_RBGS", "Testing_Application_OpticalFlow2D", "Testing_Application_OpticalFlow3D",
                 "Testing_BC_3D_Periodic", "Benchmark_Poisson3D"):
        exe = os.path.join(HOST, "_build", name)
        assert os.access(exe, os.X_OK)
        out = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
        assert "libexamg.so" in out and "not found" not in out.split("libexamg.so")[1].splitlines()[0]


@pytest.mark.gpu
@pytest.mark.parametrize("name,golden", [("Testing_Smoothers_RBGS", "Smoothers_RBGS"),
                                         ("Testing_Application_OpticalFlow2D", "Application_OpticalFlow2D"),
                                         ("Testing_Application_OpticalFlow3D", "Application_OpticalFlow3D"),
                                         ("Testing_BC_3D_Periodic", "BC_3D_Periodic")])
def test_host_program_golden_stdout(name, golden):
    _build()
    out = subprocess.run([os.path.join(HOST, "_build", name)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    with open(os.path.join(ROOT, "tests", "golden", golden + ".results")) as f:
        assert out.stdout.split() == f.read().split()


@pytest.mark.gpu
def test_benchmark_program_host_cpp(tmp_path):
    """host/Benchmark_Poisson3D.cpp (the north-star program as the back-end prints it) at 129^3, V(2,2): the statement-by-
    statement level loop and the library cycle (CUDA graph) print the same residual history, which equals the Python
    mirror's (MultigridProgram.Solve), converges in <= 12 cycles and writes results.json."""
    _build()
    exe = os.path.join(HOST, "_build", "Benchmark_Poisson3D")
    outs = []
    for lib_cycle in ("0", "1"):
        env = dict(os.environ, EXAMG_HOST_LIBRARY_CYCLE=lib_cycle)
        out = subprocess.run([exe, "7", "2"], capture_output=True, text=True, timeout=300, env=env, cwd=str(tmp_path))
        assert out.returncode == 0, out.stderr
        lines = [l for l in out.stdout.splitlines() if l.startswith(("Starting residual", "Residual after"))]
        outs.append(lines)
        assert os.path.exists(os.path.join(str(tmp_path), "results.json"))
    assert outs[0] == outs[1]
    assert 3 <= len(outs[0]) - 1 <= 12
    from exastencils_b200 import MultigridProgram, runtime as rt
    gpu = MultigridProgram(3, (1, 1, 1), 2, 7, op="scaled", order="bench", smoother="rbgs", omega=0.8, nu_pre=2,
                           nu_post=2, cg_max_it=128, cg_rel_tol=1e-3)
    f = gpu.f_solution[7]
    gpu.ctx.field_set_bc_expr(f, [(rt.OP_X, 0), (rt.OP_X, 0), (rt.OP_MUL, 0),
                                  (rt.OP_CONST, 0.5), (rt.OP_Y, 0), (rt.OP_MUL, 0), (rt.OP_Y, 0), (rt.OP_MUL, 0), (rt.OP_SUB, 0),
                                  (rt.OP_CONST, 0.5), (rt.OP_Z, 0), (rt.OP_MUL, 0), (rt.OP_Z, 0), (rt.OP_MUL, 0), (rt.OP_SUB, 0)])
    gpu.ctx.apply_bc(f)
    hist = gpu.Solve(max_it=100, target_reduction=1e-6)
    gpu.close()
    printed = [float(outs[0][0].split()[-1])] + [float(l.split()[5]) for l in outs[0][1:]]
    assert len(printed) == len(hist)
    assert all(abs(a - b) <= 1e-5 * abs(b) for a, b in zip(printed, hist))  # std::cout prints 6 significant digits
