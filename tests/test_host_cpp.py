"""The C++ host programs under host/ -- the code the generator back-end prints, calling the C ABI directly.

CPU: they compile and link against libexamg.so.  GPU: their stdout equals the reference's golden .results files
(Testing/Smoothers/RBGS.results, Testing/Application/OpticalFlow2D.results, Testing/BC/3D_Periodic.results)."""
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
    for name in ("Testing_Smoothers_RBGS", "Testing_Application_OpticalFlow2D", "Testing_BC_3D_Periodic"):
        exe = os.path.join(HOST, "_build", name)
        assert os.access(exe, os.X_OK)
        out = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
        assert "libexamg.so" in out and "not found" not in out.split("libexamg.so")[1].splitlines()[0]


@pytest.mark.gpu
@pytest.mark.parametrize("name,golden", [("Testing_Smoothers_RBGS", "Smoothers_RBGS"),
                                         ("Testing_Application_OpticalFlow2D", "Application_OpticalFlow2D"),
                                         ("Testing_BC_3D_Periodic", "BC_3D_Periodic")])
def test_host_program_golden_stdout(name, golden):
    _build()
    out = subprocess.run([os.path.join(HOST, "_build", name)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    with open(os.path.join(ROOT, "tests", "golden", golden + ".results")) as f:
        assert out.stdout.split() == f.read().split()
