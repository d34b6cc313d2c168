"""Pins the CPU oracle against the reference's own golden vectors
(Testing/**/*.results, copied to tests/golden/ by tests/golden/make_golden.py).

The reference's comparator (Testing/run_test.py:12-43) accepts a line if the
strings are equal or the numbers differ by <= 1e-6 absolute; we require string
equality of every printed token (4 significant digits), which is stricter.
"""
import pytest

import programs
from oracle_driver import make_oracle_solver


def _run(prog, og):
    s = make_oracle_solver(prog.cfg)
    prog.setup(s)
    s.sync_slots()
    out = programs.run_solve(prog, s, og.print_reduced_prec, og.max_abs_diff)
    s.close()
    return out


@pytest.mark.parametrize("prog", [programs.POISSON_2D, programs.SISC_2D_CONST, programs.SISC_2D_VAR],
                         ids=lambda p: p.name)
def test_golden_2d(prog, oracle):
    assert _run(prog, oracle) == prog.expected()


@pytest.mark.slow
@pytest.mark.parametrize("prog", [programs.SMOOTHERS_RBGS, programs.SMOOTHERS_JAC, programs.SISC_3D_CONST,
                                  programs.SISC_3D_VAR], ids=lambda p: p.name)
def test_golden_3d(prog, oracle):
    assert _run(prog, oracle) == prog.expected()


def test_print_reduced_prec(oracle):
    p = oracle.print_reduced_prec
    assert p(6.714e7 + 1.0) == "6.714e+07"
    assert p(2812.2) == "2812"
    assert p(129.83) == "129.8"
    assert p(0.00078212) == "0.0007821"
    assert p(5e-13) == "EFFECTIVELY ZERO"
    assert p(5.55e-12) == "6e-12"
    assert p(5.55e-11) == "5.5e-11" or p(5.55e-11) == "5.6e-11"
    assert p(5.554e-10) == "5.55e-10"
