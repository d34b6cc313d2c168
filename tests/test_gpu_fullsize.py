"""Parity at BASELINE.json's full sizes, two ways.

1. Directly against the CPU oracle (test_oracle_parity_at_full_size): the oracle runs a 512^3 V(2,2) cycle in about
   half a second on the box's host cores, so three cycles of BASELINE config 2 are compared node by node
   (np.array_equal on the 513^3 Solution) and norm by norm (<= 1e-12 relative, the north star's tolerance; they agree
   to the last bits).  The 27-point stencil-field program (config 3's operator) at 256^3 likewise, at the tolerance
   its exp() allows (device libm vs host libm in the coefficients).
2. Against the library's own generic kernel set in a second process with every fast path switched off
   (test_fused_equals_generic_at_full_size): residual histories as hex floats, CG counts and a SHA-256 of the final
   Solution are identical -- fusing sweeps, residual+restriction, the fast prolongation and the CUDA graph never change
   a bit, at any size, also for the Jacobi programs the oracle comparison above does not cover.
"""
import os
import subprocess
import sys

import pytest

import numpy as np

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

GENERIC = {"EXAMG_NO_FUSED": "1", "EXAMG_LEX": "0", "EXAMG_NO_FUSED_RESTRICT": "1", "EXAMG_NO_FAST_PROLONG": "1", "EXAMG_NO_GRAPH": "1",
           "EXAMG_JACV_FUSE": "0", "EXAMG_JAC_FUSE_SWEEPS": "0"}


def _run(kind, level, cycles, env_extra):
    env = dict(os.environ)
    env.update(env_extra)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "fullsize_worker.py"), kind, str(level), str(cycles)],
                         capture_output=True, text=True, timeout=900, env=env)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    rec = dict(line.split(" ", 1) for line in out.stdout.splitlines() if line.split(" ", 1)[0] in
               ("HISTORY", "CG", "SHA256", "LAUNCHES"))
    return rec


@pytest.mark.parametrize("kind,level,cycles", [
    ("poisson7", 9, 3),    # BASELINE config 2: 512^3, V(2,2) RBGS -- k_rbw<4>, k_resrestrict3, k_prolong3_fast, graph
    ("jacobi7", 9, 2),     # 512^3, V(2,3) Jacobi -- k_jacw<2>, k_jacw<3>
    ("varcoef27", 8, 2),   # BASELINE config 3 at 256^3 -- k_jacv2<27>
])
def test_fused_equals_generic_at_full_size(kind, level, cycles):
    fast = _run(kind, level, cycles, {})
    slow = _run(kind, level, cycles, GENERIC)
    assert int(fast["LAUNCHES"]) < int(slow["LAUNCHES"]), "the fast paths did not run"
    assert fast["HISTORY"] == slow["HISTORY"]
    assert fast["CG"] == slow["CG"]
    assert fast["SHA256"] == slow["SHA256"]
    hist = [float.fromhex(v) for v in fast["HISTORY"].split()]
    assert hist[-1] < 0.3 ** cycles * hist[0]  # the cycle converges (damped Jacobi: 0.21 per V(2,3))


def test_oracle_parity_at_full_size(oracle):
    """BASELINE config 2 (512^3 cells, levels 2..9, V(2,2) RBGS, CG): libexamg against the oracle, same inputs, three
    cycles: norms within 1e-12 relative, Solution bit-identical, CG iteration counts equal."""
    import bench
    from exastencils_b200 import MultigridProgram
    og = oracle
    level, cycles = 9, 3
    prog = MultigridProgram(3, (1, 1, 1), 2, level, op="scaled", order="bench", smoother="rbgs", omega=0.8, nu_pre=2,
                            nu_post=2, cg_max_it=128, cg_rel_tol=1e-3)
    bench.set_benchmark_problem(prog)
    ref = og.Solver(3, (1, 1, 1), 2, level, op=og.OP_LAPLACE_SCALED, order=og.ORDER_BENCH, smoother=og.SMOOTH_RBGS,
                    omega=0.8, nu_pre=2, nu_post=2, cg_max_it=128, cg_rel_tol=1e-3)
    bench._oracle_problem(ref, level)
    hg, hr = [prog.update_residual_norm()], [ref.update_residual_norm()]
    for _ in range(cycles):
        prog.vcycle()
        ref.vcycle()
        hg.append(prog.update_residual_norm())
        hr.append(ref.update_residual_norm())
        assert prog.last_cg_iters() == ref.last_cg_iters()
    for a, b in zip(hg, hr):
        assert abs(a - b) <= 1e-12 * abs(b), (hg, hr)
    ug = prog.fetch_solution()
    assert ug.shape == ref.solution(level).shape == (513, 513, 513)
    assert np.array_equal(ug, ref.solution(level))
    assert hg[-1] < 0.3 ** cycles * hg[0]
    prog.close()
    ref.close()


def test_oracle_parity_varcoef27_at_256(oracle):
    """BASELINE config 3's operator at 256^3 (27-point stencil field assembled on the device, damped Jacobi with fused
    pairs of sweeps): two cycles against the oracle.  The coefficients go through exp() -- the device's libm on one
    side, the host's on the other -- so fields and norms are compared to 1e-9 relative, as in tests/mgpu/worker.py;
    bit-level parity of this program is what test_fused_equals_generic_at_full_size pins."""
    import bench
    from exastencils_b200 import MultigridProgram
    og = oracle
    level, cycles = 8, 2
    prog = MultigridProgram(3, (1, 1, 1), 2, level, op="varcoeff27", order="testing", kappa=bench.KAPPA,
                            smoother="jacobi", omega=0.85, nu_pre=2, nu_post=2, cg_max_it=128, cg_rel_tol=1e-3)
    bench.set_varcoef_problem(prog)
    ref = bench._oracle_solver(og, "varcoef27", level)
    hg, hr = [prog.update_residual_norm()], [ref.update_residual_norm()]
    for _ in range(cycles):
        prog.vcycle()
        ref.vcycle()
        hg.append(prog.update_residual_norm())
        hr.append(ref.update_residual_norm())
    for a, b in zip(hg, hr):
        assert abs(a - b) <= 1e-9 * abs(b), (hg, hr)
    ug, ur = prog.fetch_solution(), ref.solution(level)
    assert float(np.max(np.abs(ug - ur))) <= 1e-9 * float(np.max(np.abs(ur)))
    prog.close()
    ref.close()
