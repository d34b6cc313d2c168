"""Builds an oracle Solver for a tests.programs.Program (test helper)."""
from oracle import oracle as og

_OP = {"scaled": og.OP_LAPLACE_SCALED, "unscaled": og.OP_LAPLACE_UNSCALED, "varcoeff": og.OP_VARCOEFF, "varcoeff27": og.OP_VARCOEFF27}
_ORDER = {"bench": og.ORDER_BENCH, "testing": og.ORDER_TESTING}
_SM = {"jacobi": og.SMOOTH_JACOBI, "rbgs": og.SMOOTH_RBGS}


def make_oracle_solver(cfg, **override):
    c = dict(cfg)
    c.update(override)
    og.set_periodic(c.get("periodic", (0, 0, 0)))  # process-wide switch: every solver creation sets it
    return og.Solver(c["dim"], c["F"], c["min_level"], c["max_level"], op=_OP[c["op"]], order=_ORDER[c["order"]],
                     kappa=c.get("kappa", 0.0), smoother=_SM[c["smoother"]], omega=c["omega"],
                     nu_pre=c["nu_pre"], nu_post=c["nu_post"], restrict_scale=c["restrict_scale"],
                     cg_max_it=c["cg_max_it"], cg_rel_tol=c["cg_rel_tol"])
