"""Generate the golden closed-loop vectors with the REFERENCE leg of the oracle: the unmodified acados + HPIPM +
BLASFEO of /root/reference compiled by oracle/ref.mk and driven by oracle/ref_driver.c (which stands in for the
generated acados_solver_<name>.c and for R/simulator.py's loop).  Needs /root/reference to have been built
(`make -C oracle -f ref.mk ref-driver`); the committed .npz files are what travels.

    python tests/golden/make_golden.py
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT))
from oracle import pyoracle as o  # noqa: E402

CASES = {
    # BASELINE.json configs[0]: single closed-loop run, flat surface, README initial state
    "config1_rti_n20_flat": dict(N=20, solver_type=o.SQP_RTI, surf=[0.0] * 6, nsteps=600, seed=None),
    "sqp_n20_default_surface": dict(N=20, solver_type=o.SQP, surf=None, nsteps=120, seed=None),
    "rti_n100_random": dict(N=100, solver_type=o.SQP_RTI, surf="random", nsteps=60, seed=46),
    "sqp_n30_random_wgrid": dict(N=30, solver_type=o.SQP, surf="random", nsteps=60, seed=43, w_u=0.001, w_qddot=0.05),
    "rti_n10_random": dict(N=10, solver_type=o.SQP_RTI, surf="random", nsteps=100, seed=47),
    # terminal ellipsoid x_N' P x_N <= alpha (BASELINE configs[3] ingredient): P, alpha from the LQR
    # construction of R/_terminal_constraint.py (Q = diag(10,10,100,10,10,10,1..), R = rho I); `term_scale` replaces alpha
    # by scale * x0' P x0 so that the row is ACTIVE at the solution (the reference then leans on HPIPM's LQ fall-back)
    "term_rti_n20_inactive": dict(N=20, solver_type=o.SQP_RTI, surf="random", nsteps=60, seed=51, term_rho=100.0),
    "term_rti_n20_active": dict(N=20, solver_type=o.SQP_RTI, surf="random", nsteps=60, seed=52, term_rho=100.0, term_scale=0.9),
    "term_sqp_n10_active": dict(N=10, solver_type=o.SQP, surf="random", nsteps=25, seed=53, term_rho=1000.0, term_scale=0.5),
    "term_sqp_n12_inactive": dict(N=12, solver_type=o.SQP, surf="random", nsteps=25, seed=54, term_rho=100.0),
}


def make_case(name, c):
    rng = np.random.default_rng(c["seed"]) if c["seed"] is not None else None
    if isinstance(c["surf"], str):
        surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal(6)
        z0 = np.concatenate([o.README_Q0 + rng.uniform(-0.05, 0.05, 6), o.README_QD0 + rng.uniform(-0.1, 0.1, 6)])
    else:
        surf = np.array([-0.15, 0.15, -0.01, 0.01, 0.01, 0.0]) if c["surf"] is None else np.array(c["surf"], float)
        z0 = np.concatenate([o.README_Q0, o.README_QD0])
    kw = {k: c[k] for k in ("w_u", "w_qddot") if k in c}
    term = {}
    if "term_rho" in c:
        import importlib

        tc = importlib.import_module("robotic_mpc_b200.terminal_constraint")
        P, alpha = tc.default_terminal_set(wcv=200.0, Ts=0.01, rho=c["term_rho"])
        if "term_scale" in c:
            alpha = c["term_scale"] * float(z0 @ P @ z0)
        term = dict(term_P=P, term_alpha=alpha, term_lh=-alpha)
    p = o.make_problem(N=c["N"], solver_type=c["solver_type"], surf=surf, **kw, **term)
    r = o.RefLeg().closed_loop(p, z0, c["nsteps"], full_traj=True)
    keep = [0, 1, c["nsteps"] // 2, c["nsteps"] - 1]
    return dict(N=c["N"], solver_type=c["solver_type"], surf=surf, z0=z0, nsteps=c["nsteps"], w_u=kw.get("w_u", 0.01),
                w_qddot=kw.get("w_qddot", 0.02), u=r["u"], z=r["z"], status=r["status"], sqp_iter=r["sqp_iter"],
                qp_iter=r["qp_iter"], res=r["res"], cost=r["cost"], traj_steps=np.array(keep),
                xtraj=r["xtraj"][keep], utraj=r["utraj"][keep], pitraj=r["pitraj"][keep], **term)


if __name__ == "__main__":
    out = Path(__file__).resolve().parent
    only = sys.argv[1:]  # optional: regenerate just the named cases
    for name, c in CASES.items():
        if only and name not in only:
            continue
        d = make_case(name, c)
        np.savez_compressed(out / f"{name}.npz", **d)
        print(name, "nfail", int((d["status"] != 0).sum()), "qp_iter mean", d["qp_iter"].mean())
