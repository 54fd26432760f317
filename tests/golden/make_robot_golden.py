"""Golden closed-loop vectors for a chain other than the UR10, generated with the REFERENCE leg of the
oracle: the reference's compiled acados + HPIPM + BLASFEO driven by oracle/ref_driver.c, model functions evaluated on the UR5
chain of R/ur_description/urdf/ur5.urdf (frame tool0, R/prediction_model.py:28-29).  The chain table stored in each file is
read from that URDF here, so the test also pins robotic_mpc_b200.robots' built-in table to the file.

    python tests/golden/make_robot_golden.py        (needs /root/reference and `make -C oracle -f ref.mk ref-driver`)
"""
import importlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT))
from oracle import pyoracle as o  # noqa: E402

rb = importlib.import_module("robotic_mpc_b200.robots")
URDF = Path("/root/reference/ur_description/urdf/ur5.urdf")
Q0 = np.array([0.0, -1.8, 1.9, -1.7, -1.57, 0.0])
QD0 = np.array([0.2, 0.1, -0.1, 0.0, 0.0, 0.0])
CASES = {
    "ur5_rti_n20": dict(N=20, solver_type=o.SQP_RTI, nsteps=80, seed=61),
    "ur5_sqp_n15": dict(N=15, solver_type=o.SQP, nsteps=30, seed=62),
    "ur5_rti_n50_term": dict(N=50, solver_type=o.SQP_RTI, nsteps=40, seed=63, term_rho=100.0),
}


def make_case(c):
    rng = np.random.default_rng(c["seed"])
    surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal(6)
    z0 = np.concatenate([Q0 + rng.uniform(-0.05, 0.05, 6), QD0 + rng.uniform(-0.1, 0.1, 6)])
    xyzR, axis = rb.chain_from_urdf(URDF, "tool0")
    term = {}
    if "term_rho" in c:
        tc = importlib.import_module("robotic_mpc_b200.terminal_constraint")
        P, alpha = tc.default_terminal_set(wcv=200.0, Ts=0.01, rho=c["term_rho"])
        term = dict(term_P=P, term_alpha=alpha, term_lh=-alpha)
    p = o.make_problem(N=c["N"], solver_type=c["solver_type"], surf=surf, **term)
    leg = o.RefLeg()
    leg.set_chain(xyzR, axis)
    try:
        r = leg.closed_loop(p, z0, c["nsteps"], full_traj=True)
    finally:
        leg.set_chain(None)
    keep = [0, 1, c["nsteps"] // 2, c["nsteps"] - 1]
    return dict(N=c["N"], solver_type=c["solver_type"], surf=surf, z0=z0, nsteps=c["nsteps"], chain_xyzR=xyzR, chain_axis=axis,
                u=r["u"], z=r["z"], status=r["status"], sqp_iter=r["sqp_iter"], qp_iter=r["qp_iter"], res=r["res"], cost=r["cost"],
                traj_steps=np.array(keep), xtraj=r["xtraj"][keep], utraj=r["utraj"][keep], pitraj=r["pitraj"][keep], **term)


if __name__ == "__main__":
    out = Path(__file__).resolve().parent / "robots"
    out.mkdir(exist_ok=True)
    for name, c in CASES.items():
        d = make_case(c)
        np.savez_compressed(out / f"{name}.npz", **d)
        print(name, "nfail", int((d["status"] != 0).sum()), "sqp_iter max", d["sqp_iter"].max(), "qp_iter mean", d["qp_iter"].mean(),
              "move", np.abs(d["z"][-1] - d["z"][0]).max())
