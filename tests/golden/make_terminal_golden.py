"""Golden vectors for the terminal-set construction, produced by the REFERENCE's own Python functions.

R/_terminal_constraint.py cannot be imported as a module here (it imports the reference's pinocchio-based model at
import time), but the functions on the path are pure numpy/scipy: this script parses the file, compiles ONLY the
function definitions `discrete_lqr ... tune_Q_diagonal_for_state_slacks` from the reference source where it lies
(/root/reference, nothing is copied into the repo) and evaluates them on the UR10 inputs of BASELINE configs[3].
The committed sets/terminal_set.npz is what travels.

    python tests/golden/make_terminal_golden.py
"""
import ast
import sys
from pathlib import Path

import numpy as np

REF = Path("/root/reference/_terminal_constraint.py")
WANT = ["discrete_lqr", "build_constraints", "alpha_from_constraints_over_ellipsoide", "ellipsoid_metric_J",
        "terminal_ellipsoid_from_lqr", "tune_Q_diagonal_for_state_slacks"]


def load_reference_functions():
    tree = ast.parse(REF.read_text())
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in WANT]
    mod = ast.Module(body=body, type_ignores=[])
    from scipy.linalg import solve_discrete_are

    ns = {"np": np, "solve_discrete_are": solve_discrete_are}
    exec(compile(mod, str(REF), "exec"), ns)
    return ns


def zoh(wcv, Ts):
    w = np.broadcast_to(np.asarray(wcv, float), (6,))
    a22 = np.exp(-w * Ts)
    a12 = (1 - a22) / w
    Ad = np.block([[np.eye(6), np.diag(a12)], [np.zeros((6, 6)), np.diag(a22)]])
    Bd = np.vstack([np.diag(Ts - a12), np.diag(1 - a22)])
    return Ad, Bd


if __name__ == "__main__":
    f = load_reference_functions()
    qdm = np.array([2.16, 2.16, np.pi, 3.20, 3.20, 3.20])
    q_min, q_max = np.full(6, -2 * np.pi), np.full(6, 2 * np.pi)
    Cx, dx, Cu, du = f["build_constraints"](q_min, q_max, -qdm, qdm, nx=12, nu=6)
    Qd = np.array([10.0, 10.0, 100.0, 10.0, 10.0, 10.0] + [1.0] * 6)
    out = dict(Cx=Cx, dx=dx, Cu=Cu, du=du, Q_diag=Qd)
    cases = [(200.0, 0.01, 1.0), (200.0, 0.01, 100.0), (np.array([228.9, 262.09, 517.3, 747.44, 429.9, 1547.76]), 0.0005, 37.0)]
    for k, (wcv, Ts, rho) in enumerate(cases):
        Ad, Bd = zoh(wcv, Ts)
        r = f["terminal_ellipsoid_from_lqr"](Ad, Bd, np.diag(Qd), rho, Cx, dx, Cu, du, safety=0.98)
        out.update({f"c{k}_wcv": np.broadcast_to(np.asarray(wcv, float), (6,)).copy(), f"c{k}_Ts": Ts, f"c{k}_rho": rho,
                    f"c{k}_Ad": Ad, f"c{k}_Bd": Bd, f"c{k}_K": r["K"], f"c{k}_P": r["P"], f"c{k}_alpha": r["alpha"],
                    f"c{k}_alpha_i": r["alpha_i"], f"c{k}_idx_active": r["idx_active"], f"c{k}_J": r["J"],
                    f"c{k}_slack": r["slack"]})
    Ad, Bd = zoh(200.0, 0.01)
    Qt, hist = f["tune_Q_diagonal_for_state_slacks"](Ad, Bd, Qd, 100.0, Cx, dx, Cu, du, 6, n_iter=5)
    out.update(tune_Q=Qt, tune_alpha=np.array([h["alpha"] for h in hist]), tune_J=np.array([h["J"] for h in hist]))
    np.savez_compressed(Path(__file__).resolve().parent / "sets" / "terminal_set.npz", **out)
    print("alpha", [float(out[f"c{k}_alpha"]) for k in range(3)], "tuned Q", Qt[:6])
