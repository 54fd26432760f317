import numpy as np, sys
sys.path.insert(0,'.')
import robotic_mpc_b200 as rm
from pathlib import Path
for p in sorted(Path('tests/golden').glob('term*.npz')):
    g=np.load(p)
    N, ns = int(g["N"]), int(g["nsteps"])
    solver = "SQP_RTI" if int(g["solver_type"]) == 1 else "SQP"
    par = rm.pack_params(1, 0.01, surf=g["surf"][None], w_u=float(g["w_u"]), w_qddot=float(g["w_qddot"]))
    eng = rm.Engine(1, N, solver, dt=0.01, nlog=ns)
    eng.set_params(par); eng.set_terminal(g["term_P"], float(g["term_alpha"]), float(g["term_lh"])); eng.set_state(g["z0"][None]); eng.reset(); eng.closed_loop(ns)
    u=eng.get("ulog").cpu().numpy()[0]; sl=eng.get("slog").cpu().numpy()[0]; fl=eng.get("flags").cpu().numpy()[0]
    print(p.stem, 'u rel %.2e'%(np.abs(u-g["u"]).max()/np.abs(g["u"]).max()), 'qp equal', np.array_equal(sl[:,2].astype(int), g["qp_iter"]), 'max diff', np.abs(sl[:,2]-g["qp_iter"]).max(), 'lq', fl[2])
