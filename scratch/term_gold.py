import sys; sys.path.insert(0,'.')
import numpy as np
import robotic_mpc_b200.params as pm
from oracle import pyoracle as o
from tests import emu
from pathlib import Path
for f in sorted(Path('tests/golden').glob('term_*.npz')):
    g=np.load(f)
    N,ns,st=int(g['N']),int(g['nsteps']),int(g['solver_type'])
    par=pm.pack_params(1,0.01,surf=g['surf'][None],w_u=float(g['w_u']),w_qddot=float(g['w_qddot']))
    term=np.concatenate([g['term_P'].ravel(),[float(g['term_lh'])],[float(g['term_alpha'])]])[None]
    r=emu.closed_loop(par,g['z0'][None],N,st,ns,seed=1,term=term)
    p=o.make_problem(N=N,solver_type=st,surf=g['surf'],term_P=g['term_P'],term_alpha=float(g['term_alpha']),term_lh=float(g['term_lh']))
    q=o.PortLeg().closed_loop(p,g['z0'],ns)
    rel=lambda a,b: np.abs(a-b).max()/np.abs(b).max()
    print(f.stem,'EMU u',rel(r['u'][0],g['u']),'status eq',np.array_equal(r['slog'][0,:,0].astype(int),g['status']),'sqp eq',np.array_equal(r['slog'][0,:,1].astype(int),g['sqp_iter']),'qp maxdiff',np.abs(r['slog'][0,:,2]-g['qp_iter']).max(),'lq',r['flags'][0,2],'cost',rel(r['slog'][0,:,3],g['cost']))
    print(f.stem,'PORT u',rel(q['u'],g['u']),'status eq',np.array_equal(q['status'],g['status']),'sqp eq',np.array_equal(q['sqp_iter'],g['sqp_iter']),'qp maxdiff',np.abs(q['qp_iter']-g['qp_iter']).max(), 'status', g['status'][:8], 'sqp', g['sqp_iter'][:8])
