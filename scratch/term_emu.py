import sys; sys.path.insert(0,'.')
import numpy as np
import robotic_mpc_b200.terminal_constraint as tc
import robotic_mpc_b200.params as pm
from oracle import pyoracle as o
from tests import emu
rho=float(sys.argv[1]) if len(sys.argv)>1 else 100.0
scale=float(sys.argv[2]) if len(sys.argv)>2 else None
B=2
rng=np.random.default_rng(3)
surf=np.array([-0.1,0.1,-0.01,0.01,0.01,0.0])+0.01*rng.standard_normal((B,6))
z0=np.concatenate([o.README_Q0+rng.uniform(-0.05,0.05,(B,6)),o.README_QD0+rng.uniform(-0.1,0.1,(B,6))],1)
par=pm.pack_params(B,0.01,surf=surf)
P,a=tc.default_terminal_set(rho=rho)
port=o.PortLeg(); ref=o.RefLeg()
for solver,N,ns in ((1,10,12),(0,8,6),(1,33,5)):
    term=np.zeros((B,146))
    al=[]
    for i in range(B):
        ai = a if scale is None else scale*(z0[i]@P@z0[i])
        al.append(ai)
        term[i,:144]=P.ravel(); term[i,144]=-ai; term[i,145]=ai
    r=emu.closed_loop(par,z0,N,solver,ns,seed=1,term=term)
    for i in range(B):
        p=o.make_problem(N=N,solver_type=solver,surf=surf[i],term_P=P,term_alpha=al[i])
        g=port.closed_loop(p,z0[i],ns)
        gr=ref.closed_loop(p,z0[i],ns)
        print(solver,N,i,'emu-port u',np.abs(g['u']-r['u'][i]).max(),'emu-ref u',np.abs(gr['u']-r['u'][i]).max(),'status',g['status'][:4],r['slog'][i,:4,0],'qp',g['qp_iter'][:6],gr['qp_iter'][:6],r['slog'][i,:6,2].astype(int),'lq',r['flags'][i,2])
