import sys; sys.path.insert(0,'.')
import numpy as np, time
import robotic_mpc_b200.terminal_constraint as tc
from oracle import pyoracle as o
sys.path.insert(0,'profiles/tools')
from run_configs import workload
surf,z0=workload(4,45)
P,a=tc.default_terminal_set(rho=100.0)
ref=o.RefLeg()
for N in (20,50,200):
  for term in (False,True):
    for i in range(2):
        kw=dict(term_P=P,term_alpha=a) if term else {}
        pr=o.make_problem(N=N,solver_type=o.SQP,surf=surf[i],**kw)
        t=time.time(); r=ref.closed_loop(pr,z0[i],3); t=time.time()-t
        print('N',N,'term',term,'inst',i,'status',r['status'],'sqp',r['sqp_iter'],'qp',r['qp_iter'],'res',r['res'][0], 'sec',round(t,2))
