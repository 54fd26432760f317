import sys; sys.path.insert(0,'.')
import numpy as np
import robotic_mpc_b200.terminal_constraint as tc
from oracle import pyoracle as o
x0=np.concatenate([o.README_Q0,o.README_QD0])
ref=o.RefLeg(); port=o.PortLeg()
for rho,scale in ((100.0,None),(100.0,0.9),(10.0,None),(1000.0,0.5)):
    P,a=tc.default_terminal_set(rho=rho)
    h0=x0@P@x0
    if scale is not None: a=scale*h0
    for st in (o.SQP_RTI,o.SQP):
        for N in (10,20):
            pr=o.make_problem(N=N,solver_type=st,term_P=P,term_alpha=a)
            r=ref.closed_loop(pr,x0,30,full_traj=True)
            q=port.closed_loop(pr,x0,30,full_traj=True)
            eu=np.abs(r['u']-q['u']).max()/np.abs(r['u']).max()
            ex=np.abs(r['xtraj']-q['xtraj']).max()
            print(rho,scale,st,N,'u rel',eu,'x abs',ex,'status eq',(r['status']==q['status']).all(),'sqp eq',(r['sqp_iter']==q['sqp_iter']).all(),'qp eq',(r['qp_iter']==q['qp_iter']).all(), r['qp_iter'][:5], q['qp_iter'][:5])
