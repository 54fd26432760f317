import sys; sys.path.insert(0,'.')
import numpy as np, ctypes as C
import robotic_mpc_b200.terminal_constraint as tc
from oracle import pyoracle as o
x0=np.concatenate([o.README_Q0,o.README_QD0])
P,a=tc.default_terminal_set(rho=100.0); a=0.9*(x0@P@x0)
N=int(sys.argv[1]); st=int(sys.argv[2]); nstep=int(sys.argv[3])
pr=o.make_problem(N=N,solver_type=st,term_P=P,term_alpha=a)
ref=o.RefLeg().lib; port=o.PortLeg().lib
dp=C.POINTER(C.c_double)
ref.rmpc_ref_create.restype=C.c_void_p; port.rmpc_oracle_create.restype=C.c_void_p
z=np.ascontiguousarray(x0)
cr=C.c_void_p(ref.rmpc_ref_create(C.byref(pr),z.ctypes.data_as(dp)))
cp=C.c_void_p(port.rmpc_oracle_create(C.byref(pr),z.ctypes.data_as(dp)))
mod=o.Model(port)
for k in range(nstep):
    ur=np.zeros(6); up=np.zeros(6); sq=C.c_int(); qi=C.c_int(); res=np.zeros(4); cost=C.c_double(); tt=C.c_double()
    s1=ref.rmpc_ref_step(cr,z.ctypes.data_as(dp),ur.ctypes.data_as(dp),C.byref(sq),C.byref(qi),res.ctypes.data_as(dp),C.byref(cost),C.byref(tt))
    q1=qi.value
    s2=port.rmpc_oracle_step(cp,z.ctypes.data_as(dp),up.ctypes.data_as(dp),C.byref(sq),C.byref(qi),res.ctypes.data_as(dp),C.byref(cost))
    print('step',k,'status',s1,s2,'qp',q1,qi.value,'du',np.abs(ur-up).max())
    if k==nstep-1:
        stat=np.zeros((64,32)); sm=C.c_int()
        it=ref.rmpc_ref_last_qp_stat(cr,stat.ctypes.data_as(dp),64,C.byref(sm))
        stat=stat.ravel()[:64*sm.value].reshape(-1,sm.value) if sm.value else stat
        ps=np.zeros((64,8)); pit=port.rmpc_oracle_last_qp_stat(cp,ps.ctypes.data_as(dp),64)
        np.set_printoptions(linewidth=250,precision=6)
        print('ref it',it); print(stat[1:it+1,[0,1,2,3,5,6,7,8,12,14]])
        print('port it',pit); print(ps[1:pit+1])
    z=np.array(mod.plant_rk4(np.full(6,200.0),0.01,z,ur))
