import numpy as np, sys, os
sys.path.insert(0,'.'); sys.path.insert(0,'profiles/tools')
from run_configs import workload
import robotic_mpc_b200 as rm
from oracle import pyoracle as o
from multiprocessing import get_context
B,N,K,N2=2048,100,40,20
surf,z0=workload(B,44)
w=np.random.default_rng(45).uniform(0.06,0.14,(B,1))
qmin,qmax=z0[:,:6]-w,z0[:,:6]+w
eng=rm.Engine(B,N,"SQP_RTI",dt=0.01,nlog=K); eng.set_params(rm.pack_params(B,0.01,surf=surf,q_min=qmin,q_max=qmax)); eng.set_option("qp_cond_N",N2); eng.set_state(z0); eng.reset(); eng.closed_loop(K)
u=eng.get("ulog").cpu().numpy(); sl=eng.get("slog").cpu().numpy(); fl=eng.get("flags").cpu().numpy()
def ref(i, c):
    return o.RefLeg().closed_loop(o.make_problem(N=N,solver_type=o.SQP_RTI,surf=surf[i],q_min=qmin[i],q_max=qmax[i],cond_N=c),z0[i],K)
def job(i): 
    r=ref(i,N2); return np.abs(u[i]-r['u']).max(axis=1)/np.abs(r['u']).max()
with get_context("fork").Pool(16) as p: E=np.array(p.map(job, range(B)))
i=int(E.max(axis=1).argmax()); print('worst instance', i, 'per-step rel err', E[i])
r2=ref(i,N2); r0=ref(i,None)
print('ref(N2) vs ref(N) per step', np.abs(r2['u']-r0['u']).max(axis=1)/np.abs(r2['u']).max())
print('qp dev', sl[i,:,2].astype(int)); print('qp ref2', r2['qp_iter']); print('qp refN', r0['qp_iter']); print('lq flags', fl[i])
print('count >1e-7:', (E.max(axis=1)>1e-7).sum(), ' >1e-8:', (E.max(axis=1)>1e-8).sum())
# HPIPM's own per-iteration table of the FIRST QP of that instance, condensed and un-condensed (columns: alpha_aff mu_aff sigma
# alpha_prim alpha_dual mu res_stat res_eq res_ineq res_comp dual_gap obj lq_fact itref_pred itref_corr ...)
import ctypes as C
for c in (N2, None):
    s = o.RefSolver(o.make_problem(N=N, solver_type=o.SQP_RTI, surf=surf[i], q_min=qmin[i], q_max=qmax[i], cond_N=c), z0[i])
    s.step(z0[i])
    buf = np.zeros(64 * 32); sm_ = C.c_int(0)
    s.lib.rmpc_ref_last_qp_stat.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_int)]
    it = s.lib.rmpc_ref_last_qp_stat(s.h, buf.ctypes.data_as(C.POINTER(C.c_double)), 64, C.byref(sm_))
    tab = buf[:(it + 1) * sm_.value].reshape(it + 1, sm_.value)
    print('cond_N', c, 'iterations', it, 'lq_fact per iteration', tab[1:, 12].astype(int), 'itref_corr', tab[1:, 14].astype(int))
    print('   alpha_aff', np.array2string(tab[1:, 0], precision=6)); print('   mu', np.array2string(tab[1:, 5], precision=3))
    s.close()
print('instance', i)
