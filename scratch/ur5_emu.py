import sys, time; sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import numpy as np
import emu
import oracle.pyoracle as po
import robotic_mpc_b200 as rm
import robotic_mpc_b200.robots as rb
t,a = rb.chain_table('ur5')
emu.set_chain(t,a)
q = np.array([0.3,-1.2,1.4,-1.0,-1.5,0.2]); qd=np.array([0.1,0.2,-0.1,0.3,0,0.1])
par = rm.pack_params(1, 0.01)
rec,y5 = emu.plant_record(par[0], np.concatenate([q,qd]))
p,R = rb.forward_kinematics(t,a,q)
print('pose err', np.abs(rec[0:3]-p).max(), np.abs(rec[3:12].reshape(3,3)-R).max())
port = po.PortLeg(); port.set_chain(t,a)
m = po.Model(port.lib)
pj,aj,Ro,pe = m.fk(q)
print('oracle fk err', np.abs(pe-p).max(), np.abs(Ro-R).max())
# closed loop parity
q0 = np.array([0.0,-1.8,1.9,-1.7,-1.57,0.0]); 
print('start pose', rb.forward_kinematics(t,a,q0)[0])
z0 = np.concatenate([q0,[0.2,0.1,-0.1,0,0,0]])[None]
for st,name in [(1,'SQP_RTI'),(0,'SQP')]:
    prob = po.make_problem(N=10, solver_type=po.SQP_RTI if st==1 else po.SQP)
    o = port.closed_loop(prob, z0[0], 15)
    e = emu.closed_loop(par, z0, 10, st, 15)
    print(name, 'dz', np.abs(e['z'][0]-o['z']).max(), 'du', np.abs(e['u'][0]-o['u']).max(), 'status', o['status'], e['slog'][0,:,0])
    try:
        ref = po.RefLeg(); ref.set_chain(t,a)
        r = ref.closed_loop(prob, z0[0], 15)
        print('  vs ref dz', np.abs(e['z'][0]-r['z']).max(), 'du', np.abs(e['u'][0]-r['u']).max(), r['sqp_iter'], e['slog'][0,:,1])
    except FileNotFoundError as ex: print(ex)
emu.set_chain(None)
print('u', np.abs(e['u'][0]).max(axis=1)[:5], 'z move', np.abs(e['z'][0,-1]-e['z'][0,0]).max(), 'res', e['slog'][0,:3,4:8])
