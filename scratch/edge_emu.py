import sys; sys.path.insert(0,'.')
import numpy as np
from tests import emu
import oracle.pyoracle as po
import robotic_mpc_b200 as rm
z0 = np.concatenate([po.README_Q0, po.README_QD0])
zbad = z0.copy(); zbad[1] = 2*np.pi + 0.3   # outside the joint box
zbad2 = z0.copy(); zbad2[6] = 9.0            # joint speed far above the input box (state is not bounded on qd)
par = rm.pack_params(1, 0.01)
for name, z in (("q out of box", zbad), ("qd large", zbad2)):
  for st, sname in ((1,'RTI'),(0,'SQP')):
    for N in (2, 10):
        e = emu.closed_loop(par, z[None], N, st, 6)
        leg = po.RefLeg()
        o = leg.closed_loop(po.make_problem(N=N, solver_type=po.SQP_RTI if st==1 else po.SQP), z, 6)
        print(name, sname, N, 'status ref', o['status'], 'dev', e['slog'][0,:,0].astype(int), 'sqp', o['sqp_iter'], e['slog'][0,:,1].astype(int),
              'du', np.abs(e['u'][0]-o['u']).max())
