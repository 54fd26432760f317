"""Closed loop of 512 bench instances x 60 steps; saves the applied controls and (status, sqp_iter, qp_iter) for a bit-for-bit A/B of two
library builds:  RMPC_LIB=a.so python profiles/tools/bitwise_run.py out_a; RMPC_LIB=b.so python ... out_b; numpy.array_equal."""
import sys, os, numpy as np
sys.path.insert(0,'.')
sys.path.insert(0,'profiles/tools')
from run_configs import workload
import robotic_mpc_b200 as rm
B,N,K=512,100,60
surf,z0=workload(B,44)
eng=rm.Engine(B,N,"SQP_RTI",dt=0.01,nlog=K); eng.set_params(rm.pack_params(B,0.01,surf=surf)); eng.set_state(z0); eng.reset(); eng.closed_loop(K)
np.save(sys.argv[1], eng.get("ulog").cpu().numpy()); np.save(sys.argv[1]+'_s', eng.get("slog").cpu().numpy()[:,:,:3])
