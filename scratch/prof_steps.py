"""Profile helper: B=4096, N=100 SQP_RTI; `nlaunch` launches of `chunk` closed-loop steps each (ncu picks one by -s)."""
import sys, os, numpy as np, torch
sys.path.insert(0, '.')
chunk = int(sys.argv[1]); nlaunch = int(sys.argv[2]) if len(sys.argv) > 2 else 4
os.environ["RMPC_STEPS_PER_LAUNCH"] = str(chunk)
import robotic_mpc_b200 as rm
from bench import make_workload
B, N = 4096, 100
surf, z0 = make_workload(B, 44)
eng = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=0)
eng.set_params(rm.pack_params(B, 0.01, surf=surf)); eng.set_state(z0); eng.reset()
for _ in range(nlaunch):
    eng.closed_loop(chunk)
torch.cuda.synchronize()
print("ok", eng.launch_count)
