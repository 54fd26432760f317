import sys, time, numpy as np, torch
sys.path.insert(0, '.')
import robotic_mpc_b200 as rm
from bench import make_workload
B, N, K = 4096, 100, 600
surf, z0 = make_workload(B, 44)
par = rm.pack_params(B, 0.01, surf=surf)
eng = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=0)
eng.set_params(par); 
for chunk in (1, 2, 5, 20, 100, 600):
    eng.set_state(z0); eng.reset(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K // chunk): eng.closed_loop(chunk)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    s = eng.get("summary").cpu().numpy()
    print(f"chunk {chunk:4d}: {ms:9.1f} ms  {B*K/ms*1e3:10.0f} steps/s   qp iters/step {s[:, rm.RS['QP_ITERS']].sum()/(B*K):.3f} fails {int(s[:, rm.RS['NFAIL']].sum())}", flush=True)
