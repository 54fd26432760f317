import sys; sys.path.insert(0, ".")
import numpy as np, torch
import oracle.pyoracle as o
import robotic_mpc_b200 as rm
B, N = 6, 12
rng = np.random.default_rng(81)
surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal((B, 6))
z0 = np.concatenate([o.README_Q0 + rng.uniform(-0.05, 0.05, (B, 6)), o.README_QD0 + rng.uniform(-0.1, 0.1, (B, 6))], 1)
par = rm.pack_params(B, 0.01, surf=surf)
eng = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=16)
eng.set_params(par); eng.set_state(z0); eng.reset(); eng.closed_loop(7)
blob = eng.store_state()
torch.cuda.synchronize()
print("size", blob.numel(), "hdr", blob[:4].cpu().numpy(), "next", blob[4:8].cpu().numpy())
b2 = blob.cpu().numpy()
eng2 = rm.Engine(B, N, "SQP_RTI", dt=0.01, nlog=16)
try:
    eng2.load_state(b2); print("load ok", eng2.lib.rmpc_steps_done(eng2.h))
except Exception as e: print("ERR", e)
try:
    eng2.load_state(blob); print("load ok (device tensor)")
except Exception as e: print("ERR2", e)
