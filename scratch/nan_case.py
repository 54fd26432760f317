import sys; sys.path.insert(0, ".")
import numpy as np, torch
import oracle.pyoracle as o
import robotic_mpc_b200 as rm
B, N, ns = 5, 10, 5
rng = np.random.default_rng(78)
surf = np.array([-0.1, 0.1, -0.01, 0.01, 0.01, 0.0]) + 0.01 * rng.standard_normal((B, 6))
z0 = np.concatenate([o.README_Q0 + rng.uniform(-0.05, 0.05, (B, 6)), o.README_QD0 + rng.uniform(-0.1, 0.1, (B, 6))], 1)
z0[4, 3] = np.nan
eng = rm.Engine(B, N, sys.argv[1] if len(sys.argv) > 1 else "SQP_RTI", dt=0.01, nlog=ns)
eng.set_params(rm.pack_params(B, 0.01, surf=surf)); eng.set_state(z0); eng.reset()
eng.closed_loop(1)
torch.cuda.synchronize()
print(eng.get("slog").cpu().numpy()[:, 0, :3])
