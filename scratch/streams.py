"""Does splitting the batch over concurrent streams hide straggler instances?  4096 x N=100 SQP_RTI, 600 steps."""
import json, os, sys
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "profiles/tools")
import torch
import robotic_mpc_b200 as rm
from run_configs import workload
B, N, K = 4096, 100, 600
surf, z0 = workload(B, 42 + 2)
par = rm.pack_params(B, 0.01, surf=surf)
for nchunk in [int(v) for v in sys.argv[1:]] or [1, 2, 4, 8]:
    bounds = np.linspace(0, B, nchunk + 1).astype(int)
    streams = [torch.cuda.Stream() for _ in range(nchunk)]
    engs = []
    for c in range(nchunk):
        b, e = bounds[c], bounds[c + 1]
        eng = rm.Engine(e - b, N, "SQP_RTI", dt=0.01, nlog=0)
        eng.set_params(par[b:e]); eng.set_state(z0[b:e]); eng.reset()
        engs.append(eng)
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for s in streams:
        s.wait_stream(torch.cuda.current_stream())
    for k in range(K):
        for eng, s in zip(engs, streams):
            with torch.cuda.stream(s):
                eng.closed_loop(1)
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    t1.record(); t1.synchronize()
    ms = t0.elapsed_time(t1)
    fails = sum(int(e.get("summary").cpu().numpy()[:, rm.RS["NFAIL"]].sum()) for e in engs)
    print(json.dumps({"chunks": nchunk, "steps_per_launch": os.environ.get("RMPC_STEPS_PER_LAUNCH", "1"), "ms_total": ms, "mpc_steps_per_s": B * K / ms * 1e3, "failures": fails}))
    for e in engs:
        e.close()
