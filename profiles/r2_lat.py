"""Latency-mode kernel (two warps per problem, layout 0 below 2 problems per SM / layout 3 forced) against the one-warp kernel
(layout 1): bits and single-solve time on the configs[0] replay.   python profiles/r2_lat.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from g1_locomotion_b200 import BatchedMPC, presets, inputs
pr = presets.get_preset("cpp_node")
z = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "config1_replay.npz"))
rep = [torch.as_tensor(z[k], dtype=torch.float64).cuda() for k in ("x0", "x_ref", "contact", "c_horizon", "p_com")]
rnd = list(inputs.random_batch(pr, 1500, seed=3, device="cuda"))
one, lat = BatchedMPC(pr, device=0, layout=1), BatchedMPC(pr, device=0, layout=3)
print("kernel_info", one.kernel_info(1), lat.kernel_info(1))
for name, ins in (("replay", rep), ("random", rnd)):
    for kw in (dict(), dict(one_rollout=True)):
        a = one.update_batch(*ins, x_current=ins[0], want_multipliers=True, **kw)
        for rep_i in range(3):
            b = lat.update_batch(*ins, x_current=ins[0], want_multipliers=True, **kw)
            same = torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]) and all(torch.equal(a[2][k], b[2][k]) for k in ("status", "iters", "resid", "y"))
            print(name, kw, "bit-identical:", same, "iters", float(a[2]["iters"].double().mean()))
prev = torch.cat([a[0][:1], a[0][:-1]]).contiguous()
aw = one.update_batch(*rep, x_current=rep[0], u_warm=torch.cat([one.update_batch(*rep, x_current=rep[0])[0][:1], one.update_batch(*rep, x_current=rep[0])[0][:-1]]).contiguous())
bw = lat.update_batch(*rep, x_current=rep[0], u_warm=torch.cat([one.update_batch(*rep, x_current=rep[0])[0][:1], one.update_batch(*rep, x_current=rep[0])[0][:-1]]).contiguous())
print("warm bit-identical:", torch.equal(aw[0], bw[0]) and torch.equal(aw[2]["iters"], bw[2]["iters"]))
for name, s in (("one warp", one), ("latency mode", lat)):
    ts = []
    for r in range(3):
        for t in range(200):
            tick = [x[t:t + 1] for x in rep]
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda._sleep(300000)
            e0.record(); s.solve_one(*tick, x_current=tick[0]); e1.record(); e1.synchronize()
            if r: ts.append(e0.elapsed_time(e1) * 1e3)
    print("%-14s replay B=1 kernel alone: p50 %.1f us  p99 %.1f us  min %.1f" % (name, np.percentile(ts, 50), np.percentile(ts, 99), min(ts)))
for B in (64, 148, 296):
    ins = [t[:B].contiguous() for t in rnd]
    for name, s in (("one warp", one), ("latency mode", lat)):
        for _ in range(3): s.update_batch(*ins, x_current=ins[0], want_resid=False)
        torch.cuda.synchronize(); ts = []
        for _ in range(30):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda._sleep(300000)
            e0.record(); s.update_batch(*ins, x_current=ins[0], want_resid=False); e1.record(); e1.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        print("B %3d %-14s p50 %.1f us" % (B, name, np.median(ts)))
