"""A/B aid: solve fixed batches with the library named by G1MPC_B200_LIB and dump every output (bit-level comparison of
two builds), plus kernel-alone timings.   python profiles/ab_dump.py out.npz"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from g1_locomotion_b200 import BatchedMPC, presets, inputs
preset = presets.get_preset("cpp_node")
out = {}
walk = [torch.as_tensor(a, dtype=torch.float64).cuda() for a in inputs.walking_batch(preset, 1024)]
rnd = list(inputs.random_batch(preset, 4097, seed=5, device="cuda"))
for layout in (1, 2):
    s = BatchedMPC(preset, device=0, layout=layout)
    for name, ins in (("walk", walk), ("rnd", rnd)):
        u, x, info = s.update_batch(*ins, x_current=ins[0], want_multipliers=True)
        for k, v in (("u", u), ("x", x), ("it", info["iters"]), ("st", info["status"]), ("res", info["resid"]), ("y", info["y"])):
            out[f"{name}{layout}_{k}"] = v.cpu().numpy()
z = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "config1_replay.npz"))
rep = [torch.as_tensor(z[k], dtype=torch.float64).cuda() for k in ("x0", "x_ref", "contact", "c_horizon", "p_com")]
s = BatchedMPC(preset, device=0)
u, x, info = s.update_batch(*rep, x_current=rep[0])
prev = torch.cat([u[:1], u[:-1]]).contiguous()
uw, xw, iw = s.update_batch(*rep, x_current=rep[0], u_warm=prev, warm_shift=1)
out["rep_u"], out["rep_it"], out["repw_u"], out["repw_it"] = u.cpu().numpy(), info["iters"].cpu().numpy(), uw.cpu().numpy(), iw["iters"].cpu().numpy()
np.savez(sys.argv[1], **out)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
big = list(inputs.random_batch(preset, 65536, seed=7, device="cuda"))
for name, ins, reps in (("B=1 replay tick 50", [t[50:51].contiguous() for t in rep], 200), ("B=1024 walking", walk, 60), ("B=65536 random", big, 5)):
    for _ in range(3):
        s.update_batch(*ins, x_current=ins[0])
    ts = []
    for k in range(reps):
        if ins[0].shape[0] == 1024:
            flush.fill_(k & 0xff)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(300000)
        e0.record(); s.update_batch(*ins, x_current=ins[0], want_resid=False); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    print("%-20s p50 %9.1f us -> %.3f M problems/s" % (name, np.median(ts), ins[0].shape[0] / np.median(ts)))
