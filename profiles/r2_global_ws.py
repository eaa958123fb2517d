"""Shared-memory workspace against global-memory workspace (persistent grid) on the long-horizon shapes: bits and throughput.
    python profiles/r2_global_ws.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from g1_locomotion_b200 import BatchedMPC, presets, inputs
def timed(s, ins, reps=3):
    for _ in range(2):
        out = s.update_batch(*ins, x_current=ins[0], want_resid=False)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        s.update_batch(*ins, x_current=ins[0], want_resid=False)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, out
for N, nc, Bs in ((20, 4, (1024, 8192)), (15, 4, (4096, 16384)), (10, 8, (4096, 16384)), (20, 8, (1024, 8192)), (40, 4, (512, 4096)), (40, 8, (256, 4096)), (30, 4, (1024, 4096)), (100, 8, (64, 1024))):
    pr = presets.get_preset("cpp_node", horizon=N, num_contacts=nc)
    solvers = {}
    for name, ws in (("shared", 1), ("global", 2)):
        try:
            solvers[name] = BatchedMPC(pr, device=0, workspace=ws)
        except RuntimeError as e:
            print("N %d nc %d %s: %s" % (N, nc, name, str(e)[:90]))
    for B in Bs:
        ins = list(inputs.random_batch(pr, B, seed=11, device="cuda"))
        res = {}
        for name, s in solvers.items():
            ms, out = timed(s, ins)
            res[name] = out
            st = torch.bincount(out[2]["status"], minlength=5).tolist()
            print("N %3d nc %d B %5d %-6s %9.3f ms  %9.0f problems/s  iters %.1f status %s  %s" % (N, nc, B, name, ms, B / (ms * 1e-3), out[2]["iters"].double().mean(), st, s.kernel_info(B)))
        if len(res) == 2:
            same = torch.equal(res["shared"][0], res["global"][0]) and torch.equal(res["shared"][1], res["global"][1]) and torch.equal(res["shared"][2]["iters"], res["global"][2]["iters"])
            print("    bit-identical:", same)
        del ins
