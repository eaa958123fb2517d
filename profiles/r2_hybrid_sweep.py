"""Batch-size sweep of the three workspace variants (shared / global / hybrid) on the long shapes: where each wins.
    python profiles/r2_hybrid_sweep.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from g1_locomotion_b200 import BatchedMPC, presets, inputs
def timed(s, ins, reps=3):
    for _ in range(2):
        s.update_batch(*ins, x_current=ins[0], want_resid=False)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        s.update_batch(*ins, x_current=ins[0], want_resid=False)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps
Bs = (64, 148, 296, 444, 592, 888, 1184, 2048, 4096, 8192)
for N, nc in ((20, 8), (40, 8), (40, 4), (30, 4), (60, 8), (24, 8), (16, 8), (20, 4)):
    pr = presets.get_preset("cpp_node", horizon=N, num_contacts=nc)
    solvers = {}
    for name, ws in (("shared", 1), ("global", 2), ("hybrid", 3), ("auto", 0)):
        try:
            solvers[name] = BatchedMPC(pr, device=0, workspace=ws, layout=1)
        except RuntimeError as e:
            pass
    print("N %d nc %d  ctas/SM: %s" % (N, nc, {k: v.kernel_info(1 << 20)["ctas_per_sm"] for k, v in solvers.items()}))
    full = list(inputs.random_batch(pr, max(Bs), seed=11, device="cuda"))
    for B in Bs:
        ins = [t[:B].contiguous() for t in full]
        row = {name: B / (timed(s, ins) * 1e-3) for name, s in solvers.items()}
        best = max((k for k in row if k != "auto"), key=row.get)
        print("   B %5d  " % B + "  ".join("%s %8.0f" % (k, v) for k, v in row.items()) + "   best: %s, auto runs: %s" % (best, solvers["auto"].kernel_info(B)["mode"].split()[0]))
