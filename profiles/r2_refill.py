"""Two problems per warp with refill (persistent grid; layout = 5) against the plain two-per-warp layout: bits and throughput.
    python profiles/r2_refill.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from g1_locomotion_b200 import BatchedMPC, presets, inputs
def timed(s, ins, reps=5):
    for _ in range(2):
        out = s.update_batch(*ins, x_current=ins[0], want_multipliers=True)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        s.update_batch(*ins, x_current=ins[0], want_resid=False)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, out
pr = presets.get_preset("cpp_node")
plain, refill = BatchedMPC(pr, device=0, layout=2), BatchedMPC(pr, device=0, layout=5)
walk = [torch.as_tensor(a, dtype=torch.float64).cuda() for a in inputs.walking_batch(pr, 16384)]
for name, full in (("random", list(inputs.random_batch(pr, 262144, seed=7, device="cuda"))), ("walking", walk)):
    for B in (1024, 1185, 2368, 4097, 16384, 65536, 262144):
        if B > full[0].shape[0]:
            continue
        ins = [t[:B].contiguous() for t in full]
        (ta, a), (tb, b) = timed(plain, ins), timed(refill, ins)
        same = all(torch.equal(a[i], b[i]) for i in (0, 1)) and all(torch.equal(a[2][k], b[2][k]) for k in ("status", "iters", "resid", "y"))
        print("%-8s B %6d  plain %8.3f ms %9.0f/s   refill %8.3f ms %9.0f/s  (%+.1f %%)  bit-identical %s  iters %.2f" %
              (name, B, ta, B / ta * 1e3, tb, B / tb * 1e3, 100 * (ta / tb - 1), same, float(a[2]["iters"].double().mean())))
