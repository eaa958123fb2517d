import os, sys
sys.path.insert(0, "/root/repo")
import torch
from g1_locomotion_b200 import BatchedMPC, presets, inputs
def timed(s, ins, reps=3):
    for _ in range(2): s.update_batch(*ins, x_current=ins[0], want_resid=False)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): s.update_batch(*ins, x_current=ins[0], want_resid=False)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps
for pname, N in (("cpp_node", 12), ("python_default", 7), ("cpp_node", 16), ("cpp_node", 14)):
    pr = presets.get_preset(pname, horizon=N)
    for B in (1024, 16384):
        ins = list(inputs.random_batch(pr, B, seed=11, device="cuda"))
        row = {}
        for name, kw in (("one", dict(layout=1)), ("two", dict(layout=2)), ("auto", {})):
            s = BatchedMPC(pr, device=0, **kw)
            row[name] = (B / (timed(s, ins) * 1e-3), s.kernel_info(B)["ctas_per_sm"])
        print(pname, N, B, {k: (round(v[0]), v[1]) for k, v in row.items()})
