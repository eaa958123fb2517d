"""Kernel variants side by side: one warp per problem (layout 1), two problems per warp (2), partner-warp mode with one (3) and
two (4) problems per main warp.  Bits against layout 1 and time per launch over a range of batch sizes.  python profiles/r2_lat2.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from g1_locomotion_b200 import BatchedMPC, presets, inputs
pr = presets.get_preset("cpp_node")
S = {k: BatchedMPC(pr, device=0, layout=k) for k in (1, 2, 3, 4)}
walk = [torch.as_tensor(a, dtype=torch.float64).cuda() for a in inputs.walking_batch(pr, 1024)]
rnd = list(inputs.random_batch(pr, 65536, seed=7, device="cuda"))
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ref = S[1].update_batch(*[t[:4097].contiguous() for t in rnd], x_current=rnd[0][:4097].contiguous(), want_multipliers=True)
for k in (2, 3, 4):
    r = S[k].update_batch(*[t[:4097].contiguous() for t in rnd], x_current=rnd[0][:4097].contiguous(), want_multipliers=True)
    print("layout", k, "bit-identical to layout 1:", torch.equal(ref[0], r[0]) and torch.equal(ref[1], r[1]) and all(torch.equal(ref[2][q], r[2][q]) for q in ("status", "iters", "resid", "y")))
for name, src, Bs in (("walking", walk, (148, 296, 444, 592, 1024)), ("random", rnd, (1024, 2048, 4096, 16384, 65536))):
    for B in Bs:
        ins = [t[:B].contiguous() for t in src]
        line = "%-8s B %6d:" % (name, B)
        for k in (1, 2, 3, 4):
            s = S[k]
            for _ in range(3): s.update_batch(*ins, x_current=ins[0], want_resid=False)
            torch.cuda.synchronize(); ts = []
            for i in range(12 if B <= 4096 else 4):
                if B <= 4096: flush.fill_(i)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda._sleep(300000)
                e0.record(); s.update_batch(*ins, x_current=ins[0], want_resid=False); e1.record(); e1.synchronize()
                ts.append(e0.elapsed_time(e1) * 1e3)
            line += "  L%d %9.1f us (%.3f M/s)" % (k, np.median(ts), B / np.median(ts))
        print(line)
