"""Two problems per warp from the hybrid workspace (experiment) against what the library runs today.
    python profiles/r2_hybrid2.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
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
for pname, N, nc, Bs in (("cpp_node", 10, 4, (1024, 4096, 65536)), ("python_default", 15, 4, (1024, 4096, 32768)), ("cpp_node", 10, 8, (1024, 4096, 16384)), ("cpp_node", 12, 8, (4096, 16384))):
    pr = presets.get_preset(pname, horizon=N, num_contacts=nc)
    solvers = {"auto": BatchedMPC(pr, device=0)}
    try:
        solvers["hybrid2"] = BatchedMPC(pr, device=0, workspace=3, layout=2)
    except RuntimeError as e:      # (the bench-shape instantiation was dropped from the library after this measurement)
        print(pname, N, nc, "hybrid2:", str(e)[:80]); continue
    print(pname, N, nc, {k: v.kernel_info(1 << 20) for k, v in solvers.items()})
    for B in Bs:
        ins = list(inputs.random_batch(pr, B, seed=11, device="cuda"))
        res = {}
        for name, s in solvers.items():
            ms, out = timed(s, ins)
            res[name] = out
            print("   B %6d %-8s %9.3f ms  %9.0f problems/s  iters %.1f  %s" % (B, name, ms, B / (ms * 1e-3), out[2]["iters"].double().mean(), s.kernel_info(B)["mode"][:40]))
        a, b = res["auto"], res["hybrid2"]
        print("      bit for bit:", torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]) and torch.equal(a[2]["iters"], b[2]["iters"]))
        del ins
