"""Hybrid workspace (per-stage region + scratch in shared memory, per-contact region in a global-memory slot per resident warp,
persistent grid) against the shared-memory and the global-memory workspace: bits and throughput.
    python profiles/r2_hybrid_ws.py"""
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
shapes = ((10, 8, (4096, 16384)), (12, 8, (4096, 16384)), (20, 8, (1024, 8192)), (40, 8, (256, 4096)), (15, 4, (4096, 16384)), (20, 4, (1024, 8192)),
          (40, 4, (512, 4096)), (30, 4, (1024, 4096)), (16, 8, (2048, 8192)), (60, 8, (256, 2048)))
for N, nc, Bs in shapes:
    pr = presets.get_preset("cpp_node", horizon=N, num_contacts=nc)
    solvers = {}
    for name, ws in (("shared", 1), ("global", 2), ("hybrid", 3)):
        try:
            solvers[name] = BatchedMPC(pr, device=0, workspace=ws, layout=1)
        except RuntimeError as e:
            print("N %d nc %d %s: %s" % (N, nc, name, str(e)[:100]))
    for B in Bs:
        ins = list(inputs.random_batch(pr, B, seed=11, device="cuda"))
        res = {}
        for name, s in solvers.items():
            ms, out = timed(s, ins)
            res[name] = out
            st = torch.bincount(out[2]["status"], minlength=5).tolist()
            ki = s.kernel_info(B)
            print("N %3d nc %d B %5d %-6s %9.3f ms  %9.0f problems/s  iters %.1f status %s  smem %d ctas/SM %d" % (N, nc, B, name, ms, B / (ms * 1e-3), out[2]["iters"].double().mean(), st, ki["smem_bytes"], ki["ctas_per_sm"]))
        names = list(res)
        for other in names[1:]:
            a, b = res[names[0]], res[other]
            same = torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]) and torch.equal(a[2]["iters"], b[2]["iters"]) and torch.equal(a[2]["status"], b[2]["status"])
            print("    %s == %s bit for bit: %s" % (names[0], other, same))
        del ins
