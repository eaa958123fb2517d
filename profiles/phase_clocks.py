"""Per-phase cycle counts of the solve kernel from a -DG1MPC_PHASE_CLOCKS build (development aid):
    G1MPC_B200_LIB=gpurun_in/libtiming.so python profiles/phase_clocks.py [B]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from g1_locomotion_b200 import BatchedMPC, presets, inputs
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
preset = presets.get_preset("cpp_node")
s = BatchedMPC(preset, device=0, layout=int(sys.argv[2]) if len(sys.argv) > 2 else 0)
ins = [torch.as_tensor(a, dtype=torch.float64).cuda() for a in inputs.walking_batch(preset, max(B, 8))]
ins = [t[:B].contiguous() for t in ins]
for _ in range(3):
    u, x, info = s.update_batch(*ins, x_current=ins[0], want_multipliers=True)
torch.cuda.synchronize()
ph = info["y"][:, :16].cpu().numpy(); it = info["iters"].cpu().numpy()
names = ["gradient", "passA", "assemble+S", "factor", "solve: misc", "passB+B2+assemble", "passC+step", "setup/loop",
         "solve: pre", "solve: backward", "solve: b", "solve: forward", "solve: output", "prologue (whole solve)", "epilogue (whole solve)"]
if os.environ.get("G1MPC_B200_LIB", "").endswith("libftiming.so"):      # -DG1MPC_FACTOR_CLOCKS: 8..12 split the factor stage
    names[3] = "factor: P22 update"; names[4] = "solves (both)"
    names[8:13] = ["factor: loads+inv6", "factor: ldl6", "factor: columns", "factor: P12 S col", "factor: stores"]
print("B =", B, "iters", it[:8], "total cycles (problem 0): %d" % ph[0].sum())
for i, n in enumerate(names):
    print("  %-20s %9.0f cycles/iter  %5.1f%%" % (n, (ph[:, i] / it).mean(), 100 * ph[:, i].sum() / ph.sum()))
