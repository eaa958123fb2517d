"""Like run_solve.py for another shape (ncu captures of the long-horizon / global-workspace kernel):
    python profiles/run_solve_shape.py N nc B [launches]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from g1_locomotion_b200 import BatchedMPC, presets, inputs  # noqa: E402

N, nc, B = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
L = int(sys.argv[4]) if len(sys.argv) > 4 else 2
preset = presets.get_preset("cpp_node", horizon=N, num_contacts=nc)
s = BatchedMPC(preset, device=0)
ins = list(inputs.random_batch(preset, B, seed=11, device="cuda"))
for _ in range(L):
    u, x, info = s.update_batch(*ins, x_current=ins[0])
torch.cuda.synchronize()
print("ok", N, nc, B, s.kernel_info(B), float(info["iters"].double().mean()), torch.bincount(info["status"]).tolist())
