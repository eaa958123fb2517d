"""Smallest command that exercises the hot path, for `ncu` captures (profiles/README.md):
    python profiles/run_solve.py [B] [launches] [walking|random]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from g1_locomotion_b200 import BatchedMPC, presets, inputs  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
L = int(sys.argv[2]) if len(sys.argv) > 2 else 4
kind = sys.argv[3] if len(sys.argv) > 3 else "walking"
preset = presets.get_preset("cpp_node")
s = BatchedMPC(preset, device=0)
if kind == "walking":
    ins = [torch.as_tensor(a, dtype=torch.float64).cuda() for a in inputs.walking_batch(preset, B)]
else:
    ins = list(inputs.random_batch(preset, B, seed=7, device="cuda"))
for _ in range(L):
    u, x, info = s.update_batch(*ins, x_current=ins[0])
torch.cuda.synchronize()
assert int((info["status"] != 0).sum()) == 0
print("ok", B, L, kind, float(info["iters"].double().mean()))
