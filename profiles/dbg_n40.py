import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
from g1_locomotion_b200 import BatchedMPC, presets, inputs
for N, nc in ((40, 4), (40, 8), (20, 4), (10, 4)):
    preset = presets.get_preset("cpp_node", horizon=N, num_contacts=nc)
    ins = list(inputs.random_batch(preset, 64, seed=11, device="cuda"))
    s = BatchedMPC(preset, device=0, max_iter=60)
    u, x, info = s.update_batch(*ins, x_current=ins[0])
    torch.cuda.synchronize()
    it = info["iters"].cpu().numpy(); st = info["status"].cpu().numpy(); r = info["resid"].cpu().numpy()
    order = np.argsort(-r[:, 1])
    print(N, nc, "rd quantiles", np.quantile(r[:, 1], [0, .5, .9, 1]), "worst problems", order[:5], r[order[:5], 1], "iters", it[order[:5]], "status", st[order[:5]])
