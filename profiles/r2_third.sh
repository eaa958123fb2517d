#!/bin/bash
out=gpurun_out/r2c; mkdir -p $out
python -m pytest tests -m gpu -q -k "runtime_shape or illconditioned or largest_dual or fleet" 2>&1 | grep -E "^E  |FAILED|passed|failed|Error" | cut -c1-260 | head -60 > $out/pytest.log; cat $out/pytest.log
python - <<'PY' 2>&1 | tee $out/rd_diag.log
import torch, numpy as np
from g1_locomotion_b200 import BatchedMPC, presets, inputs
# relative dual residual at the stop, by shape (scale = 1 + |g|inf from the emit kernel)
for N, nc, B in ((10, 4, 8192), (15, 4, 4096), (20, 4, 4096), (40, 4, 2048), (10, 8, 2048), (20, 8, 1024), (40, 8, 512), (30, 4, 1024), (64, 4, 256)):
    pr = presets.get_preset("cpp_node", horizon=N, num_contacts=nc)
    s = BatchedMPC(pr, device=0, dual_tol=1e30)
    ins = list(inputs.random_batch(pr, B, seed=17, device="cuda"))
    u, x, info = s.update_batch(*ins, x_current=ins[0])
    g = torch.cat([s.emit_qp(*[t[i:i + 128] for t in ins], which=("g",))["g"].abs().amax(1) for i in range(0, B, 128)])
    rel = info["resid"][:, 1] / (1 + g)
    q = torch.quantile(rel, torch.tensor([0.5, 0.99, 1.0], dtype=torch.float64, device="cuda")).tolist()
    print("N %2d nc %d: status %s iters %.1f  rd/(1+|g|) p50 %.1e p99 %.1e max %.1e  frac>1e-6 %.4f frac>1e-8 %.4f" % (
        N, nc, torch.unique(info["status"]).tolist(), info["iters"].double().mean(), q[0], q[1], q[2], float((rel > 1e-6).double().mean()), float((rel > 1e-8).double().mean())))
PY
