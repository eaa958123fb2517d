#!/bin/bash
out=gpurun_out/r2b; mkdir -p $out
python -m pytest tests -m gpu -q 2>&1 | tail -60 > $out/pytest.log; cat $out/pytest.log
python - <<'PY' 2>&1 | tee $out/fleet_diag.log
import torch, numpy as np
from g1_locomotion_b200 import BatchedMPC, presets
preset = presets.get_preset("cpp_node"); s = BatchedMPC(preset, device=0)
B, T, dt = 1200, 5, preset["dt"]
rng = np.random.default_rng(9)
x0 = np.zeros((B, 13))
x0[:, 3] = 0.05 + 0.01 * np.arange(B); x0[:, 4] = 7.44339342e-05; x0[:, 5] = 5.97983255e-01; x0[:, 9] = 0.35
x0[:, 12] = np.float32(-9.80665); x0[:, :12] += rng.normal(0, 0.003, (B, 12))
cur = np.zeros((B, 4, 3)); cur[:, :, 0] = x0[:, 3:4] + np.array([-0.0293, 0.1406, -0.0293, 0.1406]); cur[:, :, 1] = np.array([0.118508, 0.118508, -0.118508, -0.118508])
t0 = (rng.integers(0, 30, B) + 0.5) * dt
d = lambda a: torch.as_tensor(a, dtype=torch.float64).cuda()
X0, CUR, TT = d(x0), d(cur.reshape(B, 12)), d(t0)
xo = torch.zeros(B, 10, 13, dtype=torch.float64, device="cuda"); xr = torch.zeros_like(xo)
for t in range(T):
    xr, ct, ch, pc = s.walk_inputs(X0, CUR, TT, xo, x_ref=xr)
    u, xo, info = s.update_batch(X0, xr, ct, ch, pc, x_current=X0)
    bad = torch.nonzero(info["status"] != 0).ravel().tolist()
    print("tick", t, "bad", bad, [(int(info["status"][b]), int(info["iters"][b]), "%.2e" % float(info["resid"][b, 1]), "gap %.1e" % float(info["resid"][b, 0])) for b in bad[:8]])
    X0 = xo[:, 1].contiguous(); CUR = ch[:, 1].contiguous(); TT = TT + dt
PY
