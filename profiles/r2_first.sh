#!/bin/bash
# round-2 first GPU pass: the existing GPU suite on the refactored kernel (layout objects, warm start, new entry points),
# the generic-shape path, a quick timing of the bench shape
out=gpurun_out/r2a; mkdir -p $out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > $out/pytest.log; cat $out/pytest.log
python profiles/dev_time.py 2>&1 | tail -2 | tee $out/dev_time.log
python - <<'PY' 2>&1 | tee $out/generic.log
import torch, numpy as np, time
from g1_locomotion_b200 import BatchedMPC, presets, inputs
for N, nc in ((12, 4), (30, 4), (7, 8), (10, 4)):
    pr = presets.get_preset("cpp_node", horizon=N, num_contacts=nc)
    s = BatchedMPC(pr, device=0)
    ins = list(inputs.random_batch(pr, 512, seed=3, device="cuda"))
    u, x, info = s.update_batch(*ins, x_current=ins[0])
    torch.cuda.synchronize()
    print(N, nc, s.kernel_info(512), "status", torch.unique(info["status"]).tolist(), "iters %.2f" % info["iters"].double().mean().item(),
          "rd/g max %.2e" % float((info["resid"][:, 1]).max()))
# warm start on the config-1 replay
pr = presets.get_preset("cpp_node"); s = BatchedMPC(pr, device=0)
z = np.load("tests/golden/config1_replay.npz")
T = z["x0"].shape[0]
dev = lambda a: torch.as_tensor(a, dtype=torch.float64).cuda()
ins = [dev(z[k]) for k in ("x0", "x_ref", "contact", "c_horizon", "p_com")]
uc, xc, ic = s.update_batch(*ins, x_current=ins[0])
uprev = torch.cat([uc[:1], uc[:-1]])
uw, xw, iw = s.update_batch(*ins, x_current=ins[0], u_warm=uprev, warm_shift=1)
torch.cuda.synchronize()
ustar = dev(z["u_star"])
print("cold iters mean %.2f max %d err %.2e | warm iters mean %.2f p50 %d max %d err %.2e status %s" % (
    ic["iters"].double().mean(), ic["iters"].max(), (uc - ustar).abs().max(), iw["iters"][1:].double().mean(), iw["iters"][1:].double().median(), iw["iters"][1:].max(),
    (uw - ustar).abs().max(), torch.unique(iw["status"]).tolist()))
PY
