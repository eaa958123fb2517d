#!/bin/bash
# compute-sanitizer over a small subset of the GPU tests: memcheck (out-of-bounds / misaligned), racecheck (shared-memory
# hazards: the solve kernel aliases shared-memory regions by lifetime and synchronises with __syncwarp() only), synccheck
# (divergent barriers / shuffles: the two-problems-per-warp layout keeps control flow warp-uniform by construction).
#   bash profiles/sanitize.sh [outdir]      -> <outdir>/sanitize_{memcheck,racecheck,synccheck}.log + summary on stdout
# NOTE (round 2): compute-sanitizer is CLOSED on this GPU pool ("runs under it have left GPUs needing a reset"): the tool prints
# that notice and exits 86 without running anything (gpurun_out/r2e/sanitize_memcheck.log).  The script stays for a machine
# where the tool is available; on this pool the evidence is tests/test_round2_gpu.py::test_no_write_lands_outside_the_output_arrays
# (guard bands), the bit-identity of the two lane layouts / the two workspaces, and run-to-run determinism.
out=${1:-gpurun_out/sanitize}; mkdir -p $out
cat > /tmp/sanitize_case.py <<'PY'
import sys, numpy as np, torch
sys.path.insert(0, ".")
from g1_locomotion_b200 import BatchedMPC, presets, inputs
for layout, B in ((1, 6), (2, 7)):
    pr = presets.get_preset("cpp_node")
    s = BatchedMPC(pr, device=0, layout=layout)
    ins = [torch.as_tensor(a, dtype=torch.float64).cuda() for a in inputs.walking_batch(pr, B)]
    u, x, info = s.update_batch(*ins, x_current=ins[0], want_multipliers=True)
    uw, _, iw = s.update_batch(*ins, x_current=ins[0], u_warm=u, warm_shift=0)
    junk = torch.full_like(u, 900.0)
    uj, _, ij = s.update_batch(*ins, x_current=ins[0], u_warm=junk, warm_shift=1)          # exercises the cold restart
    s.rollout(ins[1], ins[3], ins[4], u, x_current=ins[0])
    s.emit_qp(*[t[:1] for t in ins])
    torch.cuda.synchronize()
    print("layout", layout, "status", info["status"].tolist(), "iters", info["iters"].tolist(), iw["iters"].tolist(), ij["iters"].tolist())
pr = presets.get_preset("cpp_node", horizon=12, num_contacts=8)
for ws in (1, 2):                                              # run-time layout, shared and global workspace
    s = BatchedMPC(presets.get_preset("cpp_node", horizon=13), device=0, workspace=ws)
    ins = list(inputs.random_batch(presets.get_preset("cpp_node", horizon=13), 5, seed=1, device="cuda"))
    _, _, info = s.update_batch(*ins, x_current=ins[0])
    torch.cuda.synchronize()
    print("N=13 workspace", ws, info["status"].tolist(), info["iters"].tolist())
s = BatchedMPC(presets.get_preset("cpp_node"), device=0)
x0 = torch.zeros(3, 13, dtype=torch.float64, device="cuda"); x0[:, 3] = 0.05; x0[:, 5] = 0.598; x0[:, 12] = -9.80665
cur = torch.tensor([[0.02, 0.1185, 0, 0.19, 0.1185, 0, 0.02, -0.1185, 0, 0.19, -0.1185, 0]] * 3, dtype=torch.float64, device="cuda")
r = s.closed_loop(x0, cur, torch.full((3,), 1e-9, dtype=torch.float64, device="cuda"), 3, warm_start=True)
pl = torch.tensor([[0.2, 0.3, 0.05, 0.15, 30.0]] * 3, dtype=torch.float64, device="cuda")
m = s.main_loop(pl, 3, total_duration=0.15)
torch.cuda.synchronize()
print("loops", r["status_any"].tolist(), m["status_any"].tolist())
PY
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python /tmp/sanitize_case.py > $out/sanitize_$tool.log 2>&1
  echo "== $tool rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $out/sanitize_$tool.log | tail -1)"
  grep -E "^layout|^N=13|^loops" $out/sanitize_$tool.log | head -6
done
