"""Robustness sweep of the solve kernel over many random problems per shape (no oracle: size-independent properties only):
every problem solved, duality gap and primal violation at tolerance, iteration histogram.
    python profiles/stress.py [problems_per_shape_in_units_of_65536]      -> one JSON line per shape"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from g1_locomotion_b200 import BatchedMPC, presets, inputs  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 8
SHAPES = [("cpp_node", 10, 4, 65536, reps), ("python_default", 15, 4, 65536, max(1, reps // 2)), ("cpp_node", 20, 4, 32768, max(1, reps // 2)),
          ("cpp_node", 8, 4, 65536, max(1, reps // 2)), ("cpp_node", 5, 4, 65536, max(1, reps // 2)),
          ("cpp_node", 10, 8, 32768, max(1, reps // 2)), ("cpp_node", 12, 8, 32768, max(1, reps // 2)), ("cpp_node", 40, 4, 4096, max(1, reps // 2)),
          ("cpp_node", 20, 8, 8192, max(1, reps // 2)), ("cpp_node", 40, 8, 2048, max(1, reps // 2)),
          # shapes without a compile-time instantiation (run-time layout) and the global-workspace kernel
          ("cpp_node", 12, 4, 32768, max(1, reps // 4)), ("python_default", 7, 4, 32768, max(1, reps // 4)), ("cpp_node", 30, 4, 8192, max(1, reps // 4)),
          ("cpp_node", 25, 8, 4096, max(1, reps // 4))]
for pname, N, nc, B, n_rep in SHAPES:
    preset = presets.get_preset(pname, horizon=N, num_contacts=nc)
    s = BatchedMPC(preset, device=0)
    hist = torch.zeros(64, dtype=torch.int64, device="cuda")
    bad, gap, pv, rd, total, rd_rel, flagged, warm_bad, warm_it = 0, 0.0, 0.0, 0.0, 0, 0.0, 0, 0, 0.0
    bad_seeds = []
    for r in range(n_rep):
        seed = 1000 * N + 10 * nc + r
        ins = list(inputs.random_batch(preset, B, seed=seed, device="cuda"))
        u, x, info = s.update_batch(*ins, x_current=ins[0])
        st, res = info["status"], info["resid"]
        flagged += int((st == 4).sum())          # gap converged, dual residual above dual_tol: reported as such (G1MPC_DUAL_RESIDUAL)
        st = torch.where(st == 4, torch.zeros_like(st), st)
        nb = int((st != 0).sum())
        if nb:
            bad_seeds.append({"seed": seed, "problems": torch.nonzero(st != 0).flatten()[:8].tolist(),
                              "status": st[st != 0][:8].tolist()})
        bad += nb; total += B
        ok = st == 0
        gap = max(gap, float(res[ok, 0].max())); rd = max(rd, float(res[ok, 1].max())); pv = max(pv, float(res[ok, 2].max()))
        # dual residual relative to the gradient scale of the problem (OSQP's dual tolerance is 1e-3 (1 + |g|_inf) or looser):
        # the reference-form gradient g of the 8 problems with the largest residual, from the emit kernel
        worst = torch.topk(torch.where(ok, res[:, 1], torch.zeros_like(res[:, 1])), 8).indices
        g = s.emit_qp(*[t[worst].contiguous() for t in ins], which=("g",))["g"]
        rd_rel = max(rd_rel, float((res[worst, 1] / (1.0 + g.abs().amax(dim=1))).max()))
        hist += torch.bincount(info["iters"].clamp(max=63).long(), minlength=64)
        if r == 0:      # warm start from another problem's solution (a bad start): must still solve, the safeguard bounds the cost
            uw, _, iw = s.update_batch(*ins, x_current=ins[0], u_warm=u.roll(1, 0).contiguous(), warm_shift=1, want_resid=False)
            warm_bad += int(((iw["status"] != 0) & (iw["status"] != 4)).sum())
            warm_it = float((iw["iters"] - info["iters"]).max())
        assert torch.isfinite(u[ok]).all() and torch.isfinite(x[ok]).all()
    h = hist.cpu().tolist()
    print(json.dumps({"preset": pname, "N": N, "nc": nc, "problems": total, "kernel": s.kernel_info(B), "unsolved": bad, "unsolved_where": bad_seeds,
                      "flagged_dual_residual": flagged, "warm_from_foreign_solution": {"unsolved": warm_bad, "max_extra_iters": warm_it},
                      "gap_max": gap, "dual_resid_max": rd, "dual_resid_over_1_plus_g_inf_max": rd_rel, "primal_violation_max": pv,
                      "iters_mean": sum(i * c for i, c in enumerate(h)) / total, "iters_max": max(i for i, c in enumerate(h) if c),
                      "iters_hist": {str(i): c for i, c in enumerate(h) if c}}), flush=True)
    s.close()
