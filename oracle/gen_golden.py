"""Generate tests/golden/build_*.npz from the reference's own g1_mpc/scripts/mpc.py (imported
unmodified via oracle/ref_loader.py).  Run in the build container only:

    python -m oracle.gen_golden

Each file holds the inputs of one problem, the parameter preset, and every build-half output of the
reference (A, Bd, Aqp, Bqp, H, g, Acon, l, u, psi) plus two rollouts (full / first-step-only) of a seeded
random u_opt through the reference's compute_rollout (mpc.py:379-391).
"""
import json
import os
import numpy as np
import torch

from g1_locomotion_b200 import presets, inputs
from oracle import ref_loader

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")

CASES = [
    # name, preset, overrides, generator, seed
    ("cpp_walk", "cpp_node", {}, "walk", 13),
    ("cpp_random", "cpp_node", {}, "random", 7),
    ("py_random_n15", "python_default", {}, "random", 3),
    ("py_stand_n10", "python_default", {"horizon": 10}, "stand", 5),
    ("cpp_random_n12_nc8", "cpp_node", {"horizon": 12, "num_contacts": 8}, "random", 11),
    ("cpp_fractional_contact", "cpp_node", {}, "fractional", 17),
]


# cost_scale 1: the Python class has no x2 (mpc.py:301,317); the cpp_node constants are applied to it as in ref_loader
UPDATE_CASES = [
    ("py_default_n15", "python_default", {}, "random", 23),
    ("cpp_consts_walk", "cpp_node", {"cost_scale": 1.0}, "walk", 14),
    ("cpp_consts_walk_b", "cpp_node", {"cost_scale": 1.0}, "walk", 27),
]


def make_inputs(kind, preset, seed):
    if kind == "walk":
        return inputs.walking_problem(seed, np.random.default_rng(seed), preset)
    x0, xr, ct, ch, pc = [t[1].numpy() for t in inputs.random_batch(preset, 2, seed=seed)]
    if kind == "stand":
        ct = np.ones_like(ct)
    if kind == "fractional":
        rng = np.random.default_rng(seed)
        ct = ct * rng.uniform(0.2, 1.0, ct.shape)
    return x0, xr, ct, ch, pc


def main():
    os.makedirs(OUT, exist_ok=True)
    for name, pname, over, kind, seed in CASES:
        preset = presets.get_preset(pname, **over)
        x0, xr, ct, ch, pc = make_inputs(kind, preset, seed)
        ref = ref_loader.make_reference_mpc(preset)
        out = ref_loader.reference_build(ref, preset["cost_scale"], x0, xr, ct, ch, pc)
        rng = np.random.default_rng(seed + 100)
        u = rng.uniform(-50, 200, (preset["horizon"], 3 * preset["num_contacts"]))
        x_full = ref_loader.reference_rollout(ref, u, x0, False)
        x_one = ref_loader.reference_rollout(ref, u, x0, True)
        np.savez_compressed(os.path.join(OUT, f"build_{name}.npz"), preset=json.dumps(preset), x0=x0, x_ref=xr,
                            contact=ct, c_horizon=ch, p_com=pc, u_rollout=u, x_opt_full=x_full, x_opt_one=x_one,
                            **out)
        print(name, {k: v.shape for k, v in out.items() if hasattr(v, "shape")})
    # The reference's MPC.update() run END TO END (mpc.py:360-377; solve_qp through oracle/osqp_shim.py, i.e. the OSQP
    # restatement with default settings): pins the reference's own control flow -- the hand-over of P, q, A, l, u, the
    # row-major reshape of result.x (mpc.py:352) and compute_rollout after a solve (mpc.py:379-386) -- for the three
    # call patterns its callers use: update(x_current, one_rollout=False) (run_simulation.py:106), one_rollout=True
    # (mpc_to_wbid.py:158) and x_current=None.
    for name, pname, over, kind, seed in UPDATE_CASES:
        preset = presets.get_preset(pname, **over)
        x0, xr, ct, ch, pc = make_inputs(kind, preset, seed)
        ref = ref_loader.make_reference_mpc(preset)
        res = {}
        for tag, xc, one in (("full", x0, False), ("one", x0, True), ("nocur", None, False)):
            ref.x_opt = np.zeros((preset["horizon"], 13))
            u, x = ref_loader.reference_update(ref, x0, xr, ct, ch, pc, x_current=xc, one_rollout=one)
            res["u_opt_" + tag], res["x_opt_" + tag] = u, x
        np.savez_compressed(os.path.join(OUT, f"ref_update_{name}.npz"), preset=json.dumps(preset), x0=x0, x_ref=xr,
                            contact=ct, c_horizon=ch, p_com=pc, H=ref.Hessian.copy(), g=ref.gradient.ravel().copy(), **res)
        print("update", name, res["u_opt_full"].shape)


def gait_golden():
    """GaitPlanner.plan_gait of the REFERENCE class (gait_planner.py, imported unmodified) mapped onto (contact, c_horizon)
    by main.py:259-285 (restated in oracle/node_oracle.map_gait_phases -- the mapping lives inside main() and cannot be called),
    for a few planners and query times, N = 10 and 15."""
    from oracle import node_oracle
    GP = ref_loader.load_gait_planner()
    planners = [(1000., 0.25, 0.1, 0.2, 100), (0.2, 0.8, 0.3, 0.2, 10), (0.5, 0.8, 0.3, 0.2, 10), (0.13, 0.07, 0.033, 0.11, 37),
                (0.0, 0.3, 0.25, 0.15, 5), (0.4, 0.1, 0.05, 0.1, 0)]
    rows = []
    for pl in planners:
        phases = GP(pl[0], pl[1], pl[2], pl[3], int(pl[4])).plan_gait()
        ours = node_oracle.plan_gait(*pl)
        assert len(ours) == len(phases)
        for a, b in zip(ours, phases):      # the restated planner is the reference planner, bit for bit
            assert a["start_time"] == b["start_time"] and a["end_time"] == b["end_time"] and a["support_leg"] == b["support_leg"]
            assert np.array_equal(a["left_foot"], b["left_foot"]) and np.array_equal(a["right_foot"], b["right_foot"])
        span = phases[-1]["end_time"]
        for t in list(np.linspace(0.0, min(span, 12.0) * 1.1, 23)) + [phases[1]["start_time"] if len(phases) > 1 else 0.0, -0.5]:
            for N, dt in ((10, 0.05), (15, 0.04)):
                ct, ch = node_oracle.map_gait_phases(phases, float(t), N, dt, 0.08)
                rows.append(dict(planner=np.array(pl, float), t=float(t), N=N, dt=dt, contact=ct, c_horizon=ch))
    np.savez_compressed(os.path.join(OUT, "gait_plan.npz"), planner=np.stack([r["planner"] for r in rows]), t=np.array([r["t"] for r in rows]),
                        N=np.array([r["N"] for r in rows]), dt=np.array([r["dt"] for r in rows]),
                        contact=np.stack([np.pad(r["contact"], ((0, 15 - r["N"]), (0, 0))) for r in rows]),
                        c_horizon=np.stack([np.pad(r["c_horizon"], ((0, 15 - r["N"]), (0, 0))) for r in rows]))
    print("gait_plan", len(rows), "cases")


if __name__ == "__main__":
    main()
    gait_golden()
