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


if __name__ == "__main__":
    main()
