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


def callers_golden():
    """Two CALLERS of the path run from the reference's own sources (oracle/ref_loader.py says how):
    wbid_refs.npz    -- WholeBodyID.setReference (g1_mujoco_sim/src/wbid.py:231-297) on 96 random (x_opt[1], u_opt[0], foot positions),
                        what it hands its OpenSoT tasks in the order of g1mpc_wbid_refs_batch;
    ref_main_loop.npz -- g1_mpc/scripts/main.py main() as shipped (python defaults, planner (1000, 0.25, 0.1, 0.2, 100), 4 s), first 40
                        ticks: per tick the inputs main() handed the class and the u_opt / x_opt it got back through the OSQP shim, plus
                        the x_opt[0] / u_opt[0] history main() records."""
    from oracle import node_oracle
    W = ref_loader.load_wbid()
    rng = np.random.default_rng(20251105)
    B = 96
    x1 = np.concatenate([rng.uniform(-0.4, 0.4, (B, 3)), rng.uniform(-2, 2, (B, 3)), rng.uniform(-1.5, 1.5, (B, 6)), np.full((B, 1), -9.80665)], 1)
    u0 = np.stack([rng.uniform(-80, 80, (B, 4)), rng.uniform(-80, 80, (B, 4)), rng.uniform(0, 400, (B, 4))], 2).reshape(B, 12)
    feet = x1[:, None, 3:6] + rng.uniform(-0.3, 0.3, (B, 4, 3))
    mass = 35.112
    refs = np.stack([ref_loader.reference_wbid_set_reference(W, x1[b], u0[b], feet[b], mass) for b in range(B)])
    mine = np.stack([node_oracle.wbid_refs_flat(x1[b], u0[b], feet[b], mass) for b in range(B)])
    assert np.array_equal(refs, mine), np.abs(refs - mine).max()        # the restatement IS the reference code, bit for bit
    np.savez_compressed(os.path.join(OUT, "wbid_refs.npz"), x_opt1=x1, u_opt0=u0, feet=feet, mass=mass, refs=refs)
    print("wbid_refs", refs.shape)
    tx, tu, ticks = ref_loader.run_reference_main(max_ticks=40, per_tick=True)
    np.savez_compressed(os.path.join(OUT, "ref_main_loop.npz"), planner=np.array([1000., 0.25, 0.1, 0.2, 100.]), total_duration=4.0,
                        foot_offset=0.08, traj_x=tx, traj_u=tu, **{k: np.stack([t[k] for t in ticks]) for k in ticks[0]})
    print("ref_main_loop", tx.shape, {k: np.stack([t[k] for t in ticks]).shape for k in ticks[0]})
    # SURVEY 8f-4, the simulator's (Python) halves of the message exchange, from g1_mujoco_sim/src/ros_run_simulation.py itself:
    # publish_current_state (:22-78) and G1MujocoSimulation.callback_mpc_solution (:188-218) on stand-in message objects
    from g1_locomotion_b200 import wire
    sim = ref_loader.load_ros_simulation()
    K, N = 24, 10
    x0 = rng.normal(size=(K, 13)); x0[:, 12] = -9.80665
    fpos = rng.normal(size=(K, 4, 3)); qf = rng.normal(0, 100, (K, 12))
    pub = [ref_loader.reference_publish_current_state(sim, x0[k, 0:3], x0[k, 3:6], x0[k, 6:9], x0[k, 9:12], fpos[k], qf[k], 0.05 * k) for k in range(K)]
    assert all(p["n_states"] == 1 and p["trajectory_index"] == 0 and p["frame_id"] == "SRBD" for p in pub)
    x_opt = rng.normal(size=(K, N, 13)); u_opt = rng.normal(0, 100, (K, N, 12)); c_hor = rng.normal(size=(K, N, 12))
    contact = (rng.uniform(size=(K, N, 4)) > 0.4).astype(float); landing = rng.normal(size=(K, 2))
    cb = [ref_loader.reference_callback_mpc_solution(sim, ref_loader.msg_from_dict(
        wire.pack_mpc_solution(x_opt[k], u_opt[k], c_hor[k], contact[k], landing[k], sim_time=0.05 * k))) for k in range(K)]
    np.savez_compressed(os.path.join(OUT, "wire_simulator.npz"), x0=x0, foot_positions=fpos, qp_forces=qf,
                        pub_state=np.stack([p["state"] for p in pub]), pub_positions=np.stack([p["positions"] for p in pub]),
                        pub_forces=np.stack([p["forces"] for p in pub]), pub_names=np.array(pub[0]["names"]),
                        x_opt=x_opt, u_opt=u_opt, c_horizon=c_hor, contact=contact, landing=landing,
                        cb_x_opt=np.stack([c[0] for c in cb]), cb_u_opt0=np.stack([c[1] for c in cb]),
                        cb_contact_states=np.array([c[2] for c in cb]), cb_landing=np.stack([c[3] for c in cb]))
    print("wire_simulator", K, "cases")


if __name__ == "__main__":
    import sys
    if len(sys.argv) > 1 and sys.argv[1] == "callers":
        callers_golden()
    else:
        main()
        gait_golden()
        callers_golden()
