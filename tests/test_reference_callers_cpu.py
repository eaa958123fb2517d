"""the two Python CALLERS of the path, pinned to the reference's own code.

tests/golden/wbid_refs.npz and tests/golden/ref_main_loop.npz were written by `python -m oracle.gen_golden callers` from
g1_mujoco_sim/src/wbid.py (WholeBodyID.setReference on recording stand-ins for its OpenSoT tasks) and from g1_mpc/scripts/main.py
(main() as shipped, headless, the reference MPC class behind the OSQP shim) -- oracle/ref_loader.py says how.  Here the numpy
restatements of oracle/node_oracle.py, which the GPU tests of g1mpc_wbid_refs_batch and g1mpc_main_loop_batch use as their checker,
are held to those files; with /root/reference present also to the live code."""
import os

import numpy as np
import pytest

from conftest import ROOT
from g1_locomotion_b200 import presets
from oracle import node_oracle


def _z(name):
    z = np.load(os.path.join(ROOT, "tests", "golden", name))
    return {k: z[k] for k in z.files}


def test_wbid_restatement_is_the_reference_set_reference_bit_for_bit():
    z = _z("wbid_refs.npz")
    B = len(z["refs"])
    assert B >= 64 and z["refs"].shape == (B, 42)
    for b in range(B):
        mine = node_oracle.wbid_refs_flat(z["x_opt1"][b], z["u_opt0"][b], z["feet"][b], float(z["mass"]))
        assert np.array_equal(mine, z["refs"][b]), b
    # the quirk the kernel reproduces: sum_forces = row sums of reshape(u_opt0, (3, 4)) (wbid.py:290), NOT the per-axis totals
    u0, m = z["u_opt0"][0], float(z["mass"])
    as_coded = u0.reshape(3, 4).sum(1) / m + np.array([0, 0, -9.80665])
    per_axis = u0.reshape(4, 3).sum(0) / m + np.array([0, 0, -9.80665])
    assert np.array_equal(z["refs"][0][27:30], as_coded) and not np.allclose(as_coded, per_axis)


def test_main_loop_restatement_follows_the_reference_main_tick_by_tick(orc):
    """The loop logic of main.py:200-326 (x0 <- x_opt[2], testing_double_support schedule, planner mapping, p_com = x_ref[:,3:6],
    x_opt[0] <- x_opt[1] carry, rollout without x_current) against the reference's own run: the restated loop is fed the reference's
    u_opt of every tick, so no solver sits between the two -- every input it builds for the next tick and every rollout must agree."""
    z = _z("ref_main_loop.npz")
    T = len(z["traj_x"])
    assert T >= 32
    preset = presets.get_preset("python_default")                        # main() constructs mpc.MPC() with its defaults
    worst = {}

    def on_tick(k, d):
        for key in ("x0", "x_ref", "contact", "c_horizon", "p_com", "x_opt"):
            worst[key] = max(worst.get(key, 0.0), float(np.abs(d[key] - z[key][k]).max()))

    tx, tu = node_oracle.main_loop(orc, preset, tuple(z["planner"]), T, total_duration=float(z["total_duration"]),
                                   foot_offset=float(z["foot_offset"]), solve=lambda k, m, contact: z["u_opt"][k], on_tick=on_tick)
    assert worst["x_ref"] == 0.0 and worst["contact"] == 0.0 and worst["c_horizon"] == 0.0 and worst["p_com"] == 0.0, worst
    assert worst["x0"] <= 1e-12 and worst["x_opt"] <= 1e-12, worst
    assert np.abs(tx - z["traj_x"]).max() <= 1e-12 and np.array_equal(tu, z["traj_u"])
    # the schedule of testing_double_support shows up inside the window (t_ref >= total_duration / 4 from tick 11 on)
    assert z["x_ref"][10][:, 4].min() == 0.1 and z["x_ref"][11][-1, 4] == 0.0
    # and closed on the exact minimiser instead of the reference's OSQP iterate (eps 1e-3) the same loop stays within what that
    # tolerance allows: millimetres in the CoM, centimetres per second in the velocities
    ex, eu = node_oracle.main_loop(orc, preset, tuple(z["planner"]), 24, total_duration=float(z["total_duration"]))
    d = np.abs(ex - z["traj_x"][:24]).max(0)
    assert d[:6].max() <= 5e-3 and d[6:12].max() <= 5e-2, d


def test_wire_shim_matches_the_simulator_own_packing_code():
    """SURVEY 8f-4, the simulator's halves: what g1_mujoco_sim/src/ros_run_simulation.py's publish_current_state (:22-78) put into
    /srbd_current and what its callback_mpc_solution (:188-218) read out of /mpc_solution (tests/golden/wire_simulator.npz, from the
    file itself on stand-in message objects) against g1_locomotion_b200.wire.  The header stamp (rospy.Time, third party) and the byte
    encoding (ROS 1 serialization rules) stay restated."""
    from g1_locomotion_b200 import wire
    z = _z("wire_simulator.npz")
    assert tuple(z["pub_names"]) == wire.CONTACT_NAMES
    for k in range(len(z["x0"])):
        m = wire.pack_srbd_current(z["x0"][k], z["foot_positions"][k], z["qp_forces"][k], sim_time=0.05 * k)
        assert len(m["states_horizon"]) == 1 and m["states_horizon"][0]["trajectory_index"] == 0 and m["header"]["frame_id"] == "SRBD"
        x0, now, tiled = wire.unpack_srbd_current(m, 10)
        assert np.array_equal(x0[:12], z["pub_state"][k][:12])
        assert x0[12] == float(np.float32(z["pub_state"][k][12]))            # State.gravity is float32 on the wire (State.msg:20)
        assert np.array_equal(now, z["pub_positions"][k]) and np.array_equal(tiled, np.tile(z["pub_positions"][k], (10, 1)))
        assert [c["name"] for c in m["contacts"]] == list(z["pub_names"])
        assert np.array_equal(np.array([[c["force"][a] for a in "xyz"] for c in m["contacts"]]).ravel(), z["pub_forces"][k])
        msg = wire.pack_mpc_solution(z["x_opt"][k], z["u_opt"][k], z["c_horizon"][k], z["contact"][k], z["landing"][k], sim_time=0.05 * k)
        x_opt, u0, act, land = wire.unpack_mpc_solution(wire.deserialize(wire.serialize(msg)))
        assert np.array_equal(x_opt, z["cb_x_opt"][k]) and np.array_equal(u0, z["cb_u_opt0"][k])
        assert np.array_equal(act, z["cb_contact_states"][k]) and np.array_equal(land, z["cb_landing"][k])
        assert np.array_equal(u0, z["u_opt"][k][0]) and np.array_equal(x_opt[:, :12], z["x_opt"][k][:, :12])


def test_live_reference_callers():
    from oracle import ref_loader
    if not ref_loader.available():
        pytest.skip("reference tree not present (GPU box)")
    z = _z("wbid_refs.npz")
    W = ref_loader.load_wbid()
    for b in range(0, len(z["refs"]), 7):
        live = ref_loader.reference_wbid_set_reference(W, z["x_opt1"][b], z["u_opt0"][b], z["feet"][b], float(z["mass"]))
        assert np.array_equal(live, z["refs"][b]), b
    m = _z("ref_main_loop.npz")
    tx, tu = ref_loader.run_reference_main(max_ticks=12)
    assert np.array_equal(tx, m["traj_x"][:12]) and np.array_equal(tu, m["traj_u"][:12])
    w = _z("wire_simulator.npz")
    sim = ref_loader.load_ros_simulation()
    pub = ref_loader.reference_publish_current_state(sim, w["x0"][3, 0:3], w["x0"][3, 3:6], w["x0"][3, 6:9], w["x0"][3, 9:12],
                                                     w["foot_positions"][3], w["qp_forces"][3], 0.15)
    assert np.array_equal(pub["state"], w["pub_state"][3]) and np.array_equal(pub["forces"], w["pub_forces"][3])
