"""Wire-format shim: the SRBD_state message both ways, the float32 gravity
field, the ROS 1 byte layout (restated: no ROS on this box), and the hand-over to the batched path's arrays."""
import struct

import numpy as np
import pytest

from g1_locomotion_b200 import wire, presets, inputs


def _one():
    preset = presets.get_preset("cpp_node")
    x0, x_ref, contact, c_hor, p_com = [a[3] for a in inputs.walking_batch(preset, 8)]
    return preset, x0, x_ref, contact, c_hor, p_com


def test_srbd_current_round_trip_and_float32_gravity():
    preset, x0, x_ref, contact, c_hor, p_com = _one()
    x0 = x0.copy(); x0[12] = -9.80665                       # what the simulator writes (ros_run_simulation.py:58)
    feet = c_hor[0].reshape(4, 3)
    msg = wire.pack_srbd_current(x0, feet, qp_forces=np.arange(12.0), sim_time=1.25)
    assert [c["name"] for c in msg["contacts"]] == list(wire.CONTACT_NAMES)
    assert msg["header"]["frame_id"] == "SRBD" and msg["header"]["stamp"] == {"secs": 1, "nsecs": 250000000}
    got = wire.deserialize(wire.serialize(msg))
    assert got == msg
    x0b, now, tiled = wire.unpack_srbd_current(got, preset["horizon"])
    assert np.array_equal(x0b[:12], x0[:12])
    assert x0b[12] == float(np.float32(-9.80665)) and x0b[12] != -9.80665      # State.msg:20 -- the node sees the float32 value
    assert np.array_equal(now, feet.ravel()) and tiled.shape == (preset["horizon"], 12) and np.array_equal(tiled[7], now)


def test_mpc_solution_round_trip():
    preset, x0, x_ref, contact, c_hor, p_com = _one()
    N = preset["horizon"]
    rng = np.random.default_rng(1)
    x_opt = rng.normal(size=(N, 13)); x_opt[:, 12] = np.float32(-9.80665)
    u_opt = rng.normal(size=(N, 12))
    msg = wire.pack_mpc_solution(x_opt, u_opt, c_hor, contact, (0.31, -0.12), sim_time=0.0425)
    assert len(msg["states_horizon"]) == N and [s["trajectory_index"] for s in msg["states_horizon"]] == list(range(N))
    assert msg["landing_position"] == {"x": 0.31, "y": -0.12, "z": 0.0}                # z never travels (mpc_to_wbid_node.cpp:376)
    xo, u0, act, lp = wire.unpack_mpc_solution(wire.deserialize(wire.serialize(msg)))
    assert np.array_equal(xo, x_opt) and np.array_equal(u0, u_opt[0])
    assert np.array_equal(act, contact[0] != 0) and np.array_equal(lp, [0.31, -0.12, 0.0])


def test_byte_layout_of_a_minimal_message():
    """Hand-assembled ROS 1 encoding: Header(seq, secs, nsecs, frame_id), State[] , ContactPoint[], Point."""
    msg = {"header": {"seq": 7, "stamp": {"secs": 2, "nsecs": 5}, "frame_id": "SRBD"},
           "states_horizon": [wire._state(0, np.arange(13.0))],
           "contacts": [{"name": "ab", "position": wire._vec((1, 2, 3)), "force": wire._vec((4, 5, 6)), "active": True}],
           "landing_position": wire._vec((9, 8, 0))}
    want = struct.pack("<3I", 7, 2, 5) + struct.pack("<I", 4) + b"SRBD"
    want += struct.pack("<I", 1) + struct.pack("<i", 0) + struct.pack("<12d", *np.arange(12.0)) + struct.pack("<f", 12.0)
    want += struct.pack("<I", 1) + struct.pack("<I", 2) + b"ab" + struct.pack("<6d", 1, 2, 3, 4, 5, 6) + b"\x01"
    want += struct.pack("<3d", 9, 8, 0)
    got = wire.serialize(msg)
    assert got == want and len(got) == 16 + 4 + 4 + (4 + 96 + 4) + 4 + (6 + 48 + 1) + 24
    with pytest.raises(ValueError):
        wire.deserialize(got[:-1])
    with pytest.raises(ValueError):
        wire.deserialize(got + b"\x00")


def test_batch_helpers_feed_the_batched_path_shapes():
    preset = presets.get_preset("cpp_node")
    B, N = 5, preset["horizon"]
    batch = inputs.walking_batch(preset, B)
    msgs = [wire.pack_srbd_current(batch[0][b], batch[3][b][0].reshape(4, 3), sim_time=0.04 * b) for b in range(B)]
    x0, now, t = wire.unpack_srbd_current_batch([wire.deserialize(wire.serialize(m)) for m in msgs], N)
    assert x0.shape == (B, 13) and now.shape == (B, 12) and np.allclose(t, 0.04 * np.arange(B), atol=1e-9)
    assert np.array_equal(x0[:, :12], batch[0][:, :12]) and np.array_equal(now, batch[3][:, 0])
    out = wire.pack_mpc_solution_batch(batch[1], np.zeros((B, N, 12)), batch[3], batch[2], np.zeros((B, 2)), t)
    assert len(out) == B and all(len(m["states_horizon"]) == N for m in out)


def test_stamp_carries_a_fraction_that_rounds_up_to_a_full_second():
    """sim_time accumulated by += dt lands a hair below whole seconds (2.9999999999999973 after 60 ticks of 0.05): the
    nanosecond field then rounds to 1e9 and must carry into the seconds (ros::Time::fromSec), not wrap to zero."""
    assert wire._stamp(2.9999999999999973) == {"secs": 3, "nsecs": 0}
    t, worst = 0.0, 0.0
    for k in range(4000):
        st = wire._stamp(t)
        assert 0 <= st["nsecs"] < 1000000000
        worst = max(worst, abs(st["secs"] + 1e-9 * st["nsecs"] - t))
        # the time that round-trips through the message gives the node the same gait phase (mpc_to_wbid_node.cpp:105)
        assert int(np.floor((st["secs"] + 1e-9 * st["nsecs"] + 1e-9) / 0.05)) == int(np.floor((t + 1e-9) / 0.05)), (k, t, st)
        t += 0.05
    assert worst < 1e-9
