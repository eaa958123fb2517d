"""CPU checks of oracle/node_oracle.py (restatement of the demo node's input generation and of WBID.setReference):
consistency with the independent synthetic generator in g1_locomotion_b200/inputs.py and internal properties."""
import numpy as np

from g1_locomotion_b200 import presets, inputs
from oracle import node_oracle as no


def test_node_inputs_agree_with_the_walking_generator():
    """inputs.walking_problem (noise-free) restates the same node lines from the problem-generation side: feeding its
    own x0 / current contact points / reference as previous x_opt into the node restatement reproduces it bit for bit."""
    preset = presets.get_preset("cpp_node")
    N, dt = preset["horizon"], preset["dt"]
    rng = np.random.default_rng(0)
    for phase in range(0, 45):
        x0, x_ref, contact, c_hor, p_com = inputs.walking_problem(phase, rng, preset, noise=False)
        got = no.node_update_inputs(x0, c_hor[0], (phase + 0.5) * dt, x_ref, np.zeros((N, 13)), dt, N, inputs.GRAVITY)
        assert np.array_equal(got[0], x_ref), phase
        assert np.array_equal(got[1], contact), phase
        assert np.array_equal(got[2], c_hor), phase
        assert np.array_equal(got[3], p_com), phase


def test_gait_window_and_swing_rows():
    preset = presets.get_preset("cpp_node")
    N, dt = 10, preset["dt"]
    x_prev = np.zeros((N, 13)); x_prev[:, 3] = np.arange(N); x_prev[:, 9] = 0.35
    cur = np.arange(12.0)
    for phase in (3, 12, 27):
        x_ref, contact, c_hor, p_com, land = no.node_update_inputs(np.zeros(13), cur, (phase + 0.1) * dt, x_prev, np.zeros((N, 13)), dt, N, -9.8)
        table = no.TABLE_B if phase >= 10 else no.TABLE_A
        assert np.array_equal(contact, table[phase % N:phase % N + N])
        assert np.array_equal(c_hor[0], cur)                         # row 0 is never overwritten
        assert np.all(c_hor[:, 2::3] == cur[2::3])                   # z is never overwritten
        if contact[0].sum() == 4:
            assert np.all(c_hor == cur) and np.all(land == 0)
        else:
            sub = (phase % N) % 5
            assert np.all(c_hor[:5 - sub] == cur)


def test_wbid_references():
    rng = np.random.default_rng(1)
    x1 = rng.normal(0, 0.3, 13); u0 = rng.normal(0, 50, 12); feet = rng.normal(0, 0.2, (4, 3))
    d = no.wbid_set_reference(x1, u0, feet, 35.0)
    R = d["R"]
    assert np.abs(R @ R.T - np.eye(3)).max() < 1e-15 and abs(np.linalg.det(R) - 1) < 1e-15
    # static-xyz Euler: R = Rz(yaw) Ry(pitch) Rx(roll)
    cr, sr, cp, sp, cy, sy = np.cos(x1[0]), np.sin(x1[0]), np.cos(x1[1]), np.sin(x1[1]), np.cos(x1[2]), np.sin(x1[2])
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]]); Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    assert np.abs(R - Rz @ Ry @ Rx).max() < 1e-15
    assert np.array_equal(d["wrench"], u0) and np.all(d["acceleration"][:3] == 0)
    assert np.allclose(d["com_acc"], u0.reshape(3, 4).sum(1) / 35.0 + [0, 0, -9.80665], rtol=0, atol=1e-15)
