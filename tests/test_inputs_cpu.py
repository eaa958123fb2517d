"""Synthetic input generators."""
import numpy as np

from g1_locomotion_b200 import inputs, presets


def test_walking_batch_follows_gait_tables():
    p = presets.get_preset("cpp_node")
    x0, xr, ct, ch, pc = inputs.walking_batch(p, 32)
    assert x0.shape == (32, 13) and xr.shape == (32, 10, 13) and ct.shape == (32, 10, 4)
    assert ch.shape == (32, 10, 12) and pc.shape == (32, 10, 3)
    assert set(np.unique(ct)) <= {0.0, 1.0}
    assert np.all(ct.sum(2) >= 2)                               # table B: always one foot (2 points) down
    assert np.all(ct[:, :, 0] == ct[:, :, 1]) and np.all(ct[:, :, 2] == ct[:, :, 3])
    assert np.all(x0[:, 12] == np.float32(-9.80665))            # float32 gravity over ROS (State.msg:20)
    assert np.all(xr[:, :, 12] == -9.80665) and np.all(xr[:, :, 9] == 0.35)
    assert np.allclose(pc, xr[:, :, 3:6])
    # deterministic
    again = inputs.walking_batch(p, 32)
    assert all(np.array_equal(a, b) for a, b in zip((x0, xr, ct, ch, pc), again))


def test_walking_replay_phase_zero_is_double_support():
    p = presets.get_preset("cpp_node")
    x0, xr, ct, ch, pc = inputs.walking_replay(p, 30)
    assert np.all(ct[0, :5] == 1) and np.all(ct[0, 5:, :2] == 0)     # table A rows 0..9
    assert np.all(ct[12, 0] == [1, 1, 0, 0])                          # table B, k = 2


def test_random_batch_shapes_and_stance():
    for nc in (4, 8):
        p = presets.get_preset("cpp_node", num_contacts=nc, horizon=12)
        x0, xr, ct, ch, pc = [t.numpy() for t in inputs.random_batch(p, 64, seed=3)]
        assert ct.shape == (64, 12, nc) and ch.shape == (64, 12, 3 * nc)
        assert np.all(ct.sum(2) >= nc // 2)                       # >= 1 foot in stance each step
        assert np.abs(x0[:, 2]).max() <= np.pi and np.all(np.abs(x0[:, 0:2]) <= 0.15)
