"""Synthetic input generators for the SRBD-MPC hot path.

These restate the *caller* logic that feeds MPC::updateMPC in the reference demo -- reference trajectory
(mpc_to_wbid_node.cpp:403-435), table-driven gait window (:82-127,175-187) and the Raibert footstep
overwrite (:227-309,507-527) -- only as far as needed to produce walking-shaped problems; no recorded
states ship with the reference.  Host-side numpy/torch; nothing here is on the timed path.
"""
import math
import numpy as np
import torch

from .presets import GRAVITY

# mpc_to_wbid_node.cpp:84-127 -- rows = planning steps, cols = [L-lower(heel), L-upper(toe), R-lower, R-upper]
GAIT_TABLE_A = np.array([[1, 1, 1, 1]] * 5 + [[0, 0, 1, 1]] * 5 + [[1, 1, 0, 0]] * 5 + [[0, 0, 1, 1]] * 5, float)
GAIT_TABLE_B = np.array([[1, 1, 0, 0]] * 5 + [[0, 0, 1, 1]] * 5 + [[1, 1, 0, 0]] * 5 + [[0, 0, 1, 1]] * 5, float)
COM_REF = (5.26790425e-02, 7.44339342e-05, 5.97983255e-01)   # mpc_to_wbid_node.cpp:412-413,439
V_CMD_X = 0.35                                               # :405
FOOT_OFFSET_Y, FOOT_LOWER_X, FOOT_UPPER_X = 0.118508, 0.0293, 0.1406   # :217-219
STANCE_DURATION, K_VEL_GAIN = 0.25, 0.02                     # :27, :521


def _raibert(x_opt_row, x_ref_row, dt):
    """mpc_to_wbid_node.cpp:507-527 (x, y only; the `* dt` factor is as coded)."""
    p, v, vd = x_opt_row[3:5], x_opt_row[9:11], x_ref_row[9:11]
    return p + v * STANCE_DURATION * 0.5 * dt + K_VEL_GAIN * (v - vd)


def walking_problem(gait_phase, rng, preset, noise=True):
    """One tick of the straight-line walking demo at gait phase `gait_phase` (= floor(t/dt)).

    The plant state is the nominal walk (CoM moving at v_cmd) plus seeded noise; the previous solve's
    x_opt that the Raibert rule reads is replaced by the reference trajectory plus the same noise.
    Returns x0[13], x_ref[N,13], contact[N,4], c_horizon[N,12], p_com[N,3].  Requires N <= 10 (the
    reference table only has 20 rows, :82)."""
    N, dt = preset["horizon"], preset["dt"]
    assert N <= 10 and preset["num_contacts"] == 4
    table = GAIT_TABLE_A if gait_phase < 10 else GAIT_TABLE_B
    k = gait_phase % N                                           # :105
    contact = np.array([table[i + k] for i in range(N)], float)  # :175-187
    t = gait_phase * dt
    sig = (0.005, 0.002, 0.02, 0.01) if noise else (0, 0, 0, 0)
    x0 = np.zeros(13)
    x0[3:6] = (COM_REF[0] + V_CMD_X * t, COM_REF[1], COM_REF[2])
    x0[9] = V_CMD_X
    x0[0:3] += rng.normal(0, sig[0], 3) if noise else 0
    x0[3:6] += rng.normal(0, sig[1], 3) if noise else 0
    x0[6:9] += rng.normal(0, sig[2], 3) if noise else 0
    x0[9:12] += rng.normal(0, sig[3], 3) if noise else 0
    x0[12] = np.float32(GRAVITY)                                 # State.msg:20 (float32 over ROS)
    x_ref = np.zeros((N, 13))                                    # :408-434
    for i in range(N):
        x_ref[i, 3] = x0[3] + V_CMD_X * i * dt
        x_ref[i, 4], x_ref[i, 5] = COM_REF[1], COM_REF[2]
        x_ref[i, 9] = V_CMD_X
        x_ref[i, 12] = GRAVITY
    p_com = x_ref[:, 3:6].copy()                                 # :461-468
    # current foot centres (line contacts heel/toe): the foot that is (or was last) in stance sits where
    # it landed, i.e. about where the CoM was at touch-down.
    sub = k % 5
    left_stance_now = contact[0, 0] == 1
    both = contact[0].sum() == 4
    land_bias = V_CMD_X * STANCE_DURATION * 0.5 * dt
    x_st = x0[3] - V_CMD_X * dt * sub + land_bias
    x_sw = x0[3] - V_CMD_X * dt * (5 + sub) + land_bias + (V_CMD_X * dt * 10) * (sub / 5.0)
    xl, xr = (x_st, x_sw) if left_stance_now else (x_sw, x_st)
    if both:
        xl = xr = x0[3]
    jit = rng.normal(0, 0.003, 4) if noise else np.zeros(4)
    cur = np.array([xl - FOOT_LOWER_X + jit[0], COM_REF[1] + FOOT_OFFSET_Y + jit[1], 0.0,
                    xl + FOOT_UPPER_X + jit[0], COM_REF[1] + FOOT_OFFSET_Y + jit[1], 0.0,
                    xr - FOOT_LOWER_X + jit[2], COM_REF[1] - FOOT_OFFSET_Y + jit[3], 0.0,
                    xr + FOOT_UPPER_X + jit[2], COM_REF[1] - FOOT_OFFSET_Y + jit[3], 0.0])
    c_hor = np.tile(cur, (N, 1))                                 # :50-60
    if contact[0].sum() == 2:                                    # single support (:221-227)
        x_prev = x_ref.copy()
        if noise:
            x_prev[:, 3:5] += rng.normal(0, sig[1], (N, 2))
            x_prev[:, 9:11] += rng.normal(0, sig[3], (N, 2))

        def overwrite(rows, istar):
            if istar >= N:
                return                                           # Appendix C: call skipped, range empty
            land = _raibert(x_prev[istar], x_ref[istar], dt)
            for i in rows:
                if contact[i, 0] == 0:                           # left foot in swing (:244-258)
                    c_hor[i, 0], c_hor[i, 1] = land[0] - FOOT_LOWER_X, land[1] + FOOT_OFFSET_Y
                    c_hor[i, 3], c_hor[i, 4] = land[0] + FOOT_UPPER_X, land[1] + FOOT_OFFSET_Y
                else:                                            # right foot (:259-271)
                    c_hor[i, 6], c_hor[i, 7] = land[0] - FOOT_LOWER_X, land[1] - FOOT_OFFSET_Y
                    c_hor[i, 9], c_hor[i, 10] = land[0] + FOOT_UPPER_X, land[1] - FOOT_OFFSET_Y
        overwrite(range(5 - sub, 10 - sub), 5 - sub)
        overwrite(range(10 - sub, 10), 10 - sub)
    return x0, x_ref, contact, c_hor, p_com


def _stack(problems):
    return tuple(np.stack([p[i] for p in problems]) for i in range(5))


def walking_replay(preset, n_ticks=200, seed=20251104):
    """Config 1: `n_ticks` consecutive ticks of the walking demo (gait phase = tick)."""
    rng = np.random.default_rng(seed)
    return _stack([walking_problem(t, rng, preset) for t in range(n_ticks)])


def walking_batch(preset, B=1024, seed0=1000):
    """Config 2: B independent walking problems, seed 1000+b, gait-phase offset k = b % 10, t >= 10."""
    out = []
    for b in range(B):
        rng = np.random.default_rng(seed0 + b)
        phase = 10 + 10 * int(rng.integers(0, 8)) + (b % 10)
        out.append(walking_problem(phase, rng, preset))
    return _stack(out)


def random_batch(preset, B, seed=7, device="cpu", dtype=torch.float64):
    """Configs 3/4/5: random synthetic states, yaw-ramp references and periodic two-foot gait schedules.

    Distributions follow SURVEY.md 8d (config 3).  `num_contacts` 4 = heel/toe line contact per foot,
    8 = four corners per foot (x in {-0.05,+0.12}, y in {-0.03,+0.03} about each ankle).  Generated
    with a seeded torch.Generator on `device` (so config 5 shards are produced on the GPU)."""
    N, nc, dt = preset["horizon"], preset["num_contacts"], preset["dt"]
    assert nc in (4, 8)
    gen = torch.Generator(device=device)
    gen.manual_seed(int(seed))

    def U(lo, hi, *shape):
        return lo + (hi - lo) * torch.rand(*shape, generator=gen, device=device, dtype=dtype)

    yaw = U(-math.pi, math.pi, B)
    x0 = torch.zeros(B, 13, device=device, dtype=dtype)
    x0[:, 0:2] = U(-0.15, 0.15, B, 2)
    x0[:, 2] = yaw
    centre = U(-1.0, 1.0, B, 2)                       # support centre in the world
    x0[:, 3:5] = centre + U(-0.05, 0.05, B, 2)
    x0[:, 5] = U(0.5, 0.7, B)
    x0[:, 6:9] = U(-0.5, 0.5, B, 3)
    x0[:, 9:12] = U(-0.5, 0.5, B, 3)
    x0[:, 12] = float(np.float32(GRAVITY))
    steps = torch.arange(N, device=device, dtype=dtype)
    yaw_rate = U(-0.5, 0.5, B)
    v_cmd = U(-0.5, 0.5, B, 2)
    x_ref = torch.zeros(B, N, 13, device=device, dtype=dtype)
    x_ref[:, :, 2] = yaw[:, None] + yaw_rate[:, None] * steps[None] * dt
    x_ref[:, :, 3:5] = x0[:, None, 3:5] + v_cmd[:, None, :] * steps[None, :, None] * dt
    x_ref[:, :, 5] = COM_REF[2]
    x_ref[:, :, 8] = yaw_rate[:, None]
    x_ref[:, :, 9:11] = v_cmd[:, None, :]
    x_ref[:, :, 12] = GRAVITY
    p_com = x_ref[:, :, 3:6].clone()
    # gait: period P in {6,8,10,12}; left stance for phase in [0, P/2 + ds), right stance for
    # phase in [P/2, P + ds) (mod P), ds = double-support steps
    P = torch.tensor([6, 8, 10, 12], device=device)[torch.randint(0, 4, (B,), generator=gen, device=device)]
    duty = U(0.0, 0.3, B)
    ds = torch.floor(duty * P.to(dtype) / 2)
    phase0 = torch.floor(U(0, 1, B) * P.to(dtype))
    ph = torch.remainder(phase0[:, None] + steps[None], P[:, None].to(dtype))
    half = (P // 2).to(dtype)[:, None]
    left = (ph < half + ds[:, None])
    right = (ph >= half) | (ph < ds[:, None])
    neither = ~(left | right)
    left = left | neither
    per_foot = nc // 2
    contact = torch.cat([left[:, :, None].expand(B, N, per_foot), right[:, :, None].expand(B, N, per_foot)], 2).to(dtype)
    # foot points: nominal offsets about the support centre, rotated by yaw, + N(0, 0.02)
    if nc == 4:
        offs = [(-FOOT_LOWER_X, FOOT_OFFSET_Y), (FOOT_UPPER_X, FOOT_OFFSET_Y),
                (-FOOT_LOWER_X, -FOOT_OFFSET_Y), (FOOT_UPPER_X, -FOOT_OFFSET_Y)]
    else:
        offs = [(ox, sy * FOOT_OFFSET_Y + oy) for sy in (1, -1) for ox in (-0.05, 0.12) for oy in (-0.03, 0.03)]
    offs = torch.tensor(offs, device=device, dtype=dtype)                      # [nc,2]
    c, s = torch.cos(yaw), torch.sin(yaw)
    rx = c[:, None] * offs[None, :, 0] - s[:, None] * offs[None, :, 1]
    ry = s[:, None] * offs[None, :, 0] + c[:, None] * offs[None, :, 1]
    foot = torch.zeros(B, nc, 3, device=device, dtype=dtype)
    foot[:, :, 0] = centre[:, None, 0] + rx
    foot[:, :, 1] = centre[:, None, 1] + ry
    foot[:, :, 0:2] += 0.02 * torch.randn(B, nc, 2, generator=gen, device=device, dtype=dtype)
    c_hor = foot.reshape(B, 1, 3 * nc).expand(B, N, 3 * nc).contiguous()
    return x0, x_ref, contact, c_hor, p_com
