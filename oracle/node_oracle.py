"""CPU restatement (numpy, scalar loops) of the two callers either side of the hot path -- TEST INFRASTRUCTURE,
never imported by the product (g1_locomotion_b200/).

  node_update_inputs : MPCNode::update up to the updateMPC call, g1_mpc/src/mpc_to_wbid_node.cpp:385-470
                       (callbackSrbdCurrent :30-64, gaitPlanner :71-320, RaibertHeuristic :507-527)
  wbid_set_reference : WBID.setReference, g1_mujoco_sim/src/wbid.py:231-297 (call site ros_run_simulation.py:409)

parity unpinned: neither file can run here (ROS/catkin C++ node; OpenSoT/xbot2 Python bindings), and the reference
ships no recorded inputs/outputs for them; both are restated from the source, line references beside each step.
"""
import math

import numpy as np

# mpc_to_wbid_node.cpp:84-127
TABLE_A = np.array([[1, 1, 1, 1]] * 5 + [[0, 0, 1, 1]] * 5 + [[1, 1, 0, 0]] * 5 + [[0, 0, 1, 1]] * 5)
TABLE_B = np.array([[1, 1, 0, 0]] * 5 + [[0, 0, 1, 1]] * 5 + [[1, 1, 0, 0]] * 5 + [[0, 0, 1, 1]] * 5)


def _raibert(x_opt, x_ref, i, stance_duration, dt):
    """mpc_to_wbid_node.cpp:507-527"""
    k_vel_gain = 0.02
    out = []
    for a in range(2):
        p, v, vd = x_opt[i, 3 + a], x_opt[i, 9 + a], x_ref[i, 9 + a]
        out.append(p + v * stance_duration * 0.5 * dt + k_vel_gain * (v - vd))
    return out


def node_update_inputs(x0, contacts_now, sim_time, x_opt_prev, x_ref_prev, dt, N, g):
    """One tick.  Returns x_ref[N,13], contact[N,4], c_horizon[N,12], p_com[N,3], landing[2]."""
    x0 = np.asarray(x0, float); x_opt = np.asarray(x_opt_prev, float)
    x_ref = np.array(x_ref_prev, float)                          # :398 starts from the previous reference
    for i in range(N):                                           # :408-434
        x_ref[i, 4] = 7.44339342e-05
        x_ref[i, 5] = 5.97983255e-01
        x_ref[i, 3] = x0[3] + 0.35 * i * dt
        x_ref[i, 9] = 0.35
        x_ref[i, 2] = 0
        x_ref[i, 8] = 0.0
        x_ref[i, 12] = g
    p_com = x_ref[:, 3:6].copy()                                 # :461-468
    c_hor = np.tile(np.asarray(contacts_now, float).reshape(1, 12), (N, 1))   # :50-60
    phase = int(math.floor(sim_time / dt))                       # :105
    table = TABLE_B if phase >= 10 else TABLE_A                  # :106
    k = phase % N                                                # :171
    contact = np.array([table[i + k] for i in range(N)], int)    # :175-187
    sub = k % 5                                                  # :213
    land_pub = [0.0, 0.0]
    LOW, UP, OFFY = 0.0293, 0.1406, 0.118508                     # :217-219
    if contact[0].sum() == 2:                                    # :222-227

        def put(i, lx, ly):
            if contact[i, 0] == 0:                               # :246-258
                c_hor[i, 0], c_hor[i, 1] = lx - LOW, ly + OFFY
                c_hor[i, 3], c_hor[i, 4] = lx + UP, ly + OFFY
                return [lx, ly - OFFY]
            c_hor[i, 6], c_hor[i, 7] = lx - LOW, ly - OFFY       # :259-271
            c_hor[i, 9], c_hor[i, 10] = lx + UP, ly - OFFY
            return [lx, ly + OFFY]
        i1, i2 = 5 - sub, 10 - sub
        if i1 < N:                                               # (the node reads row i1 unconditionally; i1 <= 5 < N = 10)
            lx, ly = _raibert(x_opt, x_ref, i1, 0.25, dt)        # :238-243
            for i in range(i1, min(i2, N)):                      # :245
                land_pub = put(i, lx, ly)
        if i2 < N:                                               # :276-281 (row 10 of a 10-row matrix when sub == 0: unused)
            lx, ly = _raibert(x_opt, x_ref, i2, 0.25, dt)
            for i in range(i2, min(10, N)):                      # :284
                put(i, lx, ly)
    return x_ref, contact.astype(float), c_hor, p_com, np.array(land_pub)


def map_gait_phases(phases, current_time, N, dt, foot_offset=0.08):
    """main.py:259-285 -- the lookup of the planner's phase list per horizon step and its mapping onto
    (contact_horizon [N,4], c_horizon [N,12]); `phases` is what GaitPlanner.plan_gait returns (gait_planner.py:27-62)."""
    c_horizon, contact_horizon = [], []
    for i in range(N):
        t_i = current_time + i * dt                              # :263
        phase = None
        for ph in phases:                                        # :265-268
            if ph["start_time"] <= t_i < ph["end_time"]:
                phase = ph
                break
        if phase is None:                                        # :270-271
            phase = phases[-1]
        left = np.hstack((np.array(phase["left_foot"]), 0))      # :273-274
        right = np.hstack((np.array(phase["right_foot"]), 0))
        off = np.array([foot_offset, 0, 0])
        c_horizon.append(np.concatenate((left - off, left + off, right - off, right + off)))     # :276-281
        lc = 1 if phase["support_leg"] in ["both", "left"] else 0                                # :282-283
        rc = 1 if phase["support_leg"] in ["both", "right"] else 0
        contact_horizon.append([lc, lc, rc, rc])
    return np.array(contact_horizon, float), np.array(c_horizon, float)


def plan_gait(double_support_time, single_support_time, step_length, step_width, num_steps):
    """GaitPlanner.__init__ + plan_gait restated (gait_planner.py:5-62): the phase list, with the same running sums."""
    left, right = np.array([0.0, -step_width]), np.array([0.0, step_width])                      # :19-20
    t = 0.0
    phases = [dict(start_time=t, end_time=t + double_support_time, left_foot=left.copy(), right_foot=right.copy(), support_leg="both")]
    t += double_support_time
    for i in range(int(num_steps)):
        leg = "left" if i % 2 == 0 else "right"
        phases.append(dict(start_time=t, end_time=t + single_support_time, left_foot=left.copy(), right_foot=right.copy(), support_leg=leg))
        t += single_support_time
        inc = step_length if i == 0 else 2 * step_length
        if leg == "left":
            right = right + np.array([inc, 0.0])
        else:
            left = left + np.array([inc, 0.0])
    return phases


def double_support_reference(x_ref, current_time, total_duration, dt):
    """testing_double_support, main.py:177-197 (in place on x_ref[N,13])."""
    for i in range(x_ref.shape[0]):
        t_ref = current_time + i * dt
        if t_ref < total_duration / 4:
            ref = (0.0, 0.1, 1.0)
        elif t_ref < total_duration / 2:
            ref = (0.0, 0.0, 1.0)
        elif t_ref < 3 * total_duration / 4:
            ref = (0.0, 0.0, 0.7)
        else:
            ref = (0.01, 0.0, 1.0)
        x_ref[i, 3:6] = ref
    return x_ref


def main_loop(orc, preset, planner, ticks, total_duration=4.0, foot_offset=0.08, solve=None, on_tick=None):
    """The SRBD-only loop of g1_mpc/scripts/main.py:200-326 for ONE walker on the CPU: planner phases -> mapping ->
    reference build (oracle.build) -> exact QP minimiser -> x_opt[0] carry and rollout without x_current (main.py:306-311);
    the next solve starts from x_opt[2] (main.py:228).  Returns traj_x[ticks,13] (x_opt[0] after each rollout), traj_u[ticks,12].
    solve(k, m, contact) -> u[N,12] replaces the exact minimiser (the pinning test feeds the reference run's own u_opt back, so that
    the LOOP LOGIC is compared tick by tick without a solver in between); on_tick(k, dict of this tick's inputs and x_opt) observes."""
    N, dt, g = preset["horizon"], preset["dt"], preset.get("gravity", -9.80665)
    p = orc.params_from_dict(preset)
    phases = plan_gait(*planner)
    x_ref = np.zeros((N, 13)); x_ref[:, 12] = g; x_ref[:, 5] = 1.0         # main.py:207-209
    x_opt = np.zeros((N, 13))
    t = 0.0
    tx, tu = np.zeros((ticks, 13)), np.zeros((ticks, 12))
    for k in range(ticks):
        if k == 0:
            x0 = x_ref[0].copy()                                          # main.py:226
        else:
            x0 = x_opt[2].copy()                                          # main.py:228
            double_support_reference(x_ref, t, total_duration, dt)        # main.py:230
        contact, c_hor = map_gait_phases(phases, t, N, dt, foot_offset)
        p_com = x_ref[:, 3:6].copy()                                      # main.py:286
        m = orc.build(p, x0, x_ref, contact, c_hor, p_com)
        if solve is None:
            xs, _, cert = orc.qp_exact(p, m["H"], m["g"], contact)
            assert cert["rc"] == 0, cert
            u = xs.reshape(N, 12)
        else:
            u = np.asarray(solve(k, m, contact), float).reshape(N, 12)
        x_opt[0] = x0 if k == 0 else x_opt[1].copy()                      # main.py:306-310
        x_opt = orc.rollout(p, m["A"], m["Bd"], u, None, False, x_opt=x_opt)
        tx[k], tu[k] = x_opt[0], u[0]
        if on_tick is not None:
            on_tick(k, dict(x0=x0.copy(), x_ref=x_ref.copy(), contact=contact.copy(), c_horizon=c_hor.copy(), p_com=p_com, x_opt=x_opt.copy()))
        t += dt                                                           # main.py:326
    return tx, tu


def _euler_matrix_sxyz(ai, aj, ak):
    """tf.transformations.euler_matrix(ai, aj, ak, 'sxyz')[0:3,0:3] (static x-y-z: i=0, j=1, k=2, no repetition, parity even)."""
    si, sj, sk = math.sin(ai), math.sin(aj), math.sin(ak)
    ci, cj, ck = math.cos(ai), math.cos(aj), math.cos(ak)
    cc, cs = ci * ck, ci * sk
    sc, ss = si * ck, si * sk
    return np.array([[cj * ck, sj * sc - cs, sj * cc + ss],
                     [cj * sk, sj * ss + cc, sj * cs - sc],
                     [-sj, cj * si, cj * ci]])


def wbid_set_reference(x_opt1, u_opt0, foot_positions_curr, mass):
    """wbid.py:244-297 -> dict of the references handed to the OpenSoT tasks."""
    x_opt1 = np.asarray(x_opt1, float); u_opt0 = np.asarray(u_opt0, float)
    R = _euler_matrix_sxyz(*x_opt1[0:3])                         # :249-250
    velocity = np.hstack((x_opt1[9:12], x_opt1[6:9]))            # :259-261
    inertia_torso = np.array([[8.20564e-2, 0., 0.], [0., 8.05015e-2, 0.], [0., 0., 0.32353e-2]])   # :264-268
    inertia_torso_inv = np.linalg.inv(inertia_torso)             # :271
    r = np.array(foot_positions_curr, float).reshape(4, 3) - np.tile(x_opt1[3:6], (4, 1))          # :273
    s = np.zeros((1, 3))
    for i in range(4):                                           # :275-277
        s = s + np.cross(r[i, :], x_opt1[6:9])
    acc = np.vstack((np.zeros((3, 1)), inertia_torso_inv @ s.T)).ravel()   # :280-284
    gravity = np.array([0, 0, -9.80665])                         # :288
    sum_forces = np.sum(np.reshape(u_opt0, (3, 4)), axis=1)      # :290 (as coded)
    com_acc = sum_forces / mass + gravity                        # :291
    return dict(R=R, velocity=velocity, acceleration=acc, com_pos=x_opt1[3:6].copy(), com_vel=x_opt1[9:12].copy(),
                com_acc=com_acc, wrench=u_opt0.copy())


def wbid_refs_flat(x_opt1, u_opt0, feet, mass):
    d = wbid_set_reference(x_opt1, u_opt0, feet, mass)
    return np.concatenate([d["R"].ravel(), d["velocity"], d["acceleration"], d["com_pos"], d["com_vel"], d["com_acc"], d["wrench"]])


def closed_loop(orc, preset, x0, contacts_now, sim_time, ticks):
    """SURVEY 8f-3 on the CPU for ONE walker: node inputs -> reference build (oracle.build) -> exact QP minimiser
    (oracle.qp_exact) -> reference rollout (mpc.cpp:427-446) -> x0 <- x_opt[1], contact points <- c_horizon[1], t += dt.
    Returns traj_x[ticks,13], traj_u[ticks,12]."""
    N, dt = preset["horizon"], preset["dt"]
    g = preset.get("gravity", -9.80665)
    p = orc.params_from_dict(preset)
    x0 = np.array(x0, float); cur = np.array(contacts_now, float); t = float(sim_time)
    x_ref = np.zeros((N, 13)); x_opt = np.zeros((N, 13))
    tx, tu = np.zeros((ticks, 13)), np.zeros((ticks, 12))
    for k in range(ticks):
        x_ref, contact, c_hor, p_com, _ = node_update_inputs(x0, cur, t, x_opt, x_ref, dt, N, g)
        m = orc.build(p, x0, x_ref, contact, c_hor, p_com)
        xs, _, cert = orc.qp_exact(p, m["H"], m["g"], contact)
        assert cert["rc"] == 0, cert
        u = xs.reshape(N, 12)
        x_opt = orc.rollout(p, m["A"], m["Bd"], u, x0, False, x_opt=x_opt.copy())
        tx[k], tu[k] = x0, u[0]
        x0, cur, t = x_opt[1].copy(), c_hor[1].copy(), t + dt
    return tx, tu
