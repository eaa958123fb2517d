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
