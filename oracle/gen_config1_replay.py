"""BASELINE.json configs[0] as SURVEY.md 8d specifies it -- "single G1 SRBD MPC solve on states replayed from the
straight-line walking demo" -- regenerated deterministically (no recorded states ship with the reference):

    cpp_node preset, T = 200 ticks, tick t: sim_time = t dt (gait phase t, table A for t < 10 else B,
    mpc_to_wbid_node.cpp:84-127,105-106,175-187); x_ref, contact window, contact points and the Raibert overwrite by the
    node's own logic (oracle/node_oracle.py: mpc_to_wbid_node.cpp:385-470,507-527); the loop is closed on the SRBD model
    itself with the ORACLE's solution (reference build oracle.build, exact QP minimiser oracle.qp_exact, reference rollout
    oracle.rollout): x0(t+1) = x_opt[1] + N(0, sigma), sigma = (0.005 rad, 0.002 m, 0.02 rad/s, 0.01 m/s), seed 20251104;
    contact points <- row 1 of the planner's c_horizon; the previous x_opt stays for the next Raibert evaluation.
    x0(0) = [0,0,0, 5.26790425e-02, 7.44339342e-05, 5.97983255e-01, 0...0, float32(g)].

    python -m oracle.gen_config1_replay   ->  tests/golden/config1_replay.npz

The file holds the 200 problems (x0, x_ref, contact, c_horizon, p_com), the exact minimiser u_star of each, and the
first-step forces; tests/ and bench.py's latency leg read it (TEST/MEASUREMENT INPUT; never imported by the product)."""
import os

import numpy as np

from g1_locomotion_b200 import presets
from oracle import oracle, node_oracle

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "config1_replay.npz")
SIGMA = (0.005, 0.002, 0.02, 0.01)
SEED = 20251104
TICKS = 200


def main():
    preset = presets.get_preset("cpp_node")
    N, dt, g = preset["horizon"], preset["dt"], preset.get("gravity", -9.80665)
    p = oracle.params_from_dict(preset)
    rng = np.random.default_rng(SEED)
    x0 = np.zeros(13)
    x0[3:6] = (5.26790425e-02, 7.44339342e-05, 5.97983255e-01)
    x0[12] = np.float32(g)
    cur = np.array([x0[3] - 0.0293, 0.118508, 0.0, x0[3] + 0.1406, 0.118508, 0.0,
                    x0[3] - 0.0293, -0.118508, 0.0, x0[3] + 0.1406, -0.118508, 0.0])
    x_ref, x_opt = np.zeros((N, 13)), np.zeros((N, 13))
    keep = {k: [] for k in ("x0", "x_ref", "contact", "c_horizon", "p_com", "u_star", "x_opt")}
    for t in range(TICKS):
        x_ref, contact, c_hor, p_com, _ = node_oracle.node_update_inputs(x0, cur, t * dt + 0.5 * dt * 1e-6, x_opt, x_ref, dt, N, g)
        m = oracle.build(p, x0, x_ref, contact, c_hor, p_com)
        xs, _, cert = oracle.qp_exact(p, m["H"], m["g"], contact)
        assert cert["rc"] == 0, (t, cert)
        u = xs.reshape(N, 12)
        x_opt = oracle.rollout(p, m["A"], m["Bd"], u, x0, False, x_opt=x_opt.copy())
        for k, v in (("x0", x0), ("x_ref", x_ref), ("contact", contact), ("c_horizon", c_hor), ("p_com", p_com), ("u_star", u), ("x_opt", x_opt)):
            keep[k].append(np.array(v, float).copy())
        nxt = x_opt[1].copy()
        nxt[0:3] += rng.normal(0, SIGMA[0], 3); nxt[3:6] += rng.normal(0, SIGMA[1], 3)
        nxt[6:9] += rng.normal(0, SIGMA[2], 3); nxt[9:12] += rng.normal(0, SIGMA[3], 3)
        nxt[12] = np.float32(g)
        x0, cur = nxt, c_hor[1].copy()
    out = {k: np.stack(v) for k, v in keep.items()}
    np.savez_compressed(OUT, seed=SEED, sigma=np.array(SIGMA), **out)
    x = out["x0"]
    print("ticks", TICKS, "CoM x %.3f -> %.3f m, z in [%.3f, %.3f], |roll,pitch|max %.3f rad" %
          (x[0, 3], x[-1, 3], x[:, 5].min(), x[:, 5].max(), np.abs(x[:, 0:2]).max()))


if __name__ == "__main__":
    main()
