"""Wire-format shim: the `g1_msgs/SRBD_state` message the two processes of the demo exchange,
packed from / unpacked into the arrays of the batched path, so that a ROS-side adapter has nothing left to translate.

Follows, field by field:
  * `g1_msgs/msg/State.msg`, `ContactPoint.msg`, `SRBD_state.msg` (field order; `State.gravity` is **float32**,
    everything else float64 `geometry_msgs/Vector3` / `Point`, `ContactPoint.active` is a bool byte);
  * the simulator's publisher of `/srbd_current` (`g1_mujoco_sim/src/ros_run_simulation.py:22-78`): one state, four
    contact points in the order left lower, left upper, right lower, right upper;
  * the node's reader (`g1_mpc/src/mpc_to_wbid_node.cpp:30-64`): x0 from `states_horizon[0]`, the contact points tiled
    over the horizon;
  * the node's publisher of `/mpc_solution` (`mpc_to_wbid_node.cpp:322-382`): N states from x_opt, the first-step
    forces / positions / contact flags of the four points, landing position with z = 0;
  * the simulator's reader (`ros_run_simulation.py:188-218`): x_opt, u_opt0, contact states, landing position.

Messages are plain dicts with the .msg field names (what `rospy` message objects expose as attributes).
`serialize` / `deserialize` implement the ROS 1 wire encoding of `SRBD_state` (little-endian; `string` and arrays with a
uint32 length prefix; `Header` = uint32 seq, uint32 secs, uint32 nsecs, string frame_id; `bool` one byte).
The simulator's halves (pack_srbd_current, unpack_mpc_solution) are pinned to ros_run_simulation.py's own functions run on stand-in
message objects (tests/test_reference_callers_cpu.py); there is no ROS on this box, so the node's halves (C++) and the byte encoding
are restated from the sources / the ROS 1 serialization rules: that part stays parity unpinned.

Host-side glue only: nothing here touches the device; the arrays go to `BatchedMPC.walk_inputs` / `update_*`.
"""
import struct

import numpy as np

CONTACT_NAMES = ("left_foot_line_contact_lower", "left_foot_line_contact_upper",
                 "right_foot_line_contact_lower", "right_foot_line_contact_upper")      # mpc_to_wbid_node.cpp:353-357
FRAME_ID = "SRBD"                                                                       # mpc_to_wbid_node.cpp:326


def _vec(v):
    return {"x": float(v[0]), "y": float(v[1]), "z": float(v[2])}


def _f32(x):
    """What a float64 value becomes on the wire in a float32 field (State.msg:20)."""
    return float(np.float32(x))


def _state(i, row):
    return {"trajectory_index": int(i), "orientation": _vec(row[0:3]), "position": _vec(row[3:6]),
            "angular_velocity": _vec(row[6:9]), "linear_velocity": _vec(row[9:12]), "gravity": _f32(row[12])}


def _state_row(s):
    o, p, w, v = s["orientation"], s["position"], s["angular_velocity"], s["linear_velocity"]
    return [o["x"], o["y"], o["z"], p["x"], p["y"], p["z"], w["x"], w["y"], w["z"], v["x"], v["y"], v["z"], s["gravity"]]


def _stamp(t):
    """ros::Time(double) / fromSec: seconds floored, nanoseconds rounded, and a fraction that rounds up to a full second
    carried into `secs` (ros/time.h normalizeSecNSec) -- not wrapped to zero, which would put the stamp one second early."""
    secs = int(np.floor(t))
    nsecs = int(round((t - secs) * 1e9))
    if nsecs >= 1000000000:
        secs, nsecs = secs + 1, nsecs - 1000000000
    return {"secs": secs, "nsecs": nsecs}


# ---- simulator -> node: /srbd_current ------------------------------------------------------------------------

def pack_srbd_current(x0, foot_positions, qp_forces=None, sim_time=0.0, seq=0):
    """ros_run_simulation.py:22-78: current state (one State) + the four contact points with the forces applied now."""
    x0 = np.asarray(x0, dtype=np.float64).reshape(13)
    fp = np.asarray(foot_positions, dtype=np.float64).reshape(4, 3)
    qf = np.zeros(12) if qp_forces is None else np.asarray(qp_forces, dtype=np.float64).reshape(-1)
    contacts = [{"name": n, "position": _vec(fp[i]), "force": _vec(qf[3 * i:3 * i + 3]), "active": False}
                for i, n in enumerate(CONTACT_NAMES)]
    return {"header": {"seq": int(seq), "stamp": _stamp(sim_time), "frame_id": FRAME_ID},
            "states_horizon": [_state(0, x0)], "contacts": contacts, "landing_position": _vec((0.0, 0.0, 0.0))}


def unpack_srbd_current(msg, horizon):
    """mpc_to_wbid_node.cpp:30-64: x0[13] from states_horizon[0] (gravity arrives as float32), the current contact
    points [3 * n_contacts], and their tiling over the horizon c_horizon[N, 3 * n_contacts]."""
    x0 = np.array(_state_row(msg["states_horizon"][0]), dtype=np.float64)
    now = np.array([[c["position"]["x"], c["position"]["y"], c["position"]["z"]] for c in msg["contacts"]], dtype=np.float64).reshape(-1)
    return x0, now, np.tile(now, (int(horizon), 1))


def unpack_srbd_current_batch(msgs, horizon):
    """Many walkers at once: x0 [B,13], contacts_now [B,12], sim_time [B] -- the inputs of BatchedMPC.walk_inputs."""
    x0 = np.empty((len(msgs), 13)); now = np.empty((len(msgs), 12)); t = np.empty(len(msgs))
    for b, m in enumerate(msgs):
        x0[b], now[b], _ = unpack_srbd_current(m, horizon)
        t[b] = m["header"]["stamp"]["secs"] + 1e-9 * m["header"]["stamp"]["nsecs"]
    return x0, now, t


# ---- node -> simulator: /mpc_solution ------------------------------------------------------------------------

def pack_mpc_solution(x_opt, u_opt, c_horizon, contact_horizon, landing_xy, sim_time=0.0, seq=0):
    """mpc_to_wbid_node.cpp:322-382: the N predicted states, first-step force / position / contact flag of the four
    points, landing position of the swing foot (z is always 0 on the wire)."""
    x_opt = np.asarray(x_opt, dtype=np.float64)
    u0 = np.asarray(u_opt, dtype=np.float64)[0]
    c0 = np.asarray(c_horizon, dtype=np.float64)[0]
    act = np.asarray(contact_horizon)[0]
    contacts = [{"name": n, "position": _vec(c0[3 * i:3 * i + 3]), "force": _vec(u0[3 * i:3 * i + 3]), "active": bool(act[i])}
                for i, n in enumerate(CONTACT_NAMES)]
    return {"header": {"seq": int(seq), "stamp": _stamp(sim_time), "frame_id": FRAME_ID},
            "states_horizon": [_state(i, x_opt[i]) for i in range(x_opt.shape[0])], "contacts": contacts,
            "landing_position": _vec((landing_xy[0], landing_xy[1], 0.0))}


def unpack_mpc_solution(msg):
    """ros_run_simulation.py:188-218: x_opt [len(states_horizon),13], u_opt0 [12], contact_states [4], landing [3]."""
    x_opt = np.array([_state_row(s) for s in msg["states_horizon"]], dtype=np.float64).reshape(-1, 13)
    u0 = np.array([[c["force"]["x"], c["force"]["y"], c["force"]["z"]] for c in msg["contacts"]], dtype=np.float64).reshape(-1)
    act = np.array([bool(c["active"]) for c in msg["contacts"]])
    lp = msg["landing_position"]
    return x_opt, u0, act, np.array([lp["x"], lp["y"], lp["z"]])


def pack_mpc_solution_batch(x_opt, u_opt, c_horizon, contact, landing, sim_time):
    """One /mpc_solution message per walker from the batched outputs (numpy, host)."""
    return [pack_mpc_solution(x_opt[b], u_opt[b], c_horizon[b], contact[b], landing[b], float(sim_time[b])) for b in range(len(x_opt))]


# ---- ROS 1 serialization of g1_msgs/SRBD_state ----------------------------------------------------------------

def _s_string(s):
    b = s.encode("utf-8")
    return struct.pack("<I", len(b)) + b


def _s_vec(v):
    return struct.pack("<3d", v["x"], v["y"], v["z"])


def serialize(msg):
    """bytes of one SRBD_state in the ROS 1 wire encoding (field order of the .msg files)."""
    h = msg["header"]
    out = [struct.pack("<3I", h["seq"], h["stamp"]["secs"], h["stamp"]["nsecs"]), _s_string(h["frame_id"])]
    out.append(struct.pack("<I", len(msg["states_horizon"])))
    for s in msg["states_horizon"]:                 # State.msg: int32, 4 x Vector3, float32
        out.append(struct.pack("<i", s["trajectory_index"]))
        out += [_s_vec(s["orientation"]), _s_vec(s["position"]), _s_vec(s["angular_velocity"]), _s_vec(s["linear_velocity"])]
        out.append(struct.pack("<f", s["gravity"]))
    out.append(struct.pack("<I", len(msg["contacts"])))
    for c in msg["contacts"]:                       # ContactPoint.msg: string, 2 x Vector3, bool
        out += [_s_string(c["name"]), _s_vec(c["position"]), _s_vec(c["force"]), struct.pack("<B", 1 if c["active"] else 0)]
    out.append(_s_vec(msg["landing_position"]))
    return b"".join(out)


def deserialize(buf):
    """Inverse of serialize; raises ValueError on a truncated or over-long buffer."""
    pos = 0

    def take(fmt):
        nonlocal pos
        n = struct.calcsize(fmt)
        if pos + n > len(buf):
            raise ValueError("SRBD_state: buffer too short")
        v = struct.unpack_from(fmt, buf, pos)
        pos += n
        return v

    def string():
        nonlocal pos
        (n,) = take("<I")
        if pos + n > len(buf):
            raise ValueError("SRBD_state: buffer too short")
        s = bytes(buf[pos:pos + n]).decode("utf-8")
        pos += n
        return s

    def vec():
        x, y, z = take("<3d")
        return {"x": x, "y": y, "z": z}

    seq, secs, nsecs = take("<3I")
    msg = {"header": {"seq": seq, "stamp": {"secs": secs, "nsecs": nsecs}, "frame_id": string()}, "states_horizon": [], "contacts": []}
    for _ in range(take("<I")[0]):
        (idx,) = take("<i")
        s = {"trajectory_index": idx, "orientation": vec(), "position": vec(), "angular_velocity": vec(), "linear_velocity": vec()}
        s["gravity"] = float(take("<f")[0])
        msg["states_horizon"].append(s)
    for _ in range(take("<I")[0]):
        c = {"name": string(), "position": vec(), "force": vec()}
        c["active"] = bool(take("<B")[0])
        msg["contacts"].append(c)
    msg["landing_position"] = vec()
    if pos != len(buf):
        raise ValueError("SRBD_state: %d trailing bytes" % (len(buf) - pos))
    return msg
