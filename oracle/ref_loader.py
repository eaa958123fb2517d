"""Import the reference's own g1_mpc/scripts/mpc.py UNMODIFIED (read-only, from /root/reference).

Only usable in the build container (the GPU box has no /root/reference); used by oracle/gen_golden.py to
produce tests/golden/*.npz and by tests that are skipped when the reference tree is absent.
`import osqp` at mpc.py:1 is satisfied by oracle/osqp_shim.py (OSQP().setup/.solve over the C restatement of the
OSQP algorithm), so that solve_qp() (mpc.py:323-358) and with it the whole MPC.update() (mpc.py:360-377) run
unmodified; the build steps never touch it.  `import matplotlib.pyplot` at gait_planner.py:2 is satisfied by an
empty stand-in (only plot_gait() uses it).

Round 2 also runs two CALLERS of the path from the reference's own sources, to pin the restatements in oracle/node_oracle.py:
  * g1_mujoco_sim/src/wbid.py: WholeBodyID.setReference (wbid.py:231-297) on an object whose OpenSoT tasks are recording stand-ins
    (xbot2_interface / pyopensot / ttictoc are absent here and never reached by setReference; tf.transformations.euler_matrix, a
    third-party function outside /root/reference, is restated from its published source -- static x-y-z convention);
  * g1_mpc/scripts/main.py: main() (main.py:200-326) with matplotlib / srbd_plotting stubbed and the reference MPC class behind the
    OSQP shim; what every tick leaves in x_opt[0] / u_opt[0] is recorded by wrapping compute_rollout.
"""
import importlib.util
import os
import sys
import types
import numpy as np

REF_MPC_PY = "/root/reference/g1_mpc/scripts/mpc.py"


def available():
    return os.path.exists(REF_MPC_PY)


REF_GAIT_PY = "/root/reference/g1_mpc/scripts/gait_planner.py"


def load_module():
    if "osqp" not in sys.modules:
        from . import osqp_shim
        sys.modules["osqp"] = osqp_shim
    spec = importlib.util.spec_from_file_location("g1_ref_mpc", REF_MPC_PY)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load_gait_planner():
    """The reference's GaitPlanner class (gait_planner.py:4-62), imported unmodified."""
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        mpl.pyplot = types.ModuleType("matplotlib.pyplot")
        sys.modules["matplotlib"], sys.modules["matplotlib.pyplot"] = mpl, mpl.pyplot
    spec = importlib.util.spec_from_file_location("g1_ref_gait_planner", REF_GAIT_PY)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.GaitPlanner


def reference_update(o, x0, xref, contact, c_hor, p_com, x_current=None, one_rollout=False):
    """The reference's MPC.update() (mpc.py:360-377) end to end: build, solve_qp through the osqp shim, rollout."""
    import contextlib, io
    o.x0 = np.array(x0, float).copy()
    o.x_ref_hor = np.array(xref, float).copy()
    with contextlib.redirect_stdout(io.StringIO()):      # compute_rollout prints (mpc.py:389-391)
        u, x = o.update(np.array(contact, float), np.array(c_hor, float), np.array(p_com, float),
                        None if x_current is None else np.array(x_current, float), one_rollout)
    return np.array(u, float).copy(), np.array(x, float).copy()


def make_reference_mpc(preset):
    """Instantiate the reference class and apply a preset dict (see g1_locomotion_b200/presets.py).
    The `cpp_node` preset overrides the Python object's constants with the values of mpc.cpp:25-27,90-98
    and mpc.hpp:22; maths is otherwise identical line for line."""
    mod = load_module()
    o = mod.MPC(mu=preset["mu"], fz_min=preset["fz_min"], fz_max=preset["fz_max"], dt=preset["dt"],
                HORIZON_LENGTH=preset["horizon"])
    o.ROBOT_MASS = preset["mass"]
    o.INERTIA_BODY = np.asarray(preset["inertia_body"], float).reshape(3, 3)
    o.Q_WEIGHTS = np.diag(np.asarray(preset["q_weights"], float))
    o.NUM_CONTACTS = preset["num_contacts"]
    o.NUM_CONTROLS = 3 * preset["num_contacts"]
    o.R_WEIGHTS = np.diag(np.full(o.NUM_CONTROLS, preset["r_weight"]))
    o.init_matrices()
    return o


def reference_build(o, cost_scale, x0, xref, contact, c_hor, p_com):
    """Run the reference's build steps in update() order (mpc.py:360-374), no solve."""
    o.x0 = np.array(x0, float).copy()
    o.x_ref_hor = np.array(xref, float).copy()
    o.extract_psi(); o.rotation_matrix_T(); o.set_Q(); o.set_R()
    o.calculate_A_continuous(); o.calculate_A_discrete()
    o.calculate_B_continuous(np.array(c_hor, float), np.array(p_com, float)); o.calculate_B_discrete()
    o.calculate_Aqp(); o.calculate_Bqp(); o.calculate_Ac(); o.calculate_bounds(np.array(contact, float))
    o.calculate_hessian(); o.calculate_gradient()
    return dict(A=o.A_discrete.copy(), Bd=o.B_discrete_hor.copy(), Aqp=o.Aqp.copy(), Bqp=o.Bqp.copy(),
                H=cost_scale * o.Hessian, g=cost_scale * o.gradient.ravel(), Acon=o.Ac.toarray(),
                l=o.lower_bounds_horizon.copy(), u=o.upper_bounds_horizon.copy(), psi=float(o.psi))


def reference_rollout(o, u_opt, x_current, only_first_step):
    import contextlib, io
    o.u_opt = np.array(u_opt, float).reshape(o.HORIZON_LENGTH, o.NUM_CONTROLS)
    o.x_opt = np.zeros((o.HORIZON_LENGTH, o.NUM_STATES))
    with contextlib.redirect_stdout(io.StringIO()):  # mpc.py:389-391 prints
        o.compute_rollout(None if x_current is None else np.array(x_current, float), only_first_step)
    return o.x_opt.copy()


# ----------------------------------------------------------------------------------------------------------------------
# callers of the path, run from the reference's own sources
# ----------------------------------------------------------------------------------------------------------------------
REF_WBID_PY = "/root/reference/g1_mujoco_sim/src/wbid.py"
REF_MAIN_PY = "/root/reference/g1_mpc/scripts/main.py"


class _Anything:
    """Stand-in for a third-party class the reference names at import time and setReference never touches."""
    def __init__(self, *a, **k):
        pass

    def __getattr__(self, name):
        return _Anything()

    def __call__(self, *a, **k):
        return _Anything()


def _stub_module(name, **attrs):
    mod = types.ModuleType(name)
    mod.__dict__.update(attrs)
    mod.__getattr__ = lambda attr: _Anything        # any other name: a class-like stand-in
    sys.modules[name] = mod
    return mod


def _tf_euler_matrix(ai, aj, ak, axes="sxyz"):
    """tf.transformations.euler_matrix (C. Gohlke's transformations.py, as shipped with ROS tf) for the default axes 'sxyz':
    firstaxis 0, parity 0, repetition 0, static frame -> i, j, k = 0, 1, 2.  Third-party, absent here: restated."""
    assert axes == "sxyz"
    import math
    si, sj, sk = math.sin(ai), math.sin(aj), math.sin(ak)
    ci, cj, ck = math.cos(ai), math.cos(aj), math.cos(ak)
    cc, cs = ci * ck, ci * sk
    sc, ss = si * ck, si * sk
    M = np.identity(4)
    M[0, 0] = cj * ck; M[0, 1] = sj * sc - cs; M[0, 2] = sj * cc + ss
    M[1, 0] = cj * sk; M[1, 1] = sj * ss + cc; M[1, 2] = sj * cs - sc
    M[2, 0] = -sj; M[2, 1] = cj * si; M[2, 2] = cj * ci
    return M


def load_wbid():
    """The reference's WholeBodyID class (g1_mujoco_sim/src/wbid.py), imported unmodified."""
    x = _stub_module("xbot2_interface"); x.pyxbot2_interface = _stub_module("xbot2_interface.pyxbot2_interface")
    for name in ("pyopensot", "pyopensot.tasks", "pyopensot.tasks.acceleration", "pyopensot.constraints", "pyopensot.constraints.acceleration",
                 "pyopensot.constraints.force", "pyopensot.variables", "pyopensot.solver_back_ends"):
        _stub_module(name)
    sys.modules["pyopensot"].solver_back_ends = sys.modules["pyopensot.solver_back_ends"]
    tf = _stub_module("tf"); tf.transformations = _stub_module("tf.transformations", euler_matrix=_tf_euler_matrix)
    _stub_module("ttictoc", tic=lambda *a: None, toc=lambda *a: 0.0)
    spec = importlib.util.spec_from_file_location("g1_ref_wbid", REF_WBID_PY)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.WholeBodyID


class _Affine:
    def __init__(self):
        self.linear = np.eye(3)


class _Recorder:
    def __init__(self):
        self.calls = []

    def getReference(self):
        return [_Affine(), np.zeros(6), np.zeros(6)]

    def setReference(self, *args):
        self.calls.append([np.array(getattr(a, "linear", a), float).copy() for a in args])


class _Model:
    def __init__(self, mass):
        self._m = mass

    def getMass(self):
        return self._m


def reference_wbid_set_reference(WBID, x_opt1, u_opt0, foot_positions_curr, mass, t=0.0):
    """WholeBodyID.setReference(t, x_opt1, u_opt0, foot_positions_curr) (wbid.py:231-297) on an object whose tasks record what they
    are handed.  Returns the 42 numbers in the order of g1mpc_wbid_refs_batch: R | velocity | acceleration | CoM pos, vel, acc | wrenches."""
    o = object.__new__(WBID)
    o.base, o.com, o.wrench_tasks = _Recorder(), _Recorder(), [_Recorder() for _ in range(4)]
    o.contact_frames = ["left_foot_line_contact_lower", "left_foot_line_contact_upper", "right_foot_line_contact_lower", "right_foot_line_contact_upper"]
    o.model = _Model(mass)
    o.setReference(t, np.array(x_opt1, float), np.array(u_opt0, float), [list(map(float, f)) for f in np.asarray(foot_positions_curr, float).reshape(4, 3)])
    (aff, vel, acc), = o.base.calls
    (cp, cv, ca), = o.com.calls
    wr = [w.calls[0][0] for w in o.wrench_tasks]
    return np.concatenate([aff.ravel(), vel.ravel(), acc.ravel(), cp.ravel(), cv.ravel(), ca.ravel()] + [w.ravel() for w in wr])


def run_reference_main(max_ticks=None, per_tick=False):
    """g1_mpc/scripts/main.py main() run as shipped (planner GaitPlanner(1000., 0.25, 0.1, 0.2, 100), MPC() defaults, total_duration 4 s,
    main.py:200-326), headless: matplotlib and srbd_plotting are stand-ins, `mpc` is the reference module behind the OSQP shim.
    Returns (traj_x[ticks,13], traj_u[ticks,12]) = x_opt[0] and u_opt[0] after each tick's compute_rollout -- what main() records
    (main.py:313-322).  max_ticks: stop early (raises out of the loop) to keep a test short.  per_tick: also the list of every tick's
    x0, x_ref, contact, c_horizon, p_com as main() handed them to the class, and the u_opt / x_opt it got back."""
    import contextlib, io
    mpl = _stub_module("matplotlib")
    mpl.pyplot = _stub_module("matplotlib.pyplot"); mpl.animation = _stub_module("matplotlib.animation")
    _stub_module("srbd_plotting")
    gp_spec = importlib.util.spec_from_file_location("gait_planner", REF_GAIT_PY)
    gp = importlib.util.module_from_spec(gp_spec); gp_spec.loader.exec_module(gp)
    sys.modules["gait_planner"] = gp
    ref = load_module()
    sys.modules["mpc"] = ref
    tx, tu, ticks = [], [], []
    cur = {}

    class _Stop(Exception):
        pass

    orig = {k: getattr(ref.MPC, k) for k in ("compute_rollout", "calculate_B_continuous", "calculate_bounds")}

    def recording_B(self, c_horizon, p_com_horizon):                  # what main() hands the build (main.py:295)
        cur["c_horizon"] = np.array(c_horizon, float).copy(); cur["p_com"] = np.array(p_com_horizon, float).copy()
        return orig["calculate_B_continuous"](self, c_horizon, p_com_horizon)

    def recording_bounds(self, contact):                              # (main.py:300)
        cur["contact"] = np.array(contact, float).copy()
        return orig["calculate_bounds"](self, contact)

    def recording_rollout(self, *a, **k):
        out = orig["compute_rollout"](self, *a, **k)
        tx.append(np.array(self.x_opt[0], float).copy()); tu.append(np.array(self.u_opt[0], float).copy())
        ticks.append(dict(cur, x0=np.array(self.x0, float).ravel().copy(), x_ref=np.array(self.x_ref_hor, float).copy(),
                          u_opt=np.array(self.u_opt, float).copy(), x_opt=np.array(self.x_opt, float).copy()))
        if max_ticks is not None and len(tx) >= max_ticks:
            raise _Stop()
        return out

    ref.MPC.compute_rollout, ref.MPC.calculate_B_continuous, ref.MPC.calculate_bounds = recording_rollout, recording_B, recording_bounds
    try:
        spec = importlib.util.spec_from_file_location("g1_ref_main", REF_MAIN_PY)
        mod = importlib.util.module_from_spec(spec)
        with contextlib.redirect_stdout(io.StringIO()):
            spec.loader.exec_module(mod)
            try:
                mod.main()
            except _Stop:
                pass
    finally:
        for k, f in orig.items():
            setattr(ref.MPC, k, f)
        for name in ("mpc", "gait_planner", "srbd_plotting"):
            sys.modules.pop(name, None)
    if per_tick:
        return np.array(tx), np.array(tu), ticks
    return np.array(tx), np.array(tu)


REF_ROS_SIM_PY = "/root/reference/g1_mujoco_sim/src/ros_run_simulation.py"


class RosMsg:
    """Attribute container standing in for a generated ROS message class: nested fields appear on first access, the two list fields
    of g1_msgs/SRBD_state start empty."""
    def __init__(self, *a, **k):
        self.__dict__["states_horizon"] = []
        self.__dict__["contacts"] = []

    def __getattr__(self, name):
        v = RosMsg()
        self.__dict__[name] = v
        return v


def load_ros_simulation():
    """g1_mujoco_sim/src/ros_run_simulation.py imported unmodified with stand-ins for everything it names at import time (mujoco,
    rospy, tf, message packages, its sibling modules).  Only publish_current_state (:22-78) and G1MujocoSimulation.callback_mpc_solution
    (:188-218) are used: pure field-by-field packing / unpacking."""
    for name in ("mujoco", "mujoco_viewer", "rospkg", "tf", "swing_trajectory", "pal_statistics", "rosgraph_msgs", "rosgraph_msgs.msg",
                 "config", "wbid", "viz"):
        _stub_module(name)
    rospy = _stub_module("rospy")

    class _Time:
        @staticmethod
        def from_sec(t):
            return ("rospy.Time.from_sec", float(t))         # third-party (genpy); the stamp is not part of what is pinned here
    rospy.Time = _Time
    _stub_module("ttictoc", tic=lambda *a: None, toc=lambda *a: 0.0)
    _stub_module("g1_msgs")
    _stub_module("g1_msgs.msg", SRBD_state=RosMsg, ContactPoint=RosMsg, Feet_reference=RosMsg, State=RosMsg)
    spec = importlib.util.spec_from_file_location("g1_ref_ros_sim", REF_ROS_SIM_PY)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    for name in ("config", "wbid", "viz", "swing_trajectory", "tf"):
        sys.modules.pop(name, None)
    return mod


def _v3(m):
    return [float(m.x), float(m.y), float(m.z)]


def reference_publish_current_state(mod, base_orientation, com_position, base_angular_velocity, com_linear_velocity, foot_positions, qp_forces, sim_time):
    """publish_current_state(...) (ros_run_simulation.py:22-78) with a recording publisher -> the message as plain numbers:
    state row [13] (gravity as assigned, a Python float), contact names, positions [12], forces [12]."""
    class _Pub:
        def publish(self, m):
            self.m = m
    pub = _Pub()
    mod.publish_current_state(pub, None, base_orientation, com_position, base_angular_velocity, com_linear_velocity,
                              np.asarray(foot_positions, float).reshape(4, 3), qp_forces, sim_time)
    m = pub.m
    st = m.states_horizon[0]
    row = _v3(st.orientation) + _v3(st.position) + _v3(st.angular_velocity) + _v3(st.linear_velocity) + [float(st.gravity)]
    return dict(n_states=len(m.states_horizon), trajectory_index=int(st.trajectory_index), state=np.array(row), frame_id=str(m.header.frame_id),
                names=[str(c.name) for c in m.contacts], positions=np.array([_v3(c.position) for c in m.contacts]).ravel(),
                forces=np.array([_v3(c.force) for c in m.contacts]).ravel())


def msg_from_dict(d):
    """A dict message of g1_locomotion_b200.wire as the attribute object rospy would deliver."""
    if isinstance(d, dict):
        m = RosMsg()
        for k, v in d.items():
            m.__dict__[k] = msg_from_dict(v)
        return m
    if isinstance(d, list):
        return [msg_from_dict(v) for v in d]
    return d


def reference_callback_mpc_solution(mod, msg_obj):
    """G1MujocoSimulation.callback_mpc_solution (ros_run_simulation.py:188-218) on a bare object -> x_opt, u_opt0, contact_states, landing."""
    o = object.__new__(mod.G1MujocoSimulation)
    o.u_opt0 = np.zeros(12); o.contact_states = [None] * 4
    o.callback_mpc_solution(msg_obj)
    return np.array(o.x_opt, float), np.array(o.u_opt0, float), [bool(c) for c in o.contact_states], np.array(o.foot_in_swing_final_position, float)
