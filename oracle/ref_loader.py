"""Import the reference's own g1_mpc/scripts/mpc.py UNMODIFIED (read-only, from /root/reference).

Only usable in the build container (the GPU box has no /root/reference); used by oracle/gen_golden.py to
produce tests/golden/*.npz and by tests that are skipped when the reference tree is absent.
`import osqp` at mpc.py:1 is satisfied by oracle/osqp_shim.py (OSQP().setup/.solve over the C restatement of the
OSQP algorithm), so that solve_qp() (mpc.py:323-358) and with it the whole MPC.update() (mpc.py:360-377) run
unmodified; the build steps never touch it.  `import matplotlib.pyplot` at gait_planner.py:2 is satisfied by an
empty stand-in (only plot_gait() uses it).
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
