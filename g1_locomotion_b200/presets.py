"""Parameter presets of the reference MPC (all constants are hard-coded in the reference).

`cpp_node`       : g1_mpc/src/mpc.cpp (the variant the shipped demo runs) -- ctor defaults mpc.hpp:22,
                   mass mpc.hpp:136, inertia mpc.cpp:25-27, Q mpc.cpp:90-96, R mpc.cpp:98, cost x2 mpc.cpp:361,375.
`python_default` : g1_mpc/scripts/mpc.py -- ctor defaults mpc.py:12, inertia mpc.py:24-28, Q mpc.py:47-50,
                   R mpc.py:51, cost x1 mpc.py:301,317.
"""
import copy

GRAVITY = -9.80665            # mpc.cpp:7 / mpc.py:14
ROBOT_MASS = 34.13385728      # mpc.hpp:136 / mpc.py:15

_CPP_NODE = dict(
    name="cpp_node", horizon=10, num_contacts=4, mu=0.8, fz_min=10.0, fz_max=1000.0, dt=0.05,
    mass=ROBOT_MASS, gravity=GRAVITY,
    inertia_body=[8.20564e-2, 0.0, 0.0, 0.0, 8.05015e-2, 0.0, 0.0, 0.0, 0.32353e-2],
    q_weights=[8e2, 75e0, 125e1, 3e4, 8e5, 8e5, 2e3, 4e3, 3e2, 5e2, 8e3, 5e2, 0.0],
    r_weight=0.001, cost_scale=2.0)

_PYTHON_DEFAULT = dict(
    name="python_default", horizon=15, num_contacts=4, mu=0.3, fz_min=10.0, fz_max=666.0, dt=0.04,
    mass=ROBOT_MASS, gravity=GRAVITY,
    inertia_body=[3.20564, 4.35027e-05, 0.425526, 4.35027e-05, 3.05015, -0.00065082, 0.425526, -0.00065082, 0.552353],
    q_weights=[7e5, 7e4, 1e4, 5e5, 5e5, 3e6, 3e3, 3e3, 3e3, 5e3, 1e3, 1e4, 0.0],
    r_weight=0.001, cost_scale=1.0)

PRESETS = {"cpp_node": _CPP_NODE, "python_default": _PYTHON_DEFAULT}


def get_preset(name, **overrides):
    """Return a deep copy of a preset with optional overrides (e.g. horizon=20, num_contacts=8)."""
    p = copy.deepcopy(PRESETS[name])
    for k, v in overrides.items():
        if k not in p:
            raise KeyError(f"unknown MPC parameter {k!r}")
        p[k] = v
    return p
