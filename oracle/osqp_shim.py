"""Stand-in for the third-party `osqp` Python module -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

The reference's g1_mpc/scripts/mpc.py does `import osqp` (mpc.py:1) and, in solve_qp() (mpc.py:323-358):

    self.qp_solver = osqp.OSQP()
    self.qp_solver.setup(P, q, A, l, u, verbose=False, warm_start=True)      # mpc.py:341-342
    result = self.qp_solver.solve()                                           # mpc.py:348
    self.u_opt = result.x.reshape((HORIZON_LENGTH, NUM_CONTROLS))             # mpc.py:352

pip `osqp` is not on this image and cannot be installed.  This module exposes exactly that surface over
oracle.osqp_solve (oracle/srbd_mpc_oracle.c: orc_osqp_solve, the restatement of the published OSQP algorithm with
OSQP 0.6.x default settings; constants recalled, unverified -> the solver boundary stays "parity unpinned").
With it on sys.modules["osqp"] the UNMODIFIED reference MPC.update() (mpc.py:360-377) runs end to end, so the
reference's own control flow -- the P/q/A/l/u hand-over, the row-major reshape of result.x and compute_rollout
after a solve -- is exercised by the reference's code (tests/test_reference_e2e_cpu.py) and can be timed as the
"reference path" CPU baseline (oracle/time_reference_py.py).

Semantics kept from osqp 0.6.x: P may be given as a full symmetric sparse matrix (only its upper triangle is
used -- the matrix is symmetric, so the dense solver underneath takes it whole); q of shape (n,) or (n,1);
l/u with +-inf; a fresh object is a cold start (x = z = y = 0), so warm_start=True has no effect on the first
solve, which is the only one the reference ever makes per object (mpc.py:340).
"""
import types

import numpy as np

from . import oracle

_SETTING_NAMES = ("rho", "sigma", "alpha", "eps_abs", "eps_rel", "max_iter", "check_termination", "scaling",
                  "adaptive_rho", "adaptive_rho_interval", "adaptive_rho_tolerance")


def _dense(M):
    return np.asarray(M.toarray() if hasattr(M, "toarray") else M, dtype=np.float64)


class OSQP:
    def __init__(self):
        self._data = None

    def setup(self, P=None, q=None, A=None, l=None, u=None, **settings):
        P, A = _dense(P), _dense(A)
        n, m = P.shape[0], A.shape[0]
        P = np.triu(P) + np.triu(P, 1).T                # osqp reads the upper triangle only
        q = np.asarray(q, dtype=np.float64).reshape(n)
        l = np.asarray(l, dtype=np.float64).reshape(m)
        u = np.asarray(u, dtype=np.float64).reshape(m)
        kw = {k: (int(v) if isinstance(v, (bool, np.bool_)) else v) for k, v in settings.items() if k in _SETTING_NAMES}
        self._settings = oracle.default_settings(**kw)   # verbose / warm_start / polish(False) are no-ops here
        self._data = (P, q, A, l, u)

    def solve(self):
        if self._data is None:
            raise ValueError("Workspace not initialized!")       # what osqp raises before setup()
        x, y, info = oracle.osqp_solve(*self._data, settings=self._settings)
        res = types.SimpleNamespace()
        res.x, res.y = x, y
        res.info = types.SimpleNamespace(iter=info["iters"], status="solved" if info["status"] == 1 else "maximum iterations reached",
                                         status_val=1 if info["status"] == 1 else -2, pri_res=info["prim_res"],
                                         dua_res=info["dual_res"], rho_updates=info["rho_updates"], rho_estimate=info["rho"],
                                         obj_val=float(0.5 * x @ self._data[0] @ x + self._data[1] @ x))
        return res
