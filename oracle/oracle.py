"""ctypes binding of oracle/liboracle.so (built from oracle/srbd_mpc_oracle.c by oracle/Makefile).

TEST INFRASTRUCTURE ONLY -- see oracle/__init__.py.  Every function here is a thin wrapper; the
arithmetic (and the reference file:line each step follows) is in the C file.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
NS = 13


class OrcParams(C.Structure):
    _fields_ = [("N", C.c_int32), ("nc", C.c_int32), ("mu", C.c_double), ("fz_min", C.c_double),
                ("fz_max", C.c_double), ("dt", C.c_double), ("mass", C.c_double), ("Ib", C.c_double * 9),
                ("q", C.c_double * 13), ("r", C.c_double), ("scale", C.c_double)]


class OsqpSettings(C.Structure):
    _fields_ = [("rho", C.c_double), ("sigma", C.c_double), ("alpha", C.c_double), ("eps_abs", C.c_double),
                ("eps_rel", C.c_double), ("max_iter", C.c_int32), ("check_termination", C.c_int32),
                ("scaling", C.c_int32), ("adaptive_rho", C.c_int32), ("adaptive_rho_interval", C.c_int32),
                ("adaptive_rho_tolerance", C.c_double)]


class OsqpInfo(C.Structure):
    _fields_ = [("iters", C.c_int32), ("status", C.c_int32), ("rho_updates", C.c_int32),
                ("prim_res", C.c_double), ("dual_res", C.c_double), ("rho_final", C.c_double)]


def build_lib(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "srbd_mpc_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build_lib())
        _LIB.orc_num_threads.restype = C.c_int
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


def make_params(N, nc, mu, fz_min, fz_max, dt, mass, inertia_body, q_weights, r_weight, cost_scale):
    p = OrcParams()
    p.N, p.nc, p.mu, p.fz_min, p.fz_max, p.dt, p.mass = N, nc, mu, fz_min, fz_max, dt, mass
    p.Ib[:] = list(np.asarray(inertia_body, float).ravel())
    p.q[:] = list(np.asarray(q_weights, float).ravel())
    p.r, p.scale = r_weight, cost_scale
    return p


def params_from_dict(d):
    return make_params(d["horizon"], d["num_contacts"], d["mu"], d["fz_min"], d["fz_max"], d["dt"], d["mass"],
                       d["inertia_body"], d["q_weights"], d["r_weight"], d["cost_scale"])


def default_settings(**kw):
    s = OsqpSettings()
    lib().orc_osqp_default_settings(C.byref(s))
    for k, v in kw.items():
        setattr(s, k, v)
    return s


def build(p, x0, xref, contact, c_hor, p_com):
    """Reference build half for ONE problem -> dict of dense matrices (mpc.cpp:153-376)."""
    N, nc = p.N, p.nc
    nu, n, nx, m = 3 * nc, 3 * nc * N, NS * N, 5 * nc * N
    x0, xref, contact, c_hor, p_com = map(_f64, (x0, xref, contact, c_hor, p_com))
    out = dict(A=np.zeros((NS, NS)), Bd=np.zeros((nx, nu)), Aqp=np.zeros((nx, NS)), Bqp=np.zeros((nx, n)),
               H=np.zeros((n, n)), g=np.zeros(n), Acon=np.zeros((m, n)), l=np.zeros(m), u=np.zeros(m))
    lib().orc_build(C.byref(p), _p(x0), _p(xref), _p(contact), _p(c_hor), _p(p_com), _p(out["A"]), _p(out["Bd"]),
                    _p(out["Aqp"]), _p(out["Bqp"]), _p(out["H"]), _p(out["g"]), _p(out["Acon"]), _p(out["l"]),
                    _p(out["u"]))
    return out


def osqp_solve(P, q, A, l, u, settings=None):
    """OSQP restatement (cold start) -> x, y, info dict."""
    P, q, A, l, u = map(_f64, (P, q, A, l, u))
    n, m = P.shape[0], A.shape[0]
    st = settings or default_settings()
    x, y, info = np.zeros(n), np.zeros(m), OsqpInfo()
    rc = lib().orc_osqp_solve(n, m, _p(P), _p(q), _p(A), _p(l), _p(u), C.byref(st), _p(x), _p(y), C.byref(info))
    return x, y, dict(rc=rc, iters=info.iters, status=info.status, rho_updates=info.rho_updates,
                      prim_res=info.prim_res, dual_res=info.dual_res, rho=info.rho_final)


def qp_exact(p, H, g, contact):
    """Unique minimiser x* (+ multipliers, KKT certificate) of the reference QP."""
    H, g, contact = map(_f64, (H, g, contact))
    n, m = H.shape[0], 5 * p.nc * p.N
    x, y, cert = np.zeros(n), np.zeros(m), np.zeros(4)
    rc = lib().orc_qp_exact(C.byref(p), _p(H), _p(g), _p(contact), _p(x), _p(y), _p(cert))
    return x, y, dict(rc=rc, stationarity=cert[0], primal_violation=cert[1], wrong_sign=cert[2],
                      changes=int(cert[3]))


def rollout(p, A, Bd, u_opt, x_current=None, only_first_step=False, x_opt=None):
    A, Bd, u_opt = map(_f64, (A, Bd, u_opt))
    x_opt = np.zeros((p.N, NS)) if x_opt is None else _f64(x_opt).copy()
    xc = None if x_current is None else _f64(x_current)
    lib().orc_rollout(C.byref(p), _p(A), _p(Bd), _p(u_opt), _p(xc), int(only_first_step), _p(x_opt))
    return x_opt


def update_batch(p, x0, xref, contact, c_hor, p_com, x_current=None, one_rollout=False, settings=None,
                 nthreads=0, x_opt=None):
    """MPC::updateMPC over a batch (OpenMP over problems) -> u_opt[B,N,nu], x_opt[B,N,13], iters[B]."""
    x0, xref, contact, c_hor, p_com = map(_f64, (x0, xref, contact, c_hor, p_com))
    B, N, nu = x0.shape[0], p.N, 3 * p.nc
    st = settings or default_settings()
    u_opt = np.zeros((B, N, nu))
    x_opt = np.zeros((B, N, NS)) if x_opt is None else _f64(x_opt).copy()
    iters = np.zeros(B, np.int32)
    xc = None if x_current is None else _f64(x_current)
    rc = lib().orc_update_batch(C.byref(p), B, _p(x0), _p(xref), _p(contact), _p(c_hor), _p(p_com), _p(xc),
                                int(one_rollout), C.byref(st), _p(u_opt), _p(x_opt), _p(iters), int(nthreads))
    if rc:
        raise RuntimeError("oracle: KKT factorisation failed")
    return u_opt, x_opt, iters


def num_threads():
    return int(lib().orc_num_threads())


def kkt_residuals(H, g, Acon, l, u, x, y):
    """Algorithm-independent optimality certificate (numpy): stationarity, primal violation,
    complementarity/sign violation, in OSQP's sign convention."""
    H, g, Acon, l, u, x, y = map(_f64, (H, g, Acon, l, u, x, y))
    z = Acon @ x
    stat = np.abs(H @ x + g + Acon.T @ y).max()
    prim = max(np.maximum(l - z, 0).max(), np.maximum(z - u, 0).max())
    with np.errstate(invalid="ignore"):
        gap_lo = np.where(np.isfinite(l), z - l, np.inf)
        gap_hi = np.where(np.isfinite(u), u - z, np.inf)
    ym, yp = np.minimum(y, 0), np.maximum(y, 0)
    gap_lo_f = np.where(np.isfinite(gap_lo), gap_lo, 0.0)
    gap_hi_f = np.where(np.isfinite(gap_hi), gap_hi, 0.0)
    comp = max(np.abs(np.where(np.isfinite(gap_lo), ym * gap_lo_f, np.where(ym != 0, np.inf, 0))).max(),
               np.abs(np.where(np.isfinite(gap_hi), yp * gap_hi_f, np.where(yp != 0, np.inf, 0))).max())
    return dict(stationarity=stat, primal=prim, complementarity=comp)
