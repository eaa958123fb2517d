"""Independent third-party check of the solve-half oracle: the convex-QP solver of HiGHS (the copy bundled with scipy,
`scipy.optimize._highspy._core`; scipy itself only exposes its LP/MIP side) on the reference-built matrices
    min 1/2 x'Hx + g'x   s.t.  l <= Acon x <= u            (mpc.cpp:378-414 hands exactly these to OSQP).
Test infrastructure only."""
import numpy as np
import scipy.sparse as sp


def available():
    try:
        from scipy.optimize._highspy import _core as hc
        return all(hasattr(hc, n) for n in ("HighsModel", "HighsHessian", "_Highs", "HessianFormat"))
    except Exception:
        return False


def solve(H, g, A, l, u):
    """-> (x, row_dual, status string)"""
    from scipy.optimize._highspy import _core as hc
    n, m = H.shape[0], A.shape[0]
    inf = hc.kHighsInf
    model = hc.HighsModel()
    lp = model.lp_
    lp.num_col_, lp.num_row_ = n, m
    lp.col_cost_ = np.asarray(g, float)
    lp.col_lower_ = np.full(n, -inf)
    lp.col_upper_ = np.full(n, inf)
    lp.row_lower_ = np.where(np.isfinite(l), l, -inf)
    lp.row_upper_ = np.where(np.isfinite(u), u, inf)
    Ac = sp.csc_matrix(A)
    lp.a_matrix_.format_ = hc.MatrixFormat.kColwise
    lp.a_matrix_.start_, lp.a_matrix_.index_, lp.a_matrix_.value_ = Ac.indptr, Ac.indices, Ac.data
    Hl = sp.csc_matrix(np.tril(H))
    hs = model.hessian_
    hs.dim_, hs.format_ = n, hc.HessianFormat.kTriangular
    hs.start_, hs.index_, hs.value_ = Hl.indptr, Hl.indices, Hl.data
    h = hc._Highs()
    h.setOptionValue("output_flag", False)
    h.passModel(model)
    h.run()
    sol = h.getSolution()
    return np.array(sol.col_value), np.array(sol.row_dual), str(h.getModelStatus())
