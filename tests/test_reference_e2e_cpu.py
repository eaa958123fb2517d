"""The reference's MPC.update() run END TO END (mpc.py:360-377) -- CPU only.

`oracle/osqp_shim.py` supplies the `osqp` module the reference imports (pip osqp is absent), so the unmodified class
runs build -> solve_qp -> compute_rollout.  What this pins, with the reference's own code rather than a reading of it:
the hand-over of P, q, A, l, u to the solver (mpc.py:325-342), the row-major reshape of result.x into u_opt
(mpc.py:352), and compute_rollout after a solve for the three call patterns its callers use (mpc.py:379-386).
The solver arithmetic itself stays the restated OSQP ("parity unpinned" at that boundary).

Golden files tests/golden/ref_update_*.npz were written by oracle/gen_golden.py from the imported reference; the live
tests repeat the comparison against the reference class when /root/reference is present (build container only)."""
import glob
import json
import os

import numpy as np
import pytest

from conftest import ROOT, rel_err
from g1_locomotion_b200 import presets, inputs

CASES = sorted(os.path.basename(f)[len("ref_update_"):-len(".npz")]
               for f in glob.glob(os.path.join(ROOT, "tests", "golden", "ref_update_*.npz")))


def _load(name):
    z = np.load(os.path.join(ROOT, "tests", "golden", f"ref_update_{name}.npz"))
    d = {k: z[k] for k in z.files}
    d["preset"] = json.loads(str(d["preset"]))
    return d


def test_golden_cases_exist():
    assert len(CASES) >= 3


@pytest.mark.parametrize("name", CASES)
def test_oracle_update_matches_reference_update_golden(orc, name):
    """oracle.update_batch (C: mpc.cpp build as coded + the same OSQP restatement) against the reference's update().
    Both run the same ADMM from matrices that agree to 1e-16 relative; the ADMM map of this QP (cond(H) up to 3e8)
    amplifies that rounding to ~1e-7 N over 200-350 iterations, hence 1e-5 N / 1e-9 on the states -- the control flow,
    the ordering of u_opt and every rollout variant must match for that."""
    gd = _load(name)
    p = orc.params_from_dict(gd["preset"])
    args = [gd[k][None] for k in ("x0", "x_ref", "contact", "c_horizon", "p_com")]
    for tag, xc, one in (("full", gd["x0"][None], False), ("one", gd["x0"][None], True), ("nocur", None, False)):
        u, x, iters = orc.update_batch(p, *args, x_current=xc, one_rollout=one)
        assert np.abs(u[0] - gd["u_opt_" + tag]).max() <= 1e-5, (tag, np.abs(u[0] - gd["u_opt_" + tag]).max())
        assert np.abs(x[0] - gd["x_opt_" + tag]).max() <= 1e-9, tag
    # the one_rollout call leaves rows 2.. of x_opt untouched, the x_current=None call leaves row 0 (zeros here)
    assert np.all(gd["x_opt_one"][2:] == 0.0) and np.all(gd["x_opt_nocur"][0] == 0.0)


@pytest.mark.parametrize("name", CASES)
def test_reference_solver_handover_and_reshape_bit_exact(orc, name):
    """Same solver, the REFERENCE's matrices: feeding the Hessian/gradient the reference built (stored in the golden
    file) and the reference-layout Ac, l, u to the restated OSQP must reproduce the reference's u_opt bit for bit --
    i.e. result.x.reshape((N, nu)) is step-major, contact-major, xyz (mpc.py:352) and nothing is permuted or rescaled
    on the way into the solver."""
    gd = _load(name)
    p = orc.params_from_dict(gd["preset"])
    m = orc.build(p, gd["x0"], gd["x_ref"], gd["contact"], gd["c_horizon"], gd["p_com"])
    Hu = np.triu(gd["H"]) + np.triu(gd["H"], 1).T          # osqp reads the upper triangle of P only (as the shim does)
    x, _, info = orc.osqp_solve(Hu, gd["g"], m["Acon"], m["l"], m["u"])
    N, nu = gd["preset"]["horizon"], 3 * gd["preset"]["num_contacts"]
    assert np.array_equal(x.reshape(N, nu), gd["u_opt_full"])
    assert rel_err(m["H"], gd["H"]) <= 1e-12 and rel_err(m["g"], gd["g"]) <= 1e-12
    # and the rollout the reference ran after the solve is the oracle's rollout of that u_opt
    xo = orc.rollout(p, m["A"], m["Bd"], gd["u_opt_full"], gd["x0"], False)
    assert np.abs(xo - gd["x_opt_full"]).max() <= 1e-12 * max(1.0, np.abs(xo).max())


@pytest.mark.parametrize("pname,over,kind", [("python_default", {}, "random"), ("cpp_node", {"cost_scale": 1.0}, "walk"),
                                             ("python_default", {"horizon": 10}, "walk")])
def test_live_reference_update_against_oracle(orc, pname, over, kind):
    from oracle import ref_loader
    if not ref_loader.available():
        pytest.skip("reference tree not present (GPU box)")
    preset = presets.get_preset(pname, **over)
    p = orc.params_from_dict(preset)
    ref = ref_loader.make_reference_mpc(preset)
    if kind == "walk":
        batch = inputs.walking_batch(preset, 4, seed0=77)
    else:
        batch = [t.numpy() for t in inputs.random_batch(preset, 4, seed=31)]
    for b in range(4):
        args = [a[b] for a in batch]
        ref.x_opt = np.zeros((preset["horizon"], 13))
        u, x = ref_loader.reference_update(ref, *args, x_current=args[0], one_rollout=(b % 2 == 1))
        uo, xo, _ = orc.update_batch(p, *[a[None] for a in args], x_current=args[0][None], one_rollout=(b % 2 == 1))
        assert np.abs(u - uo[0]).max() <= 1e-5 and np.abs(x - xo[0]).max() <= 1e-9
        # the reference leaves every intermediate current after update() (mpc.py:360-377): H, g are those of this call
        m = orc.build(p, *args)
        assert rel_err(ref.Hessian, m["H"]) <= 1e-12 and rel_err(ref.gradient.ravel(), m["g"]) <= 1e-12


def test_osqp_shim_surface():
    """The call sequence of mpc.py:340-348 on a tiny QP with a known solution."""
    from scipy import sparse
    from oracle import osqp_shim
    P = sparse.csc_matrix(np.array([[4.0, 1.0], [1.0, 2.0]]))
    A = sparse.csc_matrix(np.array([[1.0, 1.0], [1.0, 0.0], [0.0, 1.0]]))
    q = np.array([[1.0], [1.0]])
    s = osqp_shim.OSQP()
    with pytest.raises(ValueError):
        s.solve()
    s.setup(P, q, A, np.array([1.0, 0.0, 0.0]), np.array([1.0, 0.7, 0.7]), verbose=False, warm_start=True,
            eps_abs=1e-9, eps_rel=1e-9)
    r = s.solve()
    assert np.allclose(r.x, [0.3, 0.7], atol=1e-6) and r.info.status == "solved"
