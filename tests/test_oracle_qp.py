"""The solve-half oracle: OSQP restatement (parity unpinned: OSQP source/version not in the reference)
and the exact minimiser x* with its algorithm-independent KKT certificate.  CPU only."""
import numpy as np
import pytest

from conftest import load_golden
from g1_locomotion_b200 import presets, inputs


def _problems(orc, pname, B, seed, kind="random", **over):
    preset = presets.get_preset(pname, **over)
    p = orc.params_from_dict(preset)
    if kind == "walk":
        batch = inputs.walking_batch(preset, B, seed0=seed)
    else:
        batch = [t.numpy() for t in inputs.random_batch(preset, B, seed=seed)]
    for b in range(B):
        args = [a[b] for a in batch]
        yield preset, p, args, orc.build(p, *args)


@pytest.mark.parametrize("pname,kind", [("cpp_node", "walk"), ("cpp_node", "random"), ("python_default", "random")])
def test_exact_minimiser_has_kkt_certificate(orc, pname, kind):
    for preset, p, args, m in _problems(orc, pname, 4, 1000 if kind == "walk" else 7, kind,
                                        **({"horizon": 10} if pname == "python_default" else {})):
        x, y, info = orc.qp_exact(p, m["H"], m["g"], args[2])
        assert info["rc"] == 0, info
        r = orc.kkt_residuals(m["H"], m["g"], m["Acon"], m["l"], m["u"], x, y)
        gscale = 1.0 + np.abs(m["g"]).max()
        assert r["stationarity"] <= 1e-7 * gscale, r
        assert r["primal"] <= 1e-9, r
        assert r["complementarity"] <= 1e-6 * gscale, r
        # strong convexity: lambda_min(H) = cost_scale * r_weight  =>  ||x - x*||_2 <= ||stationarity||_2 / lambda_min
        lam_min = np.linalg.eigvalsh(m["H"])[0]
        assert lam_min >= 0.99 * preset["cost_scale"] * preset["r_weight"]
        # swing contacts carry zero force (8a a12)
        swing = np.repeat(args[2].ravel() == 0, 3)
        assert np.all(x[swing] == 0.0)


def test_osqp_restatement_terminates_like_osqp_defaults(orc):
    """Default settings (eps 1e-3, check every 25): iterations are a multiple of 25, status solved, and the
    returned point satisfies OSQP's own unscaled termination test -- but is far from x* in force space
    (SURVEY.md 7.3-1), which is why GRF parity is defined against x*."""
    far = []
    for preset, p, args, m in _problems(orc, "cpp_node", 4, 1000, "walk"):
        x, y, info = orc.osqp_solve(m["H"], m["g"], m["Acon"], m["l"], m["u"])
        assert info["status"] == 1 and info["iters"] % 25 == 0 and info["iters"] <= 4000
        z = np.clip(m["Acon"] @ x, m["l"], m["u"])
        Ax = m["Acon"] @ x
        prim = np.abs(Ax - z).max()
        dual = np.abs(m["H"] @ x + m["g"] + m["Acon"].T @ y).max()
        assert prim <= 1e-3 + 1e-3 * max(np.abs(Ax).max(), np.abs(z).max()) + 1e-9
        assert dual <= 1e-3 + 1e-3 * max(np.abs(m["H"] @ x).max(), np.abs(m["Acon"].T @ y).max(), np.abs(m["g"]).max())
        xs, _, _ = orc.qp_exact(p, m["H"], m["g"], args[2])
        far.append(np.abs(x - xs).max())
        # objective is nevertheless close
        f = lambda v: 0.5 * v @ m["H"] @ v + m["g"] @ v
        assert abs(f(x) - f(xs)) <= 5e-3 * abs(f(xs))
    assert max(far) > 1.0     # Newtons


def test_osqp_tight_tolerance_approaches_exact(orc):
    for preset, p, args, m in _problems(orc, "python_default", 2, 9, horizon=10):
        st = orc.default_settings(eps_abs=1e-9, eps_rel=1e-9, max_iter=100000, adaptive_rho_interval=50)
        x, y, info = orc.osqp_solve(m["H"], m["g"], m["Acon"], m["l"], m["u"], st)
        xs, _, cert = orc.qp_exact(p, m["H"], m["g"], args[2])
        assert cert["rc"] == 0
        assert np.abs(x - xs).max() < 5e-2, (np.abs(x - xs).max(), info)


def test_update_batch_threads_agree(orc):
    preset = presets.get_preset("cpp_node")
    p = orc.params_from_dict(preset)
    batch = inputs.walking_batch(preset, 6)
    u1, x1, it1 = orc.update_batch(p, *batch, x_current=batch[0], nthreads=1)
    u2, x2, it2 = orc.update_batch(p, *batch, x_current=batch[0], nthreads=4)
    assert np.array_equal(u1, u2) and np.array_equal(x1, x2) and np.array_equal(it1, it2)
    assert np.all(x1[:, 0] == batch[0])


@pytest.mark.parametrize("pname,kind,over", [("cpp_node", "walk", {}), ("cpp_node", "random", {}),
                                             ("python_default", "random", {"horizon": 10}), ("cpp_node", "random", {"num_contacts": 8})])
def test_exact_minimiser_agrees_with_highs_qp(orc, pname, kind, over):
    """The exact minimiser x* against an independent third-party solver: HiGHS' convex-QP solver (bundled with scipy) on
    the reference-built H, g, Acon, l, u.  Same optimal value to 1e-11 relative, x* never the worse point, and the
    forces agree to 2e-2 N -- the accuracy HiGHS reaches on cond(H) ~ 3e8 (its point differs from x* along the directions
    where H is cost_scale r_weight = 2e-3, which move the objective by 1e-7 of 2e6)."""
    import highs_qp
    if not highs_qp.available():
        pytest.skip("scipy without the bundled HiGHS core")
    for preset, p, args, m in _problems(orc, pname, 4, 1000 if kind == "walk" else 7, kind, **over):
        xs, _, info = orc.qp_exact(p, m["H"], m["g"], args[2])
        assert info["rc"] == 0
        xh, yh, status = highs_qp.solve(m["H"], m["g"], m["Acon"], m["l"], m["u"])
        assert status.endswith("kOptimal"), status
        f = lambda x: 0.5 * x @ m["H"] @ x + m["g"] @ x          # noqa: E731
        assert f(xs) <= f(xh) + 1e-12 * abs(f(xs))
        assert abs(f(xh) - f(xs)) <= 1e-11 * abs(f(xs))
        assert np.abs(xh - xs).max() <= 2e-2
        # rows HiGHS holds active (non-zero dual) are at their bound at x* as well
        ax = m["Acon"] @ xs
        act = np.abs(yh) > 1e-6 * (1.0 + np.abs(yh).max())
        at_bound = np.minimum(np.abs(ax - m["l"]), np.abs(m["u"] - ax)) <= 1e-6
        assert np.all(at_bound[act])
