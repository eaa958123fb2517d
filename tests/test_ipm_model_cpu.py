"""CPU checks of the mathematics the CUDA solve kernel relies on (tests/ipm_model.py is its numpy model):
wrench-space factorisation of the reference Hessian, stage-wise solution of the Newton systems, and the
interior point as a whole against the oracle's exact minimiser."""
import numpy as np
import pytest

from g1_locomotion_b200 import presets, inputs
from ipm_model import Problem, RiccatiFactor, StageFactor, solve_ipm
import ipm_model as im

GRF_TOL = 1e-3


def _problems(pname, over, gen, n):
    preset = presets.get_preset(pname, **over)
    batch = inputs.walking_batch(preset, n) if gen == "walk" else [t.numpy() for t in inputs.random_batch(preset, n, seed=7)]
    return preset, [[a[b] for a in batch] for b in range(n)]


@pytest.mark.parametrize("pname,over", [("cpp_node", {}), ("python_default", {"horizon": 10}), ("cpp_node", {"num_contacts": 8})])
def test_wrench_space_form_equals_reference_hessian(orc, pname, over):
    """H = s r I + W' M W and g = W' h reproduce the reference build (mpc.cpp:359-376) to rounding."""
    preset, probs = _problems(pname, over, "rand", 2)
    p = orc.params_from_dict(preset)
    for args in probs:
        m = orc.build(p, *args)
        H, g = Problem(preset, *args).dense_H_g()
        assert np.abs(H - m["H"]).max() <= 1e-12 * np.abs(m["H"]).max()
        assert np.abs(g - m["g"]).max() <= 1e-12 * np.abs(m["g"]).max()


def test_stagewise_solve_equals_dense_newton_system():
    """(M^-1 + S) y = t by the backward/forward sweeps == dense solve, for barrier weights spanning 13 decades."""
    preset, probs = _problems("cpp_node", {}, "rand", 1)
    pr = Problem(preset, *probs[0])
    M, _ = pr.M_h()
    N = pr.N
    rng = np.random.default_rng(0)
    S = np.zeros((N, 6, 6))
    for k in range(N):
        d = 10 ** rng.uniform(-3, 10, pr.nu)
        S[k] = pr.W(k) @ np.diag(1 / d) @ pr.W(k).T
    Sd = np.zeros((6 * N, 6 * N))
    for k in range(N):
        Sd[6 * k:6 * k + 6, 6 * k:6 * k + 6] = S[k]
    t = rng.normal(size=(N, 6))
    want = M @ np.linalg.solve(np.eye(6 * N) + Sd @ M, t.ravel())
    for fac in (RiccatiFactor(pr, S), StageFactor(pr, S)):
        got = fac.solve(t).ravel()
        assert np.abs(got - want).max() <= 1e-9 * np.abs(want).max()


@pytest.mark.parametrize("pname,over,gen", [("cpp_node", {}, "walk"), ("cpp_node", {}, "rand"),
                                            ("python_default", {}, "rand"), ("cpp_node", {"horizon": 20}, "rand")])
def test_interior_point_model_reaches_exact_minimiser(orc, pname, over, gen):
    preset, probs = _problems(pname, over, gen, 2)
    p = orc.params_from_dict(preset)
    for args in probs:
        m = orc.build(p, *args)
        xs, _, cert = orc.qp_exact(p, m["H"], m["g"], args[2])
        assert cert["rc"] == 0
        u, it, gap = solve_ipm(Problem(preset, *args), linear="stage")
        assert gap <= 1e-10 and it <= 25
        assert np.abs(u.ravel() - xs).max() <= GRF_TOL


def _two_cycle_problems():
    """Two of the 131,072 random problems of one config-5 shard (rank 5 of 8, seed 18, generated on the GPU) on which
    the unsafeguarded Mehrotra iteration falls into a two-cycle (gap 8 <-> 18) and runs into max_iter."""
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "ipm_two_cycle_problems.npz"))
    return [[z[k][b] for k in ("x0", "x_ref", "contact", "c_horizon", "p_com")] for b in range(len(z["shard_index"]))]


def test_corrector_safeguard_breaks_the_two_cycle(orc):
    preset = presets.get_preset("cpp_node")
    p = orc.params_from_dict(preset)
    for args in _two_cycle_problems():
        m = orc.build(p, *args)
        xs, _, cert = orc.qp_exact(p, m["H"], m["g"], args[2])
        assert cert["rc"] == 0
        u, it, gap = solve_ipm(Problem(preset, *args), linear="dense")
        assert gap <= 1e-10 and it <= 25, (it, gap)
        assert np.abs(u.ravel() - xs).max() <= GRF_TOL


def test_illconditioned_problems_need_theta_in_the_stage_form(orc):
    """Two problems met at tick 3 of a closed loop started 11 m from the origin with x_opt = 0: the node's Raibert rule
    then plants the next stance foot at the world origin (mpc_to_wbid_node.cpp:430-470 with a zero x_opt), a lever arm
    of 11 m, cond(H) 4e11 instead of the usual 3e8.  The interior point converges on them in 18-19 iterations with a
    dense or textbook-Riccati Newton solve, and with the stage form as the kernel runs it (b_k = Gamma' t - Theta_k p2,
    Theta_k = Pi (I - Gamma)); forming S_k (Gamma p2) as a product instead loses the Newton direction to rounding
    (eps |S| |P22|) and the iteration stalls -- the kernel did until Theta_k went in."""
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "ipm_illconditioned_problems.npz"))
    preset = presets.get_preset("cpp_node")
    p = orc.params_from_dict(preset)
    lost = 0
    for b in range(len(z["walker"])):
        args = [z[k][b] for k in ("x0", "x_ref", "contact", "c_horizon", "p_com")]
        m = orc.build(p, *args)
        msk = np.repeat(args[2].ravel() > 0, 3)
        assert np.linalg.cond(m["H"][np.ix_(msk, msk)]) > 1e11
        xs, _, cert = orc.qp_exact(p, m["H"], m["g"], args[2])
        assert cert["rc"] == 0
        for lin in ("dense", "riccati"):
            u, it, gap = solve_ipm(Problem(preset, *args), linear=lin)
            assert gap <= 1e-10 and it <= 25, (lin, it, gap)
            assert np.abs(u.ravel() - xs).max() <= GRF_TOL
        u, it, gap = solve_ipm(Problem(preset, *args), linear="stage")          # what the kernel runs: Theta_k = Pi (I - Gamma)
        assert gap <= 1e-10 and it <= 25, (it, gap)
        assert np.abs(u.ravel() - xs).max() <= GRF_TOL
        u, it, gap = solve_ipm(Problem(preset, *args), linear="stage_s")        # S_k (Gamma p2) as a product
        lost += int(it == 40 or np.abs(u.ravel() - xs).max() > GRF_TOL)
    assert lost >= 1            # stalls, or ends off the minimiser (which of the two depends on the rounding)


def test_warm_start_model_halves_the_iterations_and_the_safeguard_bounds_a_bad_start():
    """The kernel's warm start (srbd_solve_warp.cuh: start_point + cold restart), in the numpy model: on consecutive ticks of
    the configs[0] replay the previous minimiser shifted by one stage, pulled 0.01 m g inside every constraint row, with
    lambda = 0.1 / s, needs about half the iterations of the cold start and reaches the same minimiser; a start that has
    nothing to do with the problem stalls, is detected after 5 iterations (gap not halved) and costs at most those 5."""
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "config1_replay.npz"))
    pr = presets.get_preset("cpp_node")
    W = pr["mass"] * 9.80665
    cold, warm = [], []
    for k in range(40, 70):
        P = im.Problem(pr, z["x0"][k], z["x_ref"][k], z["contact"][k], z["c_horizon"][k], z["p_com"][k])
        u0, it0, _ = im.solve_ipm(P, linear="dense")
        u1, it1, _ = im.solve_ipm(P, linear="dense", warm=dict(u=z["u_star"][k - 1].reshape(10, 4, 3), margin=0.01 * W, mu0=0.1))
        assert np.abs(u0.reshape(10, 12) - z["u_star"][k]).max() <= 1e-3 and np.abs(u1.reshape(10, 12) - z["u_star"][k]).max() <= 1e-3
        cold.append(it0); warm.append(it1)
        s0 = im.warm_point(P, z["u_star"][k - 1].reshape(10, 4, 3), 0.01 * W)[1]
        assert s0[P.stance].min() >= 0.01 * W - 1e-12              # every slack of the warm point is at least the margin
    assert np.mean(warm) <= 0.65 * np.mean(cold), (np.mean(cold), np.mean(warm))
    rng = np.random.default_rng(0)
    P = im.Problem(pr, z["x0"][50], z["x_ref"][50], z["contact"][50], z["c_horizon"][50], z["p_com"][50])
    _, it_cold, _ = im.solve_ipm(P, linear="dense")
    u2, it2, _ = im.solve_ipm(P, linear="dense", warm=dict(u=rng.uniform(-300, 900, (10, 4, 3)), margin=0.01 * W, mu0=0.1))
    assert np.abs(u2.reshape(10, 12) - z["u_star"][50]).max() <= 1e-3 and it2 <= it_cold + 5
