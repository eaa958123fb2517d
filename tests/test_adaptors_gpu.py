"""GPU parity of the two callers either side of the hot path through the C-ABI, against the
CPU restatement oracle/node_oracle.py.  Integer/index work and +,-,*,/ arithmetic: bit-exact; sin/cos: 2 ulp."""
import numpy as np
import pytest
import torch

from g1_locomotion_b200 import presets

pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float64).cuda()


@pytest.mark.parametrize("N", [10, 8])
def test_walk_inputs_match_node_restatement(N):
    from g1_locomotion_b200 import BatchedMPC
    from oracle import node_oracle as no
    preset = presets.get_preset("cpp_node", horizon=N)
    s = BatchedMPC(preset)
    rng = np.random.default_rng(3)
    B = 257
    dt, g = preset["dt"], preset["gravity"] if "gravity" in preset else -9.80665
    x0 = rng.normal(0, 0.3, (B, 13))
    cur = rng.normal(0, 0.2, (B, 12))
    phase = rng.integers(0, 45, B)
    sim_time = (phase + rng.uniform(0.05, 0.95, B)) * dt
    x_prev = rng.normal(0, 0.5, (B, N, 13))
    x_ref0 = rng.normal(0, 0.1, (B, N, 13))                      # non-zero: the untouched columns must survive
    x_ref, contact, c_hor, p_com, landing = s.walk_inputs(_dev(x0), _dev(cur), _dev(sim_time), _dev(x_prev), x_ref=_dev(x_ref0),
                                                          want_landing=True)
    torch.cuda.synchronize()
    for b in range(B):
        w = no.node_update_inputs(x0[b], cur[b], sim_time[b], x_prev[b], x_ref0[b], dt, N, g)
        for got, want, name in zip((x_ref, contact, c_hor, p_com, landing), w, ("x_ref", "contact", "c_horizon", "p_com", "landing")):
            assert np.array_equal(got[b].cpu().numpy(), want), (b, name, int(phase[b]))


def test_walk_inputs_feed_the_solver():
    """generator -> solve on the device without touching the host: every problem solves, swing forces are zero."""
    from g1_locomotion_b200 import BatchedMPC
    preset = presets.get_preset("cpp_node")
    s = BatchedMPC(preset)
    B, N = 512, 10
    gen = torch.Generator(device="cuda"); gen.manual_seed(5)
    x0 = torch.zeros(B, 13, device="cuda", dtype=torch.float64)
    x0[:, 3] = 0.05 + 0.35 * 0.05 * torch.arange(B, device="cuda")
    x0[:, 4], x0[:, 5], x0[:, 9], x0[:, 12] = 7.44339342e-05, 5.97983255e-01, 0.35, float(np.float32(-9.80665))
    x0[:, 0:12] += 0.003 * torch.randn(B, 12, generator=gen, device="cuda", dtype=torch.float64)
    cur = torch.zeros(B, 4, 3, device="cuda", dtype=torch.float64)
    cur[:, :, 0] = x0[:, 3:4] + torch.tensor([-0.0293, 0.1406, -0.0293, 0.1406], device="cuda", dtype=torch.float64)
    cur[:, :, 1] = torch.tensor([0.118508, 0.118508, -0.118508, -0.118508], device="cuda", dtype=torch.float64)
    sim_time = (torch.arange(B, device="cuda", dtype=torch.float64) + 0.5) * 0.05
    x_prev = torch.zeros(B, N, 13, device="cuda", dtype=torch.float64)
    x_prev[:, :, 3] = x0[:, 3:4] + 0.35 * 0.05 * torch.arange(N, device="cuda")
    x_prev[:, :, 4], x_prev[:, :, 9] = 7.44339342e-05, 0.35
    x_ref, contact, c_hor, p_com = s.walk_inputs(x0, cur.reshape(B, 12).contiguous(), sim_time, x_prev)
    u, x, info = s.update_batch(x0, x_ref, contact, c_hor, p_com, x_current=x0)
    assert int((info["status"] != 0).sum()) == 0
    assert torch.all(u.view(B, N, 4, 3)[contact == 0] == 0)
    f = u.view(B, N, 4, 3)
    assert float((f[..., 0].abs() - preset["mu"] * f[..., 2]).max()) <= 1e-9 and float(f[..., 2].min()) >= 0.0
    assert torch.all(f[..., 2][contact > 0] >= preset["fz_min"] - 1e-9)


def test_wbid_refs_match_restatement():
    from g1_locomotion_b200 import BatchedMPC
    from oracle import node_oracle as no
    preset = presets.get_preset("cpp_node")
    s = BatchedMPC(preset)
    rng = np.random.default_rng(4)
    B, N = 130, 10
    u = rng.normal(0, 100, (B, N, 12)); x = rng.normal(0, 0.4, (B, N, 13)); feet = rng.normal(0, 0.3, (B, 4, 3))
    mass = 35.1127
    refs = s.wbid_refs(_dev(u), _dev(x), _dev(feet), mass).cpu().numpy()
    for b in range(B):
        want = no.wbid_refs_flat(x[b, 1], u[b, 0], feet[b], mass)
        assert np.abs(refs[b, :9] - want[:9]).max() <= 4e-16, b        # rotation: sin/cos within an ulp or two
        assert np.array_equal(refs[b, 9:], want[9:]), (b, np.abs(refs[b, 9:] - want[9:]).max())


def test_wbid_refs_match_the_reference_set_reference_golden():
    """g1mpc_wbid_refs_batch against what the reference's OWN WholeBodyID.setReference (g1_mujoco_sim/src/wbid.py:231-297) handed its
    OpenSoT tasks on the same inputs (tests/golden/wbid_refs.npz, written by `python -m oracle.gen_golden callers` from the unmodified
    file with recording stand-ins for the tasks): bit for bit, except the rotation (device sin/cos, <= 4e-16)."""
    import os
    from conftest import ROOT
    from g1_locomotion_b200 import BatchedMPC
    z = np.load(os.path.join(ROOT, "tests", "golden", "wbid_refs.npz"))
    B, N = len(z["refs"]), 10
    s = BatchedMPC(presets.get_preset("cpp_node"))
    rng = np.random.default_rng(9)
    u = rng.normal(0, 100, (B, N, 12)); x = rng.normal(0, 0.4, (B, N, 13))
    u[:, 0], x[:, 1] = z["u_opt0"], z["x_opt1"]                    # the adaptor reads row 0 of u_opt and row 1 of x_opt (ros_run_simulation.py:409)
    refs = s.wbid_refs(_dev(u), _dev(x), _dev(z["feet"]), float(z["mass"])).cpu().numpy()
    assert np.abs(refs[:, :9] - z["refs"][:, :9]).max() <= 4e-16
    assert np.array_equal(refs[:, 9:], z["refs"][:, 9:]), np.abs(refs[:, 9:] - z["refs"][:, 9:]).max()


def test_adaptors_reject_unsupported_shapes():
    from g1_locomotion_b200 import BatchedMPC
    s = BatchedMPC(presets.get_preset("cpp_node", num_contacts=8))
    with pytest.raises(RuntimeError):
        s.wbid_refs(torch.zeros(1, 10, 24, device="cuda", dtype=torch.float64), torch.zeros(1, 10, 13, device="cuda", dtype=torch.float64),
                    torch.zeros(1, 4, 3, device="cuda", dtype=torch.float64), 35.0)


def test_closed_loop_matches_cpu_composition(orc):
    """12 ticks of 6 walkers on the device == the same loop composed on the CPU from the node restatement, the
    reference build, the exact QP minimiser and the reference rollout.  Forces within 2e-3 N, states within 1e-6."""
    from g1_locomotion_b200 import BatchedMPC
    from oracle import node_oracle as no
    preset = presets.get_preset("cpp_node")
    s = BatchedMPC(preset)
    B, T, dt = 6, 12, preset["dt"]
    rng = np.random.default_rng(8)
    x0 = np.zeros((B, 13))
    x0[:, 3] = 0.05 + 0.1 * np.arange(B); x0[:, 4] = 7.44339342e-05; x0[:, 5] = 5.97983255e-01; x0[:, 9] = 0.35
    x0[:, 12] = np.float32(-9.80665)
    x0[:, :12] += rng.normal(0, 0.003, (B, 12))
    cur = np.zeros((B, 4, 3))
    cur[:, :, 0] = x0[:, 3:4] + np.array([-0.0293, 0.1406, -0.0293, 0.1406])
    cur[:, :, 1] = np.array([0.118508, 0.118508, -0.118508, -0.118508])
    t0 = (np.array([0, 3, 7, 12, 18, 26]) + 0.5) * dt
    out = s.closed_loop(_dev(x0), _dev(cur.reshape(B, 12)), _dev(t0), T)
    torch.cuda.synchronize()
    assert int(out["status_any"].abs().sum()) == 0
    tx, tu = out["traj_x"].cpu().numpy(), out["traj_u"].cpu().numpy()
    for b in range(B):
        wx, wu = no.closed_loop(orc, preset, x0[b], cur[b].ravel(), t0[b], T)
        assert np.abs(tx[b] - wx).max() <= 1e-6, (b, np.abs(tx[b] - wx).max())
        assert np.abs(tu[b] - wu).max() <= 2e-3, (b, np.abs(tu[b] - wu).max())
    assert np.array_equal(tx[:, 0], x0)


def test_closed_loop_of_a_large_fleet_equals_the_small_one():
    """1,200 walkers use the two-problems-per-warp layout, 6 walkers the one-problem layout: the trajectories of the same
    walker must be bit-identical in both runs (the layouts return the same bits, every walker is independent)."""
    from g1_locomotion_b200 import BatchedMPC
    preset = presets.get_preset("cpp_node")
    s = BatchedMPC(preset)
    B, T, dt = 1200, 5, preset["dt"]
    rng = np.random.default_rng(9)
    x0 = np.zeros((B, 13))
    # walkers up to 12 m from the origin: with x_opt = 0 at the first tick the node plants the first swing foot near the
    # WORLD origin, so the far ones get lever arms of metres and cond(H) up to 4e11 (tests/golden/ipm_illconditioned_problems.npz)
    x0[:, 3] = 0.05 + 0.01 * np.arange(B); x0[:, 4] = 7.44339342e-05; x0[:, 5] = 5.97983255e-01; x0[:, 9] = 0.35
    x0[:, 12] = np.float32(-9.80665)
    x0[:, :12] += rng.normal(0, 0.003, (B, 12))
    cur = np.zeros((B, 4, 3))
    cur[:, :, 0] = x0[:, 3:4] + np.array([-0.0293, 0.1406, -0.0293, 0.1406])
    cur[:, :, 1] = np.array([0.118508, 0.118508, -0.118508, -0.118508])
    t0 = (rng.integers(0, 30, B) + 0.5) * dt
    assert s.kernel_info(B)["problems_per_warp"] == 2 and s.kernel_info(6)["problems_per_warp"] == -1      # (-1: latency mode, one problem per two warps)
    big = s.closed_loop(_dev(x0), _dev(cur.reshape(B, 12)), _dev(t0), T)
    small = s.closed_loop(_dev(x0[:6]), _dev(cur[:6].reshape(6, 12)), _dev(t0[:6]), T)
    torch.cuda.synchronize()
    # (the far walkers' problems have cond(H) up to 4e11: a handful end with status 4 = gap converged, dual residual above
    # dual_tol -- reported, not hidden; nothing worse than that may occur)
    assert set(torch.unique(big["status_any"]).tolist()) <= {0, 4} and int((big["status_any"] != 0).sum()) <= 8
    assert torch.equal(big["traj_x"][:6], small["traj_x"]) and torch.equal(big["traj_u"][:6], small["traj_u"])
