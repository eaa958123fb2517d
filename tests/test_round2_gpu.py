"""GPU tests of the round-2 rows: rollout-only entry, gait planner (8f-1 Python half), warm start, the run-time (N, nc)
kernel, the drop-in class's attribute semantics (8a a18), the main.py-style closed loop (8f-3), the C++ adapter, and the
whole-config KKT net.  Everything goes through the C-ABI; the checker is the CPU oracle / the golden files written from
the imported reference."""
import json
import os
import subprocess

import numpy as np
import pytest
import torch

from conftest import ROOT, golden_cases, load_golden, rel_err
from g1_locomotion_b200 import presets, inputs, BatchedMPC, MPC
from oracle import node_oracle

pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float64).cuda()


def _replay():
    z = np.load(os.path.join(ROOT, "tests", "golden", "config1_replay.npz"))
    return {k: z[k] for k in z.files}


# --------------------------------------------------------------------------------------------------------------------
# 8f-1 Python half: plan_gait + mapping, bit-exact against the reference class's output
# --------------------------------------------------------------------------------------------------------------------
def test_plan_gait_kernel_matches_reference_golden_bit_for_bit():
    z = np.load(os.path.join(ROOT, "tests", "golden", "gait_plan.npz"))
    total = 0
    for N, dt in ((10, 0.05), (15, 0.04)):
        sel = np.where((z["N"] == N))[0]
        assert np.all(z["dt"][sel] == dt)
        s = BatchedMPC(presets.get_preset("cpp_node", horizon=N, dt=dt), device=0)
        contact, c_hor = s.plan_gait(_dev(z["planner"][sel]), _dev(z["t"][sel]), 0.08)
        assert np.array_equal(contact.cpu().numpy(), z["contact"][sel][:, :N])
        assert np.array_equal(c_hor.cpu().numpy(), z["c_horizon"][sel][:, :N])
        total += len(sel)
    assert total == len(z["t"]) >= 200
    # and against the restatement on a larger random set (planner parameters and query times)
    rng = np.random.default_rng(7)
    B = 2000
    pl = np.stack([rng.uniform(0, 1.5, B), rng.uniform(0.05, 1.0, B), rng.uniform(0.01, 0.4, B), rng.uniform(0.05, 0.3, B),
                   rng.integers(0, 60, B).astype(float)], 1)
    t = rng.uniform(-0.2, 30.0, B)
    s = BatchedMPC(presets.get_preset("cpp_node"), device=0)
    contact, c_hor = [a.cpu().numpy() for a in s.plan_gait(_dev(pl), _dev(t), 0.08)]
    for b in range(0, B, 7):
        ct, ch = node_oracle.map_gait_phases(node_oracle.plan_gait(*pl[b]), float(t[b]), 10, 0.05, 0.08)
        assert np.array_equal(contact[b], ct) and np.array_equal(c_hor[b], ch), b
    with pytest.raises(RuntimeError):
        BatchedMPC(presets.get_preset("cpp_node", num_contacts=8), device=0).plan_gait(_dev(pl[:1]), _dev(t[:1]))


# --------------------------------------------------------------------------------------------------------------------
# rollout-only entry (mpc.cpp:427-446) against the reference's own compute_rollout outputs
# --------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", golden_cases())
def test_rollout_batch_matches_reference_golden(name):
    gd = load_golden(name)
    s = BatchedMPC(gd["preset"], device=0)
    N = gd["preset"]["horizon"]
    xr, ch, pc, u, x0 = [_dev(gd[k][None]) for k in ("x_ref", "c_horizon", "p_com", "u_rollout", "x0")]
    full = s.rollout(xr, ch, pc, u, x_current=x0)
    one = s.rollout(xr, ch, pc, u, x_current=x0, one_rollout=True)
    assert rel_err(full[0].cpu().numpy(), gd["x_opt_full"]) <= 1e-9
    assert rel_err(one[0].cpu().numpy(), gd["x_opt_one"]) <= 1e-9 and float(one[0, 2:].abs().max()) == 0.0
    keep = s.rollout(xr, ch, pc, u, x_current=None, one_rollout=True, x_opt=torch.full((1, N, 13), 7.0, dtype=torch.float64, device="cuda"))
    assert float((keep[0, 0] - 7.0).abs().max()) == 0.0 and float((keep[0, 2:] - 7.0).abs().max()) == 0.0     # rows the reference leaves alone
    # the solve kernel's fused rollout and the rollout-only kernel agree on the solver's own u_opt
    ins = [_dev(gd[k][None]) for k in ("x0", "x_ref", "contact", "c_horizon", "p_com")]
    uo, xo, _ = s.update_batch(*ins, x_current=ins[0])
    assert float((s.rollout(xr, ch, pc, uo, x_current=x0) - xo).abs().max()) <= 1e-9 * max(1.0, float(xo.abs().max()))


def test_rollout_of_the_reference_update_goldens():
    import glob
    for f in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "ref_update_*.npz"))):
        z = np.load(f)
        pr = json.loads(str(z["preset"]))
        s = BatchedMPC(pr, device=0)
        xr, ch, pc, x0 = [_dev(z[k][None]) for k in ("x_ref", "c_horizon", "p_com", "x0")]
        got = s.rollout(xr, ch, pc, _dev(z["u_opt_full"][None]), x_current=x0)[0].cpu().numpy()
        assert np.abs(got - z["x_opt_full"]).max() <= 1e-9 * max(1.0, np.abs(got).max())
        got = s.rollout(xr, ch, pc, _dev(z["u_opt_one"][None]), x_current=x0, one_rollout=True)[0].cpu().numpy()
        assert np.abs(got - z["x_opt_one"]).max() <= 1e-9 * max(1.0, np.abs(got).max())


# --------------------------------------------------------------------------------------------------------------------
# warm start
# --------------------------------------------------------------------------------------------------------------------
def test_warm_start_reaches_the_same_minimiser_in_fewer_iterations():
    """configs[0] replay (closed-loop generator of SURVEY 8d): every tick started from the previous tick's solution shifted
    by one stage.  Same unique minimiser (1e-3 N against the oracle's x*), about half the iterations."""
    z = _replay()
    s = BatchedMPC(presets.get_preset("cpp_node"), device=0)
    ins = [_dev(z[k]) for k in ("x0", "x_ref", "contact", "c_horizon", "p_com")]
    ustar = _dev(z["u_star"])
    uc, xc, ic = s.update_batch(*ins, x_current=ins[0])
    prev = torch.cat([uc[:1], uc[:-1]]).contiguous()
    uw, xw, iw = s.update_batch(*ins, x_current=ins[0], u_warm=prev, warm_shift=1)
    assert int((ic["status"] != 0).sum()) == 0 and int((iw["status"] != 0).sum()) == 0
    assert float((uc - ustar).abs().max()) <= 1e-3 and float((uw - ustar).abs().max()) <= 1e-3
    assert float((xw - xc).abs().max()) <= 1e-7
    cold, warm = float(ic["iters"][1:].double().mean()), float(iw["iters"][1:].double().mean())
    assert warm <= 0.65 * cold, (cold, warm)
    assert int(iw["iters"].max()) <= int(ic["iters"].max()) + 5          # the safeguard bounds a bad start by 5 extra iterations
    # in place (u_warm aliases u_opt), and on both lane layouts
    for layout in (1, 2):
        s2 = BatchedMPC(presets.get_preset("cpp_node"), device=0, layout=layout)
        buf = prev.clone()
        u2, _, i2 = s2.update_batch(*ins, x_current=ins[0], u_warm=buf, warm_shift=1, u_opt=buf)
        assert torch.equal(u2, uw) and torch.equal(i2["iters"], iw["iters"]), layout
    # warm_shift = 0 from the solution itself: a re-solve converges at once
    u0, _, i0 = s.update_batch(*ins, x_current=ins[0], u_warm=uc, warm_shift=0)
    assert float((u0 - ustar).abs().max()) <= 1e-3 and float(i0["iters"].double().mean()) <= 0.6 * cold


def test_warm_start_from_garbage_is_contained():
    """A warm start that has nothing to do with the problem (random forces, NaN, another walker's solution, zeros) must not
    cost the answer: the start point is clamped into the interior, a stalled start restarts cold after 5 iterations."""
    z = _replay()
    s = BatchedMPC(presets.get_preset("cpp_node"), device=0)
    ins = [_dev(z[k]) for k in ("x0", "x_ref", "contact", "c_horizon", "p_com")]
    ustar = _dev(z["u_star"])
    uc, _, ic = s.update_batch(*ins, x_current=ins[0])
    gen = torch.Generator(device="cuda"); gen.manual_seed(1)
    B = uc.shape[0]
    junk = [2000.0 * (torch.rand(uc.shape, generator=gen, device="cuda", dtype=torch.float64) - 0.3),
            torch.full_like(uc, float("nan")), torch.zeros_like(uc), uc.flip(0).contiguous(), torch.full_like(uc, 1e30)]
    for w in junk:
        u, _, info = s.update_batch(*ins, x_current=ins[0], u_warm=w, warm_shift=1)
        assert int((info["status"] != 0).sum()) == 0
        assert float((u - ustar).abs().max()) <= 1e-3
        assert int(info["iters"].max()) <= int(ic["iters"].max()) + 6


def test_closed_loop_warm_start_halves_the_iterations():
    """the fleet loop with and without warm start across ticks -- same trajectories, fewer iterations."""
    s = BatchedMPC(presets.get_preset("cpp_node"), device=0)
    B, T = 256, 40
    def fleet():
        x0 = torch.zeros(B, 13, dtype=torch.float64, device="cuda")
        x0[:, 3], x0[:, 4], x0[:, 5], x0[:, 12] = 5.26790425e-02, 7.44339342e-05, 5.97983255e-01, float(np.float32(-9.80665))
        gen = torch.Generator(device="cuda"); gen.manual_seed(3)
        x0[:, 0:12] += 0.002 * torch.randn(B, 12, generator=gen, device="cuda", dtype=torch.float64)
        cur = torch.zeros(B, 4, 3, dtype=torch.float64, device="cuda")
        cur[:, :, 0] = x0[:, 3:4] + torch.tensor([-0.0293, 0.1406, -0.0293, 0.1406], dtype=torch.float64, device="cuda")
        cur[:, :, 1] = torch.tensor([0.118508, 0.118508, -0.118508, -0.118508], dtype=torch.float64, device="cuda")
        t = torch.full((B,), 1e-9, dtype=torch.float64, device="cuda")
        return x0, cur.reshape(B, 12).contiguous(), t
    cold = s.closed_loop(*fleet(), T)
    warm = s.closed_loop(*fleet(), T, warm_start=True)
    torch.cuda.synchronize()
    assert int(cold["status_any"].abs().sum()) == 0 and int(warm["status_any"].abs().sum()) == 0
    assert float((cold["traj_x"] - warm["traj_x"]).abs().max()) <= 1e-6 and float((cold["traj_u"] - warm["traj_u"]).abs().max()) <= 2e-3
    ic, iw = float(cold["iters_sum"].double().mean()) / T, float(warm["iters_sum"].double().mean()) / T
    assert iw <= 0.7 * ic, (ic, iw)


# --------------------------------------------------------------------------------------------------------------------
# latency mode: two warps per problem for small batches
# --------------------------------------------------------------------------------------------------------------------
def test_latency_mode_kernel_gives_the_same_bits_as_the_one_warp_kernel():
    """Up to two problems per SM the library runs ONE problem per two-warp CTA: the partner warp does the predictor's pre-pass
    and backward sweep one stage behind the factor loop (srbd_mpc_solve_lat_kernel).  Same functions, same operations: every
    output is bit-identical to the one-warp kernel (layout 1) -- cold, warm (cold restarts included), one-step rollout, no
    x_current, problems that never reach a factorisation (all contacts in swing; NaN input), B = 1 and launches beyond 2 per SM."""
    preset = presets.get_preset("cpp_node")
    one, lat, auto = BatchedMPC(preset, device=0, layout=1), BatchedMPC(preset, device=0, layout=3), BatchedMPC(preset, device=0)
    assert auto.kernel_info(1)["problems_per_warp"] == lat.kernel_info(700)["problems_per_warp"] == -1      # one problem per TWO warps
    assert auto.kernel_info(297)["problems_per_warp"] == 1 and one.kernel_info(1)["problems_per_warp"] == 1   # above 2 per SM: one warp
    lat2 = BatchedMPC(preset, device=0, layout=4)       # the same with two problems per main warp (not chosen automatically)
    z = _replay()
    rep = [_dev(z[k]) for k in ("x0", "x_ref", "contact", "c_horizon", "p_com")]
    rnd = list(inputs.random_batch(preset, 700, seed=41, device="cuda"))
    odd = [t.clone() for t in rnd]
    odd[2][3] = 0.0                                     # all swing: no factorisation at all
    odd[0][5, 4] = float("nan")                         # contained failure
    def same(a, b):
        ok = torch.equal(a[0].nan_to_num(nan=-1.0), b[0].nan_to_num(nan=-1.0)) and torch.equal(a[1].nan_to_num(nan=-1.0), b[1].nan_to_num(nan=-1.0))
        return ok and all(torch.equal(a[2][k].nan_to_num(nan=-1.0) if a[2][k].is_floating_point() else a[2][k],
                                      b[2][k].nan_to_num(nan=-1.0) if b[2][k].is_floating_point() else b[2][k]) for k in ("status", "iters", "resid", "y"))
    for ins in (rep, rnd, odd, [t[:1].contiguous() for t in rnd]):
        for kw in (dict(x_current=ins[0]), dict(x_current=ins[0], one_rollout=True), dict()):
            xp = torch.full((ins[0].shape[0], 10, 13), 2.0, dtype=torch.float64, device="cuda")
            a = one.update_batch(*ins, x_opt=xp.clone(), want_multipliers=True, **kw)
            for _ in range(2):
                assert same(a, lat.update_batch(*ins, x_opt=xp.clone(), want_multipliers=True, **kw))
            assert same(a, lat2.update_batch(*ins, x_opt=xp.clone(), want_multipliers=True, **kw))
    assert set(a[2]["status"].tolist()) == {0}
    st_odd = one.update_batch(*odd, x_current=odd[0])[2]["status"]
    assert int(st_odd[5]) == 3 and int(st_odd[3]) == 0
    # warm start (good starts, and junk that triggers the cold restart inside the kernel)
    u = one.update_batch(*rep, x_current=rep[0])[0]
    for w in (torch.cat([u[:1], u[:-1]]).contiguous(), torch.full_like(u, 700.0)):
        a = one.update_batch(*rep, x_current=rep[0], u_warm=w, want_multipliers=True)
        assert same(a, lat.update_batch(*rep, x_current=rep[0], u_warm=w, want_multipliers=True))
    # the Python default shape (N = 15: every one of the sixteen hardware barriers is in use)
    p15 = presets.get_preset("python_default")
    r15 = list(inputs.random_batch(p15, 200, seed=4, device="cuda"))
    a15, l15 = BatchedMPC(p15, device=0, layout=1), BatchedMPC(p15, device=0)
    assert l15.kernel_info(200)["problems_per_warp"] == -1
    assert same(a15.update_batch(*r15, x_current=r15[0], want_multipliers=True), l15.update_batch(*r15, x_current=r15[0], want_multipliers=True))
    # the host API with one problem (zero-copy path) lands on the latency kernel too
    host = [t[7:8].cpu().numpy() for t in rep]
    uh, xh, ih = auto.update_host(*host, x_current=host[0])
    assert np.array_equal(uh[0], u[7].cpu().numpy())


# --------------------------------------------------------------------------------------------------------------------
# run-time shapes: any (N, nc) that fits shared memory
# --------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("N,nc,pname", [(12, 4, "cpp_node"), (30, 4, "cpp_node"), (7, 8, "cpp_node"), (3, 4, "python_default"),
                                        (25, 8, "python_default"), (48, 4, "cpp_node"), (9, 2, "cpp_node")])
def test_runtime_shape_kernel_kkt(orc, N, nc, pname):
    """Shapes with no compile-time instantiation run the same kernel source with the layout in registers (g1mpc_create never
    answers UNSUPPORTED_SHAPE for a shape that fits shared memory).  Certified on reference-built matrices like config 4."""
    preset = presets.get_preset(pname, horizon=N, num_contacts=nc)
    s = BatchedMPC(preset, device=0)
    B = 24
    if nc in (4, 8):
        ins = list(inputs.random_batch(preset, B, seed=40 + N, device="cuda"))
    else:       # two line contacts (one per foot): take the heel points of the nc = 4 generator
        p4 = presets.get_preset(pname, horizon=N, num_contacts=4)
        x0, xr, ct, ch, pc = inputs.random_batch(p4, B, seed=40 + N, device="cuda")
        ins = [x0, xr, ct[:, :, ::2].contiguous(), ch.reshape(B, N, 4, 3)[:, :, ::2].reshape(B, N, 6).contiguous(), pc]
    u, x, info = s.update_batch(*ins, x_current=ins[0], want_multipliers=True)
    assert int((info["status"] != 0).sum()) == 0, info["status"]
    p = orc.params_from_dict(preset)
    host = [t.cpu().numpy() for t in ins]
    uh, yh, xh = u.cpu().numpy(), info["y"].cpu().numpy(), x.cpu().numpy()
    for b in range(0, B, 3):
        m = orc.build(p, *[a[b] for a in host])
        k = orc.kkt_residuals(m["H"], m["g"], m["Acon"], m["l"], m["u"], uh[b].ravel(), yh[b])
        gs = 1.0 + np.abs(m["g"]).max()
        assert k["stationarity"] <= 5e-6 * gs and k["primal"] <= 1e-9 and k["complementarity"] <= 1e-6 * gs, (b, k)
        xo = orc.rollout(p, m["A"], m["Bd"], uh[b], host[0][b], False)
        assert np.abs(xo - xh[b]).max() <= 1e-8 * max(1.0, np.abs(xo).max())      # (closed form against a recurrence of N steps)
    if N * nc <= 60:        # small enough for the exact active-set minimiser
        for b in range(0, B, 6):
            m = orc.build(p, *[a[b] for a in host])
            xs, _, cert = orc.qp_exact(p, m["H"], m["g"], host[2][b])
            assert cert["rc"] == 0 and np.abs(uh[b].ravel() - xs).max() <= 1e-3


def test_shape_beyond_shared_memory_runs_from_the_global_workspace(orc):
    """N = 60, nc = 8 needs 240 KB per problem: refused with workspace = 1 (shared memory), run with the default (the
    global-memory workspace behind L1/L2, persistent grid).  Horizons this long sit on the accuracy floor of FP64
    (profiles/r2_a_dual_residual_by_shape.txt): what is checked is that the STATUS is honest -- a problem reported solved is
    stationary to 1e-5 (1 + |g|inf) on the reference-built QP, anything worse comes back flagged (4 or 1) -- and that every
    returned point is primal feasible."""
    preset = presets.get_preset("cpp_node", horizon=60, num_contacts=8)
    with pytest.raises(RuntimeError, match="exceeds shared memory"):
        BatchedMPC(preset, device=0, workspace=1)
    with pytest.raises(RuntimeError):
        BatchedMPC(presets.get_preset("cpp_node", horizon=1), device=0)
    s = BatchedMPC(preset, device=0)
    # 24 problems fit one wave of the hybrid workspace (stage factors in shared memory, per-contact state in global memory);
    # a batch of several waves runs from the global-memory workspace
    assert "hybrid" in s.kernel_info(24)["mode"] and 0 < s.kernel_info(24)["smem_bytes"] < 240096
    assert s.kernel_info(4096)["smem_bytes"] == 0
    ins = list(inputs.random_batch(preset, 24, seed=3, device="cuda"))
    u, x, info = s.update_batch(*ins, x_current=ins[0], want_multipliers=True)
    st = info["status"].tolist()
    assert set(st) <= {0, 1, 4} and st.count(0) >= 6, st
    assert float(info["resid"][:, 2].max()) == 0.0 and bool(torch.isfinite(u).all())
    p = orc.params_from_dict(preset)
    for b in [i for i, v in enumerate(st) if v == 0][:4]:
        m = orc.build(p, *[t[b].cpu().numpy() for t in ins])
        k = orc.kkt_residuals(m["H"], m["g"], m["Acon"], m["l"], m["u"], u[b].cpu().numpy().ravel(), info["y"][b].cpu().numpy())
        assert k["primal"] <= 1e-9 and k["stationarity"] <= 1e-5 * (1 + np.abs(m["g"]).max()), (b, k)


def test_no_write_lands_outside_the_output_arrays():
    """compute-sanitizer is closed on this GPU pool, so out-of-bounds writes are looked for with guard bands: every output
    array of the solve, rollout, emit, gait-plan and closed-loop entry points is a window into a larger buffer filled with a
    sentinel; after the calls (odd batch sizes, every lane layout, shared / global / hybrid workspace) the bands are untouched."""
    SENT = -7.25e301
    def window(*shape, dtype=torch.float64):
        n = int(np.prod(shape))
        buf = torch.full((n + 512,), SENT if dtype == torch.float64 else -77, dtype=dtype, device="cuda")
        return buf, buf[256:256 + n].view(*shape)
    def intact(buf, n):
        v = SENT if buf.dtype == torch.float64 else -77
        return bool((buf[:256] == v).all()) and bool((buf[256 + n:] == v).all())
    for preset, kw, B in ((presets.get_preset("cpp_node"), dict(layout=1), 5), (presets.get_preset("cpp_node"), dict(layout=2), 7),
                          (presets.get_preset("cpp_node", horizon=13), dict(workspace=1), 3),
                          (presets.get_preset("cpp_node", horizon=21, num_contacts=8), dict(workspace=2), 301),
                          (presets.get_preset("cpp_node", horizon=21, num_contacts=8), dict(workspace=3), 1187),      # hybrid, > one wave
                          (presets.get_preset("python_default"), dict(), 1301),          # two per warp from the hybrid workspace (odd batch)
                          (presets.get_preset("cpp_node"), dict(layout=5), 2371)):       # refill: persistent grid, odd batch
        N, nu = preset["horizon"], 3 * preset["num_contacts"]
        s = BatchedMPC(preset, device=0, **kw)
        ins = list(inputs.random_batch(preset, B, seed=2, device="cuda"))
        bu, u = window(B, N, nu); bx, x = window(B, N, 13)
        x.zero_()
        uo, xo, info = s.update_batch(*ins, x_current=ins[0], u_opt=u, x_opt=x, want_multipliers=True)
        br, xr2 = window(B, N, 13); xr2.zero_()
        s.rollout(ins[1], ins[3], ins[4], u, x_current=ins[0], x_opt=xr2)
        torch.cuda.synchronize()
        assert intact(bu, B * N * nu) and intact(bx, B * N * 13) and intact(br, B * N * 13), (kw, B)
        assert bool(torch.isfinite(u).all()) and float((xr2 - x).abs().max()) <= 1e-8 * float(x.abs().max())


@pytest.mark.parametrize("N,nc,B", [(20, 8, 2048), (40, 4, 1500), (30, 4, 777)])
def test_global_workspace_gives_the_same_bits_as_shared_memory(N, nc, B):
    """The long-horizon kernel variant keeps a problem's working set in global memory (one slot per resident warp, persistent
    grid, eight warps per SM instead of one or two): same code, same arithmetic, so every output must be bit-identical to the
    shared-memory variant -- including a launch larger than the grid (each warp walks several problems)."""
    preset = presets.get_preset("cpp_node", horizon=N, num_contacts=nc)
    ins = list(inputs.random_batch(preset, B, seed=29, device="cuda"))
    a = BatchedMPC(preset, device=0, workspace=1).update_batch(*ins, x_current=ins[0], want_multipliers=True)
    sg = BatchedMPC(preset, device=0, workspace=2)
    assert sg.kernel_info(B)["smem_bytes"] == 0
    b = sg.update_batch(*ins, x_current=ins[0], want_multipliers=True)
    c = sg.update_batch(*ins, x_current=ins[0], want_multipliers=True)          # and again on the used workspace
    for r in (b, c):
        assert torch.equal(a[0], r[0]) and torch.equal(a[1], r[1])
        for k in ("status", "iters", "resid", "y"):
            assert torch.equal(a[2][k], r[2][k]), k
    # the host API's chunks (four streams) take turns on the one workspace
    host = [t.cpu().numpy() for t in ins]
    uh, xh, ih = sg.update_host(*host, x_current=host[0])
    assert np.array_equal(uh, a[0].cpu().numpy()) and np.array_equal(ih["iters"], a[2]["iters"].cpu().numpy())
    # the hybrid workspace (stance list, stage factors and scratch in shared memory, per-contact state in a global-memory slot per
    # resident warp): same code again, same bits, forced and as the automatic choice picks it
    sh = BatchedMPC(preset, device=0, workspace=3)
    ki = sh.kernel_info(B)
    assert "hybrid" in ki["mode"] and 0 < ki["smem_bytes"] < BatchedMPC(preset, device=0, workspace=1).kernel_info(B)["smem_bytes"]
    for r in (sh.update_batch(*ins, x_current=ins[0], want_multipliers=True), sh.update_batch(*ins, x_current=ins[0], want_multipliers=True)):
        assert torch.equal(a[0], r[0]) and torch.equal(a[1], r[1])
        for k in ("status", "iters", "resid", "y"):
            assert torch.equal(a[2][k], r[2][k]), k
    auto = BatchedMPC(preset, device=0)
    modes = set()
    for Bsub in (100, ki["ctas_per_sm"] * 148, B):
        sub = [t[:Bsub].contiguous() for t in ins]
        r = auto.update_batch(*sub, x_current=sub[0])
        modes.add(auto.kernel_info(Bsub)["mode"].split()[0])
        assert torch.equal(a[0][:Bsub], r[0]) and torch.equal(a[2]["iters"][:Bsub], r[2]["iters"]), Bsub
    assert {"one", "hybrid"} <= modes, modes          # shared memory for one wave of it, the hybrid for one wave of the hybrid


@pytest.mark.parametrize("pname,N,nc,B", [("python_default", 15, 4, 4001), ("cpp_node", 10, 8, 4001), ("cpp_node", 12, 8, 3333)])
def test_two_problems_per_warp_from_the_hybrid_workspace_give_the_same_bits(pname, N, nc, B):
    """Short horizons whose two-problem CTAs would leave schedulers empty if both working sets lay in shared memory (N = 15, or
    eight contacts per stage) run two problems per warp with only the stage factors on chip and the per-contact state in a
    global-memory slot per half-warp (persistent grid; above 4 problems per SM, whole-batch launches).  Same code, same bits as
    one problem per warp from shared memory -- odd batch (the last half-warp idles) and more than two waves of the grid."""
    preset = presets.get_preset(pname, horizon=N, num_contacts=nc)
    ins = list(inputs.random_batch(preset, B, seed=31, device="cuda"))
    one = BatchedMPC(preset, device=0, layout=1, workspace=1)
    a = one.update_batch(*ins, x_current=ins[0], want_multipliers=True)
    assert one.kernel_info(B)["mode"] == "one problem per warp"
    auto = BatchedMPC(preset, device=0)
    ki = auto.kernel_info(B)
    assert ki["mode"].startswith("hybrid workspace, two problems per warp") and ki["problems_per_warp"] == 2, ki
    assert B > 2 * 2 * 148 * ki["ctas_per_sm"]                               # more than two waves of the persistent grid
    assert "hybrid" not in auto.kernel_info(500)["mode"]
    forced = BatchedMPC(preset, device=0, workspace=3, layout=2)
    for s in (auto, forced, auto):
        r = s.update_batch(*ins, x_current=ins[0], want_multipliers=True)
        assert torch.equal(a[0], r[0]) and torch.equal(a[1], r[1])
        for k in ("status", "iters", "resid", "y"):
            assert torch.equal(a[2][k], r[2][k]), k
    # the host API cuts a batch into chunks on several streams: those keep the shared-memory kernel (no shared workspace to queue on)
    host = [t.cpu().numpy() for t in ins]
    uh, xh, ih = auto.update_host(*host, x_current=host[0])
    assert np.array_equal(uh, a[0].cpu().numpy()) and np.array_equal(ih["iters"], a[2]["iters"].cpu().numpy())
    # warm start (a previous solution shifted by one stage) through the same variants
    prev = torch.roll(a[0], 1, 0).contiguous()
    w1 = one.update_batch(*ins, x_current=ins[0], u_warm=prev, warm_shift=1)
    w2 = auto.update_batch(*ins, x_current=ins[0], u_warm=prev, warm_shift=1)
    assert torch.equal(w1[0], w2[0]) and torch.equal(w1[2]["iters"], w2[2]["iters"]) and int((w2[2]["status"] != 0).sum()) == 0


# --------------------------------------------------------------------------------------------------------------------
# 8a a18: the drop-in class keeps every intermediate current after update(), like the reference object
# --------------------------------------------------------------------------------------------------------------------
def _check_attributes(orc, mpc, preset, x0, xr, ct, ch, pc):
    p = orc.params_from_dict(preset)
    m = orc.build(p, x0, xr, ct, ch, pc)
    N, nc = preset["horizon"], preset["num_contacts"]
    assert rel_err(mpc.Hessian, m["H"]) <= 1e-9 and rel_err(mpc.gradient.ravel(), m["g"]) <= 1e-9 and mpc.gradient.shape == (3 * nc * N, 1)
    assert rel_err(mpc.Aqp, m["Aqp"]) <= 1e-9 and rel_err(mpc.Bqp, m["Bqp"]) <= 1e-9
    assert rel_err(mpc.A_discrete, m["A"]) <= 1e-9 and rel_err(mpc.B_discrete_hor, m["Bd"]) <= 1e-9
    assert rel_err(mpc.Ac.toarray(), m["Acon"]) == 0.0
    assert np.array_equal(mpc.lower_bounds_horizon, m["l"]) and np.array_equal(mpc.upper_bounds_horizon, m["u"])
    assert rel_err(mpc.A_continuous, (m["A"] - np.eye(13)) / preset["dt"]) <= 1e-9
    assert rel_err(mpc.B_continuous_hor, m["Bd"] / preset["dt"]) <= 1e-9
    assert rel_err(mpc.B_discrete, m["Bd"][-13:]) <= 1e-9 and rel_err(mpc.B_continuous, m["Bd"][-13:] / preset["dt"]) <= 1e-9
    psi = float(np.mean(xr[:, 2]))
    T = np.array([[np.cos(psi), np.sin(psi), 0], [-np.sin(psi), np.cos(psi), 0], [0, 0, 1.0]])      # mpc.py:124-128
    assert np.abs(mpc.rotation_z_T - T).max() <= 1e-12 and abs(mpc.psi - psi) <= 1e-15
    assert np.array_equal(mpc.r, ch[N - 1, 3 * (nc - 1):] - pc[N - 1])                               # mpc.py:176, last pass of the loop
    assert mpc.Q.shape == (13 * N, 13 * N) and mpc.R.shape == (3 * nc * N, 3 * nc * N)
    assert np.array_equal(np.diag(mpc.Q)[:13], np.asarray(preset["q_weights"], float))
    return m


def test_mpc_class_run_simulation_sequence_and_attributes(orc):
    """g1_mujoco_sim/src/run_simulation.py:71-106: in-place writes into x0 (13,1) and x_ref_hor, lists of arrays as arguments,
    update(..., x_current=MPC.x0, one_rollout=True); then every attribute the reference leaves current is read."""
    preset = presets.get_preset("python_default")
    mpc = MPC()
    mpc.init_matrices()
    assert np.all(mpc.Hessian == 0) and mpc.Aqp.shape == (13 * 15, 13)        # as after the reference's init_matrices
    rng = np.random.default_rng(2)
    for tick in range(2):
        mpc.x0[0:3] = rng.normal(0, 0.05, (3, 1))
        mpc.x0[3:6] = np.array([5.2e-2, 1e-4, 0.6]).reshape(3, 1) + rng.normal(0, 0.01, (3, 1))
        mpc.x0[6:9] = rng.normal(0, 0.1, (3, 1))
        mpc.x0[9:12] = rng.normal(0, 0.1, (3, 1))
        mpc.x0[12] = mpc.g
        mpc.x_ref_hor[0, :] = mpc.x0[:].copy().reshape(13)
        mpc.x_ref_hor[0:, 3:6] = [5.26790425e-02, 7.44339342e-05, 5.97983255e-01]
        mpc.x_ref_hor[0:, -1] = mpc.g
        feet = [np.array([-0.03, 0.12, 0.0]), np.array([0.14, 0.12, 0.0]), np.array([-0.03, -0.12, 0.0]), np.array([0.14, -0.12, 0.0])]
        c_horizon = [np.concatenate(feet) for _ in range(mpc.HORIZON_LENGTH)]
        contact_horizon = [np.array([1, 1, 1, 1]) for _ in range(mpc.HORIZON_LENGTH)]
        p_com_horizon = mpc.x_ref_hor[:, 3:6].copy()
        u_opt0, x_opt1 = mpc.update(contact_horizon, c_horizon, p_com_horizon, x_current=mpc.x0, one_rollout=True)
        assert u_opt0.shape == (15, 12) and x_opt1.shape == (15, 13) and mpc.qp_info["status"] == 0
        args = (mpc.x0.ravel().copy(), mpc.x_ref_hor.copy(), np.ones((15, 4)), np.array(c_horizon), p_com_horizon)
        m = _check_attributes(orc, mpc, preset, *args)
        p = orc.params_from_dict(preset)
        xs, _, cert = orc.qp_exact(p, m["H"], m["g"], args[2])
        assert cert["rc"] == 0 and np.abs(u_opt0.ravel() - xs).max() <= 1e-3
        assert np.abs(x_opt1[1] - orc.rollout(p, m["A"], m["Bd"], xs.reshape(15, 12), args[0], True)[1]).max() <= 1e-6
        assert np.all(x_opt1[2:] == 0.0) and np.array_equal(mpc.u_opt0.ravel(), u_opt0[0])


def test_mpc_class_main_py_sequence_and_rollout_of_a_modified_u_opt(orc):
    """g1_mpc/scripts/main.py:290-310: the fourteen step methods, solve_qp(), a manual x_opt[0] carry and compute_rollout()
    without x_current; compute_rollout must roll out the CURRENT u_opt (a caller-modified one included), not re-solve."""
    preset = presets.get_preset("cpp_node")
    mpc = MPC.cpp_node()
    mpc.init_matrices()
    x0, xr, ct, ch, pc = inputs.walking_problem(23, np.random.default_rng(5), preset)
    mpc.x0 = x0.copy()
    mpc.x_ref_hor = xr.copy()
    mpc.extract_psi(); mpc.rotation_matrix_T(); mpc.set_Q(); mpc.set_R(); mpc.calculate_A_continuous(); mpc.calculate_A_discrete()
    mpc.calculate_B_continuous([r for r in ch], pc); mpc.calculate_B_discrete(); mpc.calculate_Aqp(); mpc.calculate_Bqp(); mpc.calculate_Ac()
    mpc.calculate_bounds([list(r) for r in ct]); mpc.calculate_hessian(); mpc.calculate_gradient()
    res = mpc.solve_qp()
    m = _check_attributes(orc, mpc, preset, x0, xr, ct, ch, pc)
    p = orc.params_from_dict(preset)
    xs, _, cert = orc.qp_exact(p, m["H"], m["g"], ct)
    assert cert["rc"] == 0 and np.abs(res.x - xs).max() <= 1e-3 and np.all(mpc.x_opt == 0.0)      # solve_qp leaves x_opt alone
    mpc.x_opt[0, :] = np.squeeze(mpc.x0.copy())
    mpc.compute_rollout()
    want = orc.rollout(p, m["A"], m["Bd"], mpc.u_opt, x0, False)
    assert np.abs(mpc.x_opt - want).max() <= 1e-9 * np.abs(want).max()
    mpc.u_opt = 0.5 * mpc.u_opt
    mpc.compute_rollout(x0.reshape(13, 1), True)
    want = orc.rollout(p, m["A"], m["Bd"], mpc.u_opt, x0, True, x_opt=want)
    assert np.abs(mpc.x_opt - want).max() <= 1e-9 * np.abs(want).max()
    # warm start across update() calls: same answer, fewer iterations on the neighbouring tick
    u_cold, _ = mpc.update(ct, ch, pc, x_current=x0)
    it_cold = mpc.qp_info["iters"]
    mpc.warm_start = True
    mpc.update(ct, ch, pc, x_current=x0)              # first warm-enabled call has nothing to start from yet
    x0b = x0.copy(); x0b[3] += 0.002
    mpc.x0 = x0b
    u_warm, _ = mpc.update(ct, ch, pc, x_current=x0b)
    it_warm = mpc.qp_info["iters"]
    mpc.warm_start = False
    u_ref, _ = mpc.update(ct, ch, pc, x_current=x0b)
    assert np.abs(u_warm - u_ref).max() <= 1e-4 and it_warm < it_cold and mpc.qp_info["iters"] >= it_warm


def test_mpc_class_rejects_what_it_cannot_represent():
    m = MPC()
    m.Q_WEIGHTS = m.Q_WEIGHTS.copy(); m.Q_WEIGHTS[0, 1] = 1.0
    with pytest.raises(ValueError, match="diagonal"):
        m.init_matrices()
    m = MPC()
    m.R_WEIGHTS = np.diag(np.linspace(1e-3, 2e-3, 12))
    with pytest.raises(ValueError, match="uniform"):
        m.init_matrices()
    m = MPC(HORIZON_LENGTH=12)                          # a horizon with no compile-time kernel: runs (run-time layout)
    m.init_matrices()
    m.NUM_CONTACTS = 8
    m.R_WEIGHTS = np.diag(np.full(24, 1e-3))
    m.init_matrices()
    assert m.Bqp.shape == (13 * 12, 24 * 12)


# --------------------------------------------------------------------------------------------------------------------
# 8f-3 as main.py runs it
# --------------------------------------------------------------------------------------------------------------------
def test_main_loop_matches_cpu_composition(orc):
    preset = presets.get_preset("python_default", horizon=10)
    s = BatchedMPC(preset, device=0)
    planners = np.array([[1000., 0.25, 0.1, 0.2, 100], [0.2, 0.3, 0.05, 0.15, 30], [0.1, 0.2, 0.08, 0.2, 50]])
    T = 14
    out = s.main_loop(_dev(planners), T, total_duration=0.04 * T)
    torch.cuda.synchronize()
    assert int(out["status_any"].abs().sum()) == 0
    for b in range(len(planners)):
        tx, tu = node_oracle.main_loop(orc, preset, planners[b], T, total_duration=0.04 * T)
        assert np.abs(out["traj_x"][b].cpu().numpy() - tx).max() <= 1e-5, b
        assert np.abs(out["traj_u"][b].cpu().numpy() - tu).max() <= 5e-3, b
    warm = s.main_loop(_dev(planners), T, total_duration=0.04 * T, warm_start=True)
    assert int(warm["status_any"].abs().sum()) == 0
    assert float((warm["traj_x"] - out["traj_x"]).abs().max()) <= 1e-5
    # main.py feeds x_opt[2] back and switches its reference by quarters: consecutive solutions differ by 100-250 N, the warm
    # start stalls and the safeguard restarts cold after 5 iterations -- a loss, bounded (an IPM need not gain; measured, reported)
    assert int(warm["iters_sum"].sum()) <= int(out["iters_sum"].sum()) + 5 * T * len(planners)


def test_main_loop_on_the_scenario_main_py_ships(orc):
    """g1mpc_main_loop_batch on main.py's own scenario (mpc.MPC() defaults, GaitPlanner(1000, 0.25, 0.1, 0.2, 100), total_duration 4 s):
    (i) against the restated loop closed on the exact minimiser -- the loop whose logic tests/test_reference_callers_cpu.py holds to
    the reference's main() tick by tick; (ii) against the trajectory the reference's own main() produced through the OSQP shim
    (tests/golden/ref_main_loop.npz): that run stops every solve at eps 1e-3, so the two closed loops agree to millimetres in the
    CoM and centimetres per second in the velocities, not better."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "ref_main_loop.npz"))
    preset = presets.get_preset("python_default")
    s = BatchedMPC(preset, device=0)
    T = 32
    out = s.main_loop(_dev(z["planner"][None]), T, total_duration=float(z["total_duration"]), foot_offset=float(z["foot_offset"]))
    torch.cuda.synchronize()
    assert int(out["status_any"].abs().sum()) == 0
    tx, tu = out["traj_x"][0].cpu().numpy(), out["traj_u"][0].cpu().numpy()
    ex, eu = node_oracle.main_loop(orc, preset, tuple(z["planner"]), T, total_duration=float(z["total_duration"]))
    assert np.abs(tx - ex).max() <= 1e-5 and np.abs(tu - eu).max() <= 5e-3, (np.abs(tx - ex).max(), np.abs(tu - eu).max())
    d = np.abs(tx - z["traj_x"][:T]).max(0)
    assert d[:6].max() <= 5e-3 and d[6:12].max() <= 5e-2, d
    assert np.abs(tu - z["traj_u"][:T]).max() <= 40.0                  # (forces of an eps = 1e-3 OSQP iterate: tens of newtons off)


# --------------------------------------------------------------------------------------------------------------------
# C++ adapter (include/g1mpc_mpc.hpp) -- compiled by __graft_entry__.build(), run here against the Python path
# --------------------------------------------------------------------------------------------------------------------
def test_cpp_adapter_against_the_python_path(tmp_path, orc):
    exe = os.path.join(ROOT, "tests", "cpp", "test_mpc_adapter")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "tests", "cpp"), "test_mpc_adapter"])
    preset = presets.get_preset("cpp_node")
    x0, xr, ct, ch, pc = inputs.walking_problem(37, np.random.default_rng(9), preset)
    fin, fout = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    np.concatenate([x0.ravel(), xr.ravel(), ct.ravel(), ch.ravel(), pc.ravel()]).tofile(fin)
    r = subprocess.run([exe, fin, fout], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    o = np.fromfile(fout)
    N = 10
    cuts = np.cumsum([N * 12, N * 13, 12, 3, N * 13, N * 12, N * 13, N * 13])
    u1, x1, u0, rr, x_one, sol, x_roll, x_half = np.split(o, cuts)[:8]
    s = BatchedMPC(preset, device=0)
    ins = [_dev(a[None]) for a in (x0, xr, ct, ch, pc)]
    u, x, _ = s.update_batch(*ins, x_current=ins[0])
    assert np.array_equal(u1, u[0].cpu().numpy().ravel()) and np.array_equal(x1, x[0].cpu().numpy().ravel())      # same kernel, same bits
    assert np.array_equal(u0, u1[:12]) and np.array_equal(rr, ch[N - 1, 9:12] - pc[N - 1])
    _, xo, _ = s.update_batch(*ins, x_current=ins[0], one_rollout=True)
    assert np.array_equal(x_one, xo[0].cpu().numpy().ravel()) and np.array_equal(sol, u1)
    p = orc.params_from_dict(preset)
    m = orc.build(p, x0, xr, ct, ch, pc)
    want = orc.rollout(p, m["A"], m["Bd"], u1.reshape(N, 12), x0, False)
    assert np.abs(x_roll.reshape(N, 13) - want).max() <= 1e-9 * np.abs(want).max()
    want = orc.rollout(p, m["A"], m["Bd"], 0.5 * u1.reshape(N, 12), x0, False)
    assert np.abs(x_half.reshape(N, 13) - want).max() <= 1e-9 * np.abs(want).max()


# --------------------------------------------------------------------------------------------------------------------
# status: solved means gap AND dual residual
# --------------------------------------------------------------------------------------------------------------------
def test_status_reports_a_dual_residual_that_did_not_converge(orc):
    preset = presets.get_preset("cpp_node")
    ins = list(inputs.random_batch(preset, 512, seed=21, device="cuda"))
    strict = BatchedMPC(preset, device=0, dual_tol=1e-30)        # unreachable: every problem must come back flagged, not "solved"
    u, _, info = strict.update_batch(*ins, x_current=ins[0])
    assert set(torch.unique(info["status"]).tolist()) == {4}
    normal = BatchedMPC(preset, device=0)
    u2, _, info2 = normal.update_batch(*ins, x_current=ins[0])
    assert int((info2["status"] != 0).sum()) == 0
    assert torch.equal(u, u2) and torch.equal(info["iters"], info2["iters"])      # the flagged result is the same point
    assert float(info2["resid"][:, 1].max()) <= 1e-2             # absolute dual residual of the accepted points (|g| ~ 1e4)
