"""Whole-config parity net ("weak" 1-3): every problem of configs 1, 2 and 4 and 4,096 random picks of configs 3
and 5 get an ALGORITHM-INDEPENDENT optimality certificate on REFERENCE-built matrices -- oracle.build restates
mpc.cpp:153-376 and is pinned to the reference's own mpc.py at 1e-12 (tests/test_oracle_golden.py); oracle.kkt_residuals
is plain numpy.  No active-set solve is needed for that, so it scales to whole batches; the exact minimiser x* is compared
on a few hundred problems on top.  Thresholds: primal violation 0 (1e-9), stationarity |H u + g + A'y|inf and
complementarity below 1e-6 (1 + |g|inf) -- three orders under the reference solver's own eps_rel = 1e-3 (mpc.cpp:385)."""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest
import torch

from conftest import ROOT
from g1_locomotion_b200 import presets, inputs, BatchedMPC, MPC

pytestmark = pytest.mark.gpu
THREADS = max(4, min(32, len(os.sched_getaffinity(0))))


def _dev(arrs):
    return [torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float64).cuda() for a in arrs]


def _certify(orc, preset, host, u, y, idx, stat_tol=1e-6, x_star=None, n_exact=0):
    """KKT certificate of problems `idx` (threads: the C build releases the GIL).  Returns the worst relative stationarity."""
    p = orc.params_from_dict(preset)

    def one(b):
        m = orc.build(p, *[a[b] for a in host])
        xg = u[b].ravel()
        k = orc.kkt_residuals(m["H"], m["g"], m["Acon"], m["l"], m["u"], xg, y[b])
        gs = 1.0 + np.abs(m["g"]).max()
        sw = np.repeat(host[2][b].ravel() == 0, 3)
        err = None
        if x_star is not None:
            err = np.abs(xg - x_star[b].ravel()).max()
        elif b in exact_set:
            xs, _, cert = orc.qp_exact(p, m["H"], m["g"], host[2][b])
            assert cert["rc"] == 0, (b, cert)
            err = np.abs(xg - xs).max()
        return b, k["stationarity"] / gs, k["primal"], k["complementarity"] / gs, bool(np.all(xg[sw] == 0.0)), err

    exact_set = set(list(idx)[:n_exact])
    with ThreadPoolExecutor(THREADS) as ex:
        res = list(ex.map(one, idx))
    worst = max(r[1] for r in res)
    for b, st, pr, cp, swing_zero, err in res:
        assert pr <= 1e-9 and st <= stat_tol and cp <= stat_tol and swing_zero, (b, st, pr, cp)
        if err is not None:
            assert err <= 1e-3, (b, err)
    return worst


def test_config1_every_tick_of_the_8d_replay(orc):
    """configs[0] as SURVEY 8d specifies it (tests/golden/config1_replay.npz: 200 ticks closed on the oracle's x_opt[1] + noise):
    as one batch through the device API, and tick by tick through the drop-in MPC class; every tick against the stored x*."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "config1_replay.npz"))
    host = [z[k] for k in ("x0", "x_ref", "contact", "c_horizon", "p_com")]
    preset = presets.get_preset("cpp_node")
    s = BatchedMPC(preset, device=0)
    ins = _dev(host)
    u, x, info = s.update_batch(*ins, x_current=ins[0], want_multipliers=True)
    assert int((info["status"] != 0).sum()) == 0
    un, yn = u.cpu().numpy(), info["y"].cpu().numpy()
    T = host[0].shape[0]
    assert T == 200
    _certify(orc, preset, host, un, yn, range(T), x_star=z["u_star"])
    assert np.abs(x.cpu().numpy() - z["x_opt"]).max() <= 1e-6              # the oracle's rollout of its own minimiser
    m = MPC.cpp_node()
    m.init_matrices()
    for t in range(0, T, 4):
        m.x0 = host[0][t].copy(); m.x_ref_hor = host[1][t].copy()
        ut, xt = m.update(host[2][t], host[3][t], host[4][t], x_current=host[0][t].copy())
        assert np.array_equal(ut, un[t]) and m.qp_info["status"] == 0      # B = 1 through the host API: the same bits


def test_config2_every_walking_problem(orc):
    preset = presets.get_preset("cpp_node")
    host = list(inputs.walking_batch(preset, 1024))
    s = BatchedMPC(preset, device=0)
    ins = _dev(host)
    u, x, info = s.update_batch(*ins, x_current=ins[0], want_multipliers=True)
    assert int((info["status"] != 0).sum()) == 0
    worst = _certify(orc, preset, host, u.cpu().numpy(), info["y"].cpu().numpy(), range(1024), n_exact=256)
    print("config 2: worst relative stationarity %.1e over 1,024 problems, x* on 256" % worst)


@pytest.mark.parametrize("seed,B,name", [(7, 65536, "config 3"), (13 + 5, 131072, "config 5, shard of rank 5")])
def test_configs_3_and_5_random_picks(orc, seed, B, name):
    preset = presets.get_preset("cpp_node")
    ins = list(inputs.random_batch(preset, B, seed=seed, device="cuda"))
    s = BatchedMPC(preset, device=0)
    u, x, info = s.update_batch(*ins, x_current=ins[0], want_multipliers=True)
    assert int((info["status"] != 0).sum()) == 0
    pick = np.sort(np.random.default_rng(seed).choice(B, 4096, replace=False))
    tp = torch.as_tensor(pick, device="cuda")
    host = [t[tp].cpu().numpy() for t in ins]
    worst = _certify(orc, preset, host, u[tp].cpu().numpy(), info["y"][tp].cpu().numpy(), range(4096), n_exact=64)
    print("%s: worst relative stationarity %.1e over 4,096 picks of %d" % (name, worst, B))


@pytest.mark.parametrize("N,B", [(10, 4096), (20, 1024), (40, 256)])
def test_config4_every_problem(orc, N, B):
    """Config 4 at SURVEY 8d's sizes (nc = 8, seed 11).  The N = 20 and N = 40 batches run the global-workspace kernel.
    At N = 40 the dual residual of a few problems in a thousand sits on the FP64 floor of the stage-wise solves
    (profiles/r2_a_dual_residual_by_shape.txt): those come back with status 4 and are held to 1e-4 here -- still ten times
    under OSQP's own tolerance -- every other problem to 1e-6."""
    preset = presets.get_preset("cpp_node", horizon=N, num_contacts=8)
    ins = list(inputs.random_batch(preset, B, seed=11, device="cuda"))
    s = BatchedMPC(preset, device=0)
    u, x, info = s.update_batch(*ins, x_current=ins[0], want_multipliers=True)
    st = info["status"].cpu().numpy()
    assert set(np.unique(st)) <= ({0} if N < 40 else {0, 4}), np.unique(st, return_counts=True)
    assert (st == 4).sum() <= 0.02 * B
    host = [t.cpu().numpy() for t in ins]
    un, yn = u.cpu().numpy(), info["y"].cpu().numpy()
    _certify(orc, preset, host, un, yn, np.where(st == 0)[0])
    if (st == 4).any():
        _certify(orc, preset, host, un, yn, np.where(st == 4)[0], stat_tol=1e-4)
