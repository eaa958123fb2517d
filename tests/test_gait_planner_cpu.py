"""SURVEY 8f-1, Python half: GaitPlanner.plan_gait (gait_planner.py:27-62) and its mapping onto (contact, c_horizon)
(main.py:259-285).  tests/golden/gait_plan.npz was produced by oracle/gen_golden.py from the REFERENCE class imported
unmodified (matplotlib stubbed); here the numpy restatement (oracle/node_oracle.py) is pinned to it bit for bit, and --
when /root/reference is present -- to the live class.  The CUDA kernel is compared with the same file in
tests/test_round2_gpu.py."""
import os

import numpy as np
import pytest

from conftest import ROOT
from oracle import node_oracle


def _cases():
    z = np.load(os.path.join(ROOT, "tests", "golden", "gait_plan.npz"))
    return [dict(planner=z["planner"][i], t=float(z["t"][i]), N=int(z["N"][i]), dt=float(z["dt"][i]),
                 contact=z["contact"][i][:int(z["N"][i])], c_horizon=z["c_horizon"][i][:int(z["N"][i])]) for i in range(len(z["t"]))]


def test_restated_planner_and_mapping_match_reference_golden():
    cases = _cases()
    assert len(cases) >= 200
    kinds = set()
    for c in cases:
        phases = node_oracle.plan_gait(*c["planner"])
        ct, ch = node_oracle.map_gait_phases(phases, c["t"], c["N"], c["dt"], 0.08)
        assert np.array_equal(ct, c["contact"]) and np.array_equal(ch, c["c_horizon"])
        kinds |= {tuple(r) for r in ct}
    assert kinds == {(1, 1, 1, 1), (1, 1, 0, 0), (0, 0, 1, 1)}          # both, left and right support all occur


def test_planner_semantics_spelled_out():
    """The facts the kernel relies on: running-sum phase boundaries, the first step half as long as the others, left foot
    at -step_width (as coded, gait_planner.py:19), the last phase used past the end (main.py:270-271)."""
    ph = node_oracle.plan_gait(0.2, 0.8, 0.3, 0.2, 4)
    assert [p["support_leg"] for p in ph] == ["both", "left", "right", "left", "right"]
    assert ph[1]["start_time"] == 0.2 and ph[2]["start_time"] == 0.2 + 0.8 and ph[3]["start_time"] == (0.2 + 0.8) + 0.8
    assert ph[0]["left_foot"][1] == -0.2 and ph[0]["right_foot"][1] == 0.2
    assert ph[2]["right_foot"][0] == 0.3 and ph[3]["left_foot"][0] == 0.6 and ph[4]["right_foot"][0] == 0.3 + 0.6
    ct, ch = node_oracle.map_gait_phases(ph, 100.0, 3, 0.05)
    assert np.all(ct == [0, 0, 1, 1]) and ch[0, 0] == ph[4]["left_foot"][0] - 0.08 and ch[0, 9] == ph[4]["right_foot"][0] + 0.08


def test_live_reference_planner():
    from oracle import ref_loader
    if not ref_loader.available():
        pytest.skip("reference tree not present (GPU box)")
    GP = ref_loader.load_gait_planner()
    rng = np.random.default_rng(3)
    for _ in range(40):
        pl = (float(rng.uniform(0, 1)), float(rng.uniform(0.05, 1)), float(rng.uniform(0.01, 0.4)), float(rng.uniform(0.05, 0.3)), int(rng.integers(0, 40)))
        ref = GP(*pl).plan_gait()
        ours = node_oracle.plan_gait(*pl)
        assert len(ref) == len(ours)
        for a, b in zip(ours, ref):
            assert a["start_time"] == b["start_time"] and a["end_time"] == b["end_time"] and a["support_leg"] == b["support_leg"]
            assert np.array_equal(a["left_foot"], b["left_foot"]) and np.array_equal(a["right_foot"], b["right_foot"])
