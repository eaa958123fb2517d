import glob
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def rel_err(a, b):
    """max-norm relative error: ||a-b||_max / max(||b||_max, tiny)."""
    a, b = np.asarray(a, float), np.asarray(b, float)
    both_inf = np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b))
    d = np.where(both_inf, 0.0, a - np.where(both_inf, 0.0, b))
    fin = np.abs(np.where(np.isfinite(b), b, 0.0)).max() if b.size else 0.0
    return float(np.abs(d).max() / max(fin, 1e-300)) if a.size else 0.0


def golden_cases():
    out = []
    for f in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "build_*.npz"))):
        out.append(os.path.basename(f)[len("build_"):-len(".npz")])
    return out


def load_golden(name):
    z = np.load(os.path.join(ROOT, "tests", "golden", f"build_{name}.npz"))
    d = {k: z[k] for k in z.files}
    d["preset"] = json.loads(str(d["preset"]))
    return d


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle
    oracle.lib()
    return oracle
