"""CPU-only checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/g1mpc_b200.h declares, the params struct mirrors the header, and the product refuses to run
without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest
import torch

from g1_locomotion_b200 import _lib, presets

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "g1mpc_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(g1mpc_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    syms = _declared_symbols()
    assert set(syms) == set(_lib.EXPORTS), (syms, _lib.EXPORTS)
    for s in syms:
        assert getattr(lib, s) is not None
    assert lib.g1mpc_abi_version() == 7 == _lib.ABI_VERSION


def test_params_struct_matches_header_layout():
    # 2 int32 + 6 double + 9 + 13 + 2 double + 1 double + 2 int32 + 3 double + 2 int32
    assert C.sizeof(_lib.G1mpcParams) == 8 + 8 * 6 + 8 * 9 + 8 * 13 + 8 * 2 + 8 + 8 + 8 * 3 + 8
    assert _lib.G1mpcParams.layout.offset == 260 and _lib.G1mpcParams.dual_tol.offset == 264
    assert _lib.G1mpcParams.mu.offset == 8 and _lib.G1mpcParams.inertia_body.offset == 56


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    lib = _lib.load()
    from g1_locomotion_b200.mpc import _c_params, BatchedMPC, MPC
    h = C.c_void_p()
    rc = lib.g1mpc_create(C.byref(_c_params(presets.get_preset("cpp_node"))), 0, C.byref(h))
    assert rc == -4 and not h.value                      # G1MPC_ERR_NO_DEVICE
    assert b"no CPU fallback" in lib.g1mpc_last_error()
    with pytest.raises(RuntimeError):
        BatchedMPC("cpp_node")
    m = MPC()
    with pytest.raises(RuntimeError):
        m.init_matrices()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "g1_locomotion_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.replace("no oracle", ""), f"{f} mentions the oracle"


def test_presets_match_reference_constants():
    c, p = presets.get_preset("cpp_node"), presets.get_preset("python_default")
    assert (c["mu"], c["fz_min"], c["fz_max"], c["dt"], c["horizon"]) == (0.8, 10.0, 1000.0, 0.05, 10)   # mpc.hpp:22
    assert (p["mu"], p["fz_min"], p["fz_max"], p["dt"], p["horizon"]) == (0.3, 10.0, 666.0, 0.04, 15)    # mpc.py:12
    assert c["cost_scale"] == 2.0 and p["cost_scale"] == 1.0
    assert c["q_weights"][12] == 0.0 and p["q_weights"][12] == 0.0
    with pytest.raises(KeyError):
        presets.get_preset("cpp_node", bogus=1)
