"""GPU test of the no-allocation entry point g1mpc_solve_batch_static_ws (include/g1mpc_b200.h): the global-workspace
variant of the solve kernel on the library's file-scope `static __device__` scratch array.  Called through the C-ABI
(ctypes) on device buffers and compared with the existing entry point g1mpc_solve_batch on the same inputs, on two calls in
a row (the second on the scratch the first -- and a launch of ANOTHER shape in between -- left behind).

Collected last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C

import pytest
import torch

from g1_locomotion_b200 import _lib, presets, inputs, BatchedMPC

pytestmark = pytest.mark.gpu

STATIC_WS_BYTES = 32 << 20          # G1MPC_STATIC_WS_BYTES


def _solve(entry, s, ins, B):
    """One call of a g1mpc_solve_batch-shaped entry point on the current stream; fresh outputs (x_opt poisoned: with x_current
    and a full rollout every row is written)."""
    N, nc = s.N, s.nc
    u = torch.full((B, N, 3 * nc), float("nan"), device="cuda", dtype=torch.float64)
    x = torch.full((B, N, 13), float("nan"), device="cuda", dtype=torch.float64)
    status = torch.full((B,), -7, device="cuda", dtype=torch.int32)
    iters = torch.full((B,), -7, device="cuda", dtype=torch.int32)
    resid = torch.full((B, 4), float("nan"), device="cuda", dtype=torch.float64)
    y = torch.full((B, 5 * nc * N), float("nan"), device="cuda", dtype=torch.float64)
    p = lambda t: C.c_void_p(t.data_ptr())
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    rc = entry(s._h, B, p(ins[0]), p(ins[1]), p(ins[2]), p(ins[3]), p(ins[4]), p(ins[0]), 0, p(u), p(x), p(status), p(iters),
               p(resid), p(y), stream)
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    return rc, dict(u=u, x=x, status=status, iters=iters, resid=resid, y=y), free_before - free_after


# (N, nc, B, bit_identical): N = 10 / nc = 4 is the bench shape -- the existing entry point runs its compile-time
# shared-memory instantiation, the new one the run-time-layout global-workspace instantiation of the same source; N = 12 / nc = 4
# has no compile-time instantiation (run-time layout on both sides); N = 20 / nc = 8 has a compile-time global-workspace
# instantiation.  B is larger than the number of resident warps, so every warp reuses its slot for several problems.
CASES = [(10, 4, 3000), (12, 4, 1500), (20, 8, 2100)]


def test_static_scratch_entry_point_matches_solve_batch_twice_in_a_row():
    lib = _lib.load()
    handles = {}
    for N, nc, B in CASES:
        preset = presets.get_preset("cpp_node", horizon=N, num_contacts=nc)
        handles[(N, nc)] = (BatchedMPC(preset, device=0), list(inputs.random_batch(preset, B, seed=101 + N, device="cuda")))
    want = {}
    for (N, nc, B) in CASES:
        s, ins = handles[(N, nc)]
        rc, ref, _ = _solve(lib.g1mpc_solve_batch, s, ins, B)
        assert rc == 0, lib.g1mpc_last_error()
        assert int((ref["status"] != 0).sum()) == 0
        want[(N, nc)] = ref
    # two rounds: in the second every shape finds the scratch as the LAST shape of the first round left it
    for rnd in range(2):
        for (N, nc, B) in CASES:
            s, ins = handles[(N, nc)]
            rc, got, grew = _solve(lib.g1mpc_solve_batch_static_ws, s, ins, B)
            assert rc == 0, lib.g1mpc_last_error()
            ref = want[(N, nc)]
            # the tolerance of this suite's variant-against-variant tests: the same arithmetic, so the same bits
            # (test_global_workspace_gives_the_same_bits_as_shared_memory, test_latency_mode_kernel_gives_the_same_bits...)
            for k in ("status", "iters", "u", "x", "resid", "y"):
                assert torch.equal(got[k], ref[k]), (rnd, N, nc, k, float((got[k].double() - ref[k].double()).abs().max()))
            if rnd == 1:
                assert grew <= 0, f"device memory in use grew by {grew} B across a call that must not allocate"


def test_static_scratch_entry_point_rejects_what_does_not_fit():
    lib = _lib.load()
    # one slot per SM must fit the 32 MiB array: N = 60 / nc = 8 needs 240,096 B per problem in flight, the limit on 148 SMs is 226,719 B
    preset = presets.get_preset("cpp_node", horizon=60, num_contacts=8)
    s = BatchedMPC(preset, device=0)
    ins = list(inputs.random_batch(preset, 8, seed=5, device="cuda"))
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    assert 240096 > STATIC_WS_BYTES // sms
    rc, out, _ = _solve(lib.g1mpc_solve_batch_static_ws, s, ins, 8)
    assert rc == -2, rc                                             # G1MPC_ERR_UNSUPPORTED_SHAPE
    msg = lib.g1mpc_last_error().decode()
    assert "240096" in msg and str(STATIC_WS_BYTES) in msg and "g1mpc_solve_batch" in msg, msg
    assert bool(torch.isnan(out["u"]).all()) and int((out["status"] != -7).sum()) == 0         # nothing ran
    # the existing entry point still serves the shape
    rc, out, _ = _solve(lib.g1mpc_solve_batch, s, ins, 8)
    assert rc == 0 and bool(torch.isfinite(out["u"]).all())
    # argument errors, and an empty batch
    s10 = BatchedMPC(presets.get_preset("cpp_node"), device=0)
    assert lib.g1mpc_solve_batch_static_ws(s10._h, 4, None, None, None, None, None, None, 0, None, None, None, None, None, None, None) == -1
    assert lib.g1mpc_solve_batch_static_ws(None, 4, None, None, None, None, None, None, 0, None, None, None, None, None, None, None) == -1
    assert lib.g1mpc_solve_batch_static_ws(s10._h, 0, None, None, None, None, None, None, 0, None, None, None, None, None, None, None) == 0
