"""ctypes loader for the in-tree C-ABI library libg1mpc_b200.so (include/g1mpc_b200.h).

No fallback of any kind: if the library is missing or no CUDA device is present the product raises.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("G1MPC_B200_LIB") or os.path.join(_HERE, "libg1mpc_b200.so")   # override: A/B runs of two builds
_lib = None

EXPORTS = ["g1mpc_create", "g1mpc_destroy", "g1mpc_solve_batch", "g1mpc_solve_batch_warm", "g1mpc_solve_batch_static_ws", "g1mpc_rollout_batch", "g1mpc_rollout_host", "g1mpc_emit_qp_batch",
           "g1mpc_update_host", "g1mpc_update_host_begin", "g1mpc_update_host_end", "g1mpc_set_host_warm_start", "g1mpc_layout_info",
           "g1mpc_walk_inputs_batch", "g1mpc_wbid_refs_batch", "g1mpc_closed_loop_batch", "g1mpc_closed_loop_batch_ex", "g1mpc_plan_gait_batch",
           "g1mpc_main_loop_batch", "g1mpc_kernel_info", "g1mpc_fp64_peak_tflops", "g1mpc_last_error", "g1mpc_abi_version"]
ABI_VERSION = 7


class G1mpcParams(C.Structure):
    """Mirror of struct g1mpc_params (include/g1mpc_b200.h)."""
    _fields_ = [("horizon", C.c_int32), ("num_contacts", C.c_int32), ("mu", C.c_double), ("fz_min", C.c_double),
                ("fz_max", C.c_double), ("dt", C.c_double), ("mass", C.c_double), ("gravity", C.c_double),
                ("inertia_body", C.c_double * 9), ("q_weights", C.c_double * 13), ("r_weight", C.c_double),
                ("cost_scale", C.c_double), ("gap_tol", C.c_double), ("max_iter", C.c_int32), ("layout", C.c_int32),
                ("dual_tol", C.c_double), ("warm_margin", C.c_double), ("warm_mu0", C.c_double),
                ("workspace", C.c_int32), ("reserved", C.c_int32)]


def build(force=False, extra=""):
    """Compile csrc/g1mpc_kernels.cu for sm_100a into the package directory (nvcc cross-compiles without a GPU)."""
    src = os.path.join(_HERE, "csrc", "g1mpc_kernels.cu")
    hdr = os.path.join(_HERE, "..", "include", "g1mpc_b200.h")
    deps = [src, hdr] + [os.path.join(_HERE, "csrc", f) for f in ("srbd_solve_warp.cuh", "node_adaptors.cuh")]
    stale = (not os.path.exists(LIB_PATH)) or os.path.getmtime(LIB_PATH) < max(os.path.getmtime(d) for d in deps)
    if force or stale:
        cmd = ["make", "-C", os.path.join(_HERE, "csrc"), "-B", "../libg1mpc_b200.so"]
        if extra:
            cmd.append(f"EXTRA={extra}")
        subprocess.check_call(cmd)
    return LIB_PATH


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(g1_locomotion_b200 has no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    vp, i64, i32p, dp = C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p
    lib.g1mpc_create.argtypes = [C.POINTER(G1mpcParams), C.c_int, C.POINTER(vp)]
    lib.g1mpc_create.restype = C.c_int
    lib.g1mpc_destroy.argtypes = [vp]
    lib.g1mpc_destroy.restype = None
    lib.g1mpc_solve_batch.argtypes = [vp, i64, dp, dp, dp, dp, dp, dp, C.c_int, dp, dp, i32p, i32p, dp, dp, vp]
    lib.g1mpc_solve_batch.restype = C.c_int
    lib.g1mpc_solve_batch_static_ws.argtypes = [vp, i64, dp, dp, dp, dp, dp, dp, C.c_int, dp, dp, i32p, i32p, dp, dp, vp]
    lib.g1mpc_solve_batch_static_ws.restype = C.c_int
    lib.g1mpc_solve_batch_warm.argtypes = [vp, i64, dp, dp, dp, dp, dp, dp, C.c_int, dp, C.c_int, dp, dp, i32p, i32p, dp, dp, vp]
    lib.g1mpc_solve_batch_warm.restype = C.c_int
    lib.g1mpc_rollout_batch.argtypes = [vp, i64, dp, dp, dp, dp, dp, C.c_int, dp, vp]
    lib.g1mpc_rollout_batch.restype = C.c_int
    lib.g1mpc_rollout_host.argtypes = [vp, i64, dp, dp, dp, dp, dp, C.c_int, dp]
    lib.g1mpc_rollout_host.restype = C.c_int
    lib.g1mpc_set_host_warm_start.argtypes = [vp, C.c_int, C.c_int]
    lib.g1mpc_set_host_warm_start.restype = C.c_int
    lib.g1mpc_closed_loop_batch_ex.argtypes = [vp, i64, C.c_int32, C.c_int] + [dp] * 8 + [i32p, i32p, vp]
    lib.g1mpc_closed_loop_batch_ex.restype = C.c_int
    lib.g1mpc_plan_gait_batch.argtypes = [vp, i64, dp, dp, C.c_double, dp, dp, vp]
    lib.g1mpc_plan_gait_batch.restype = C.c_int
    lib.g1mpc_main_loop_batch.argtypes = [vp, i64, C.c_int32, dp, C.c_double, C.c_double, C.c_int, dp, dp, dp, dp, dp, dp, i32p, i32p, vp]
    lib.g1mpc_main_loop_batch.restype = C.c_int
    lib.g1mpc_emit_qp_batch.argtypes = [vp, i64] + [dp] * 14 + [vp]
    lib.g1mpc_emit_qp_batch.restype = C.c_int
    lib.g1mpc_update_host.argtypes = [vp, i64, dp, dp, dp, dp, dp, dp, C.c_int, dp, dp, i32p, i32p, dp]
    lib.g1mpc_update_host.restype = C.c_int
    lib.g1mpc_update_host_begin.argtypes = [vp, C.c_int, i64, dp, dp, dp, dp, dp, dp, C.c_int, dp, dp, i32p, i32p, dp]
    lib.g1mpc_update_host_begin.restype = C.c_int
    lib.g1mpc_update_host_end.argtypes = [vp, C.c_int]
    lib.g1mpc_update_host_end.restype = C.c_int
    lib.g1mpc_walk_inputs_batch.argtypes = [vp, i64] + [dp] * 9 + [vp]
    lib.g1mpc_walk_inputs_batch.restype = C.c_int
    lib.g1mpc_wbid_refs_batch.argtypes = [vp, i64, dp, dp, dp, C.c_double, dp, vp]
    lib.g1mpc_wbid_refs_batch.restype = C.c_int
    lib.g1mpc_closed_loop_batch.argtypes = [vp, i64, C.c_int32] + [dp] * 8 + [i32p, vp]
    lib.g1mpc_closed_loop_batch.restype = C.c_int
    lib.g1mpc_kernel_info.argtypes = [vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    lib.g1mpc_kernel_info.restype = C.c_int
    lib.g1mpc_layout_info.argtypes = [vp, i64, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    lib.g1mpc_layout_info.restype = C.c_int
    lib.g1mpc_fp64_peak_tflops.argtypes = [C.c_int, C.POINTER(C.c_double)]
    lib.g1mpc_fp64_peak_tflops.restype = C.c_int
    lib.g1mpc_last_error.restype = C.c_char_p
    lib.g1mpc_abi_version.restype = C.c_int
    if lib.g1mpc_abi_version() != ABI_VERSION:
        raise RuntimeError(f"{LIB_PATH}: ABI version {lib.g1mpc_abi_version()}, this package expects {ABI_VERSION}: rebuild "
                           "(python -c 'import __graft_entry__ as g; g.build()')")
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().g1mpc_last_error().decode(errors="replace")
        raise RuntimeError(f"{what} failed (g1mpc_error {rc}): {msg}")
