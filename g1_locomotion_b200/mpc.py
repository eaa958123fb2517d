"""Host-side mirror of the reference MPC interface over the B200 C-ABI.

`BatchedMPC`  -- the batched solver: update_batch() on CUDA float64 tensors (device pointers go
                 straight into g1mpc_solve_batch), update_host() on numpy arrays (H2D/D2H inside the call).
`MPC`         -- drop-in for the reference Python class g1_mpc/scripts/mpc.py:10-391 (same constructor,
                 attributes and step-method names; one problem = batch of 1).  The C++ class
                 g1_mpc::MPC (mpc.hpp:15-199) maps onto the same calls.
There is no CPU path: every numeric result comes from the CUDA kernels.
"""
import ctypes as C

import contextlib

import numpy as np
import torch

from . import _lib
from .presets import get_preset, GRAVITY, ROBOT_MASS

NS = 13


def _c_params(p, gap_tol=0.0, max_iter=0, layout=0, dual_tol=0.0, warm_margin=0.0, warm_mu0=0.0, workspace=0):
    cp = _lib.G1mpcParams()
    cp.horizon, cp.num_contacts = int(p["horizon"]), int(p["num_contacts"])
    cp.mu, cp.fz_min, cp.fz_max, cp.dt = float(p["mu"]), float(p["fz_min"]), float(p["fz_max"]), float(p["dt"])
    cp.mass, cp.gravity = float(p["mass"]), float(p.get("gravity", GRAVITY))
    cp.inertia_body[:] = [float(v) for v in np.asarray(p["inertia_body"], float).ravel()]
    cp.q_weights[:] = [float(v) for v in np.asarray(p["q_weights"], float).ravel()]
    cp.r_weight, cp.cost_scale = float(p["r_weight"]), float(p["cost_scale"])
    cp.gap_tol, cp.max_iter, cp.layout = float(gap_tol), int(max_iter), int(layout)
    cp.dual_tol, cp.warm_margin, cp.warm_mu0 = float(dual_tol), float(warm_margin), float(warm_mu0)
    cp.workspace, cp.reserved = int(workspace), 0
    return cp


class BatchedMPC:
    """Batch of independent SRBD-MPC problems sharing one parameter set, on one CUDA device."""

    def __init__(self, preset="cpp_node", device=None, gap_tol=0.0, max_iter=0, layout=0, dual_tol=0.0, warm_margin=0.0,
                 warm_mu0=0.0, workspace=0, nvtx=False, **overrides):
        """layout: 0 = lane layout chosen per launch, 1 / 2 = one / two problems per warp (bit-identical results);
        workspace: 0 = auto, 1 = shared memory, 2 = global memory behind the caches, 3 = hybrid (long horizons; bit-identical results);
        nvtx: wrap every entry point in an NVTX range "g1mpc.<call>" (what the reference's chrono / ros::WallTime timers around
        updateMPC become on a GPU timeline, mpc.cpp:382,404-412, mpc_to_wbid_node.cpp:476,496);
        dual_tol, warm_margin, warm_mu0: see include/g1mpc_b200.h (0 = defaults)."""
        self.params = get_preset(preset, **overrides) if isinstance(preset, str) else dict(preset, **overrides)
        if not torch.cuda.is_available():
            raise RuntimeError("g1_locomotion_b200 needs a CUDA device (no CPU fallback)")
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else
                                   (device if isinstance(device, int) else torch.device(device).index or 0))
        self.N, self.nc = self.params["horizon"], self.params["num_contacts"]
        self.nu = 3 * self.nc
        self._lib = _lib.load()
        self._h = C.c_void_p()
        self._pinned_keep, self._inflight = [], {}
        self._nvtx = bool(nvtx)
        cp = _c_params(self.params, gap_tol, max_iter, layout, dual_tol, warm_margin, warm_mu0, workspace)
        self._b1 = None
        _lib.check(self._lib.g1mpc_create(C.byref(cp), self.device.index, C.byref(self._h)), "g1mpc_create")

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.g1mpc_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- helpers ---------------------------------------------------------------------------------
    def _range(self, name):
        return torch.cuda.nvtx.range("g1mpc." + name) if self._nvtx else contextlib.nullcontext()

    def _chk(self, t, shape, name):
        if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.float64 and t.is_contiguous()):
            raise TypeError(f"{name}: expected a contiguous CUDA float64 tensor")
        if t.device != self.device:
            raise ValueError(f"{name}: tensor on {t.device}, solver on {self.device}")
        if tuple(t.shape) != tuple(shape):
            raise ValueError(f"{name}: expected shape {tuple(shape)}, got {tuple(t.shape)}")
        return C.c_void_p(t.data_ptr())

    def kernel_info(self, B=None):
        """Launch configuration of the solve kernel; with B, that of the lane layout a batch of B problems uses."""
        a, b, c = C.c_int32(), C.c_int32(), C.c_int32()
        _lib.check(self._lib.g1mpc_kernel_info(self._h, C.byref(a), C.byref(b), C.byref(c)), "g1mpc_kernel_info")
        if B is None:
            return dict(smem_bytes=a.value, threads=b.value, ctas_per_sm=c.value)
        ppw = C.c_int32()
        _lib.check(self._lib.g1mpc_layout_info(self._h, int(B), C.byref(ppw), C.byref(a), C.byref(c)), "g1mpc_layout_info")
        mode = ("latency: one problem per two-warp CTA" if ppw.value == -1 else
                "hybrid workspace (stage factors in shared, per-contact state in global memory), persistent grid" if ppw.value == -2 else
                "hybrid workspace, two problems per warp, persistent grid" if ppw.value == -3 else
                "global-memory workspace, persistent grid" if a.value == 0 else
                "two problems per warp" if ppw.value == 2 else "one problem per warp")
        return dict(smem_bytes=a.value, threads=64 if ppw.value == -1 else b.value, ctas_per_sm=c.value,
                    problems_per_warp={-2: 1, -3: 2}.get(ppw.value, ppw.value), mode=mode)

    # -- device API ------------------------------------------------------------------------------
    def update_batch(self, x0, x_ref, contact, c_horizon, p_com, x_current=None, *, one_rollout=False,
                     u_opt=None, x_opt=None, want_resid=True, want_multipliers=False, u_warm=None, warm_shift=1):
        """MPC.update over a batch of CUDA tensors (mpc.py:360-377 / mpc.cpp:448-473).

        Returns (u_opt[B,N,nu], x_opt[B,N,13], info) with info = {status[B], iters[B], resid[B,4], y[B,5ncN]?}.
        u_warm [B,N,nu]: start from that previous solution, shifted by warm_shift stages (g1mpc_solve_batch_warm).
        Asynchronous on the current torch CUDA stream."""
        B, N, nc, nu = x0.shape[0], self.N, self.nc, self.nu
        px0 = self._chk(x0, (B, NS), "x0"); pxr = self._chk(x_ref, (B, N, NS), "x_ref")
        pct = self._chk(contact, (B, N, nc), "contact"); pch = self._chk(c_horizon, (B, N, nu), "c_horizon")
        ppc = self._chk(p_com, (B, N, 3), "p_com")
        pxc = self._chk(x_current, (B, NS), "x_current") if x_current is not None else None
        if u_opt is None:
            u_opt = torch.empty(B, N, nu, device=self.device, dtype=torch.float64)
        if x_opt is None:
            # initMatrices zero-fills x_opt_; the rollout overwrites every row when x_current is given and one_rollout
            # is off (mpc.py:379-386), so the fill kernel is only queued when a row would survive
            alloc = torch.empty if (x_current is not None and not one_rollout) else torch.zeros
            x_opt = alloc(B, N, NS, device=self.device, dtype=torch.float64)
        pu = self._chk(u_opt, (B, N, nu), "u_opt"); pxo = self._chk(x_opt, (B, N, NS), "x_opt")
        status = torch.empty(B, device=self.device, dtype=torch.int32)
        iters = torch.empty(B, device=self.device, dtype=torch.int32)
        resid = torch.empty(B, 4, device=self.device, dtype=torch.float64) if want_resid else None
        y = torch.empty(B, 5 * nc * N, device=self.device, dtype=torch.float64) if want_multipliers else None
        pw = self._chk(u_warm, (B, N, nu), "u_warm") if u_warm is not None else None
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        with torch.cuda.device(self.device), self._range("solve_batch"):
            rc = self._lib.g1mpc_solve_batch_warm(self._h, B, px0, pxr, pct, pch, ppc, pxc, int(one_rollout), pw, int(warm_shift),
                                                  pu, pxo, C.c_void_p(status.data_ptr()), C.c_void_p(iters.data_ptr()),
                                                  C.c_void_p(resid.data_ptr()) if want_resid else None,
                                                  C.c_void_p(y.data_ptr()) if want_multipliers else None, stream)
        _lib.check(rc, "g1mpc_solve_batch_warm")
        info = dict(status=status, iters=iters, resid=resid)
        if want_multipliers:
            info["y"] = y
        return u_opt, x_opt, info

    def emit_qp(self, x0, x_ref, contact, c_horizon, p_com, which=("A", "Bd", "Aqp", "Bqp", "H", "g", "Acon", "l", "u")):
        """Materialise the reference's dense intermediates on the device (verify mode / attribute access)."""
        B, N, nc, nu = x0.shape[0], self.N, self.nc, self.nu
        n, nx, m = nu * N, NS * N, 5 * nc * N
        px0 = self._chk(x0, (B, NS), "x0"); pxr = self._chk(x_ref, (B, N, NS), "x_ref")
        pct = self._chk(contact, (B, N, nc), "contact"); pch = self._chk(c_horizon, (B, N, nu), "c_horizon")
        ppc = self._chk(p_com, (B, N, 3), "p_com")
        shapes = dict(A=(B, NS, NS), Bd=(B, nx, nu), Aqp=(B, nx, NS), Bqp=(B, nx, n), H=(B, n, n), g=(B, n),
                      Acon=(B, m, n), l=(B, m), u=(B, m))
        which = set(which)
        if ("l" in which) != ("u" in which):
            which |= {"l", "u"}
        out = {k: torch.empty(shapes[k], device=self.device, dtype=torch.float64) for k in shapes if k in which}
        ptr = lambda k: C.c_void_p(out[k].data_ptr()) if k in out else None
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        with torch.cuda.device(self.device), self._range("emit_qp_batch"):
            rc = self._lib.g1mpc_emit_qp_batch(self._h, B, px0, pxr, pct, pch, ppc, ptr("A"), ptr("Bd"), ptr("Aqp"),
                                               ptr("Bqp"), ptr("H"), ptr("g"), ptr("Acon"), ptr("l"), ptr("u"), stream)
        _lib.check(rc, "g1mpc_emit_qp_batch")
        return out

    def rollout(self, x_ref, c_horizon, p_com, u_opt, x_current=None, *, one_rollout=False, x_opt=None):
        """MPC.compute_rollout of a given u_opt (mpc.py:379-391 / mpc.cpp:427-446) over a batch of CUDA tensors; x_opt is
        in/out (zeros if None).  No solve."""
        B, N, nu = u_opt.shape[0], self.N, self.nu
        pxr = self._chk(x_ref, (B, N, NS), "x_ref"); pch = self._chk(c_horizon, (B, N, nu), "c_horizon")
        ppc = self._chk(p_com, (B, N, 3), "p_com"); pu = self._chk(u_opt, (B, N, nu), "u_opt")
        pxc = self._chk(x_current, (B, NS), "x_current") if x_current is not None else None
        if x_opt is None:
            x_opt = torch.zeros(B, N, NS, device=self.device, dtype=torch.float64)
        pxo = self._chk(x_opt, (B, N, NS), "x_opt")
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        with torch.cuda.device(self.device), self._range("rollout_batch"):
            rc = self._lib.g1mpc_rollout_batch(self._h, B, pxr, pch, ppc, pu, pxc, int(one_rollout), pxo, stream)
        _lib.check(rc, "g1mpc_rollout_batch")
        return x_opt

    def solve_one(self, x0, x_ref, contact, c_horizon, p_com, x_current=None, *, one_rollout=False, u_warm=None, warm_shift=1):
        """Low-overhead single-problem call (the reference's use: one updateMPC per control tick,
        mpc_to_wbid_node.cpp:494): B = 1 device tensors in, results in buffers that are allocated once and reused
        (u_opt[1,N,nu], x_opt[1,N,13], status[1], iters[1]) -- no allocation, no shape checks beyond the first call."""
        if self._b1 is None:
            mk = lambda *s, dt=torch.float64: torch.zeros(*s, device=self.device, dtype=dt)
            self._b1 = dict(u=mk(1, self.N, self.nu), x=mk(1, self.N, NS), st=mk(1, dt=torch.int32), it=mk(1, dt=torch.int32))
            self._b1["ptrs"] = tuple(C.c_void_p(self._b1[k].data_ptr()) for k in ("u", "x", "st", "it"))
        pu, pxo, pst, pit = self._b1["ptrs"]
        vp = C.c_void_p
        rc = self._lib.g1mpc_solve_batch_warm(self._h, 1, vp(x0.data_ptr()), vp(x_ref.data_ptr()), vp(contact.data_ptr()),
                                              vp(c_horizon.data_ptr()), vp(p_com.data_ptr()),
                                              vp(x_current.data_ptr()) if x_current is not None else None, int(one_rollout),
                                              vp(u_warm.data_ptr()) if u_warm is not None else None, int(warm_shift), pu, pxo, pst, pit,
                                              None, None, vp(torch.cuda.current_stream(self.device).cuda_stream))
        if rc:
            _lib.check(rc, "g1mpc_solve_batch_warm")
        return self._b1["u"], self._b1["x"], self._b1

    def set_host_warm_start(self, enable=True, shift=1):
        """Host API (update_host / update_host_begin): start every call from the solution of the previous call on the same slot
        (g1mpc_set_host_warm_start).  Off by default = the reference's cold solve per call."""
        _lib.check(self._lib.g1mpc_set_host_warm_start(self._h, int(bool(enable)), int(shift)), "g1mpc_set_host_warm_start")

    def plan_gait(self, planner, current_time, foot_offset=0.08):
        """GaitPlanner.plan_gait + the mapping of main.py:259-285 for a batch: planner [B,5] = (double_support_time,
        single_support_time, step_length, step_width, num_steps), current_time [B] -> contact [B,N,4], c_horizon [B,N,12]."""
        B, N = planner.shape[0], self.N
        pp = self._chk(planner, (B, 5), "planner"); pt = self._chk(current_time, (B,), "current_time")
        contact = torch.empty(B, N, 4, device=self.device, dtype=torch.float64)
        c_hor = torch.empty(B, N, 12, device=self.device, dtype=torch.float64)
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        with torch.cuda.device(self.device):
            rc = self._lib.g1mpc_plan_gait_batch(self._h, B, pp, pt, float(foot_offset), C.c_void_p(contact.data_ptr()),
                                                 C.c_void_p(c_hor.data_ptr()), stream)
        _lib.check(rc, "g1mpc_plan_gait_batch")
        return contact, c_hor

    def main_loop(self, planner, ticks, total_duration=4.0, foot_offset=0.08, warm_start=False, x_ref=None, record=True):
        """The SRBD-only loop of g1_mpc/scripts/main.py:200-326 for B walkers on the device (g1mpc_main_loop_batch).
        Returns dict(traj_x[B,ticks,13], traj_u[B,ticks,12], status_any[B], iters_sum[B], x_opt, u_opt, x_ref, current_time)."""
        B, N = planner.shape[0], self.N
        pp = self._chk(planner, (B, 5), "planner")
        mk = lambda *s, dt=torch.float64: torch.zeros(*s, device=self.device, dtype=dt)
        if x_ref is None:                                   # main.py:207-209
            x_ref = mk(B, N, NS); x_ref[:, :, 12] = self.params.get("gravity", GRAVITY); x_ref[:, :, 5] = 1.0
        pxr = self._chk(x_ref, (B, N, NS), "x_ref")
        x_opt, u_opt, t = mk(B, N, NS), mk(B, N, 12), mk(B)
        traj_x = mk(B, ticks, NS) if record else None
        traj_u = mk(B, ticks, 12) if record else None
        status_any, iters_sum = mk(B, dt=torch.int32), mk(B, dt=torch.int32)
        p = lambda a: None if a is None else C.c_void_p(a.data_ptr())
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        with torch.cuda.device(self.device):
            rc = self._lib.g1mpc_main_loop_batch(self._h, B, int(ticks), pp, float(total_duration), float(foot_offset), int(bool(warm_start)),
                                                 p(t), pxr, p(x_opt), p(u_opt), p(traj_x), p(traj_u), p(status_any), p(iters_sum), stream)
        _lib.check(rc, "g1mpc_main_loop_batch")
        return dict(traj_x=traj_x, traj_u=traj_u, status_any=status_any, iters_sum=iters_sum, x_opt=x_opt, u_opt=u_opt, x_ref=x_ref,
                    current_time=t)

    # -- the callers either side of the path ---------------------------------
    def walk_inputs(self, x0, contacts_now, sim_time, x_opt_prev, x_ref=None, want_landing=False):
        """What the demo node builds before updateMPC (mpc_to_wbid_node.cpp:385-470) for a batch of CUDA tensors.
        x_ref is in/out (None = zeros, as after initMatrices).  Returns x_ref, contact, c_horizon, p_com[, landing]."""
        B, N = x0.shape[0], self.N
        px0 = self._chk(x0, (B, NS), "x0"); pc = self._chk(contacts_now, (B, 12), "contacts_now")
        pt = self._chk(sim_time, (B,), "sim_time"); pxo = self._chk(x_opt_prev, (B, N, NS), "x_opt_prev")
        if x_ref is None:
            x_ref = torch.zeros(B, N, NS, device=self.device, dtype=torch.float64)
        pxr = self._chk(x_ref, (B, N, NS), "x_ref")
        mk = lambda *s: torch.empty(*s, device=self.device, dtype=torch.float64)
        contact, c_hor, p_com = mk(B, N, 4), mk(B, N, 12), mk(B, N, 3)
        landing = mk(B, 2) if want_landing else None
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        with torch.cuda.device(self.device):
            rc = self._lib.g1mpc_walk_inputs_batch(self._h, B, px0, pc, pt, pxo, pxr, C.c_void_p(contact.data_ptr()),
                                                   C.c_void_p(c_hor.data_ptr()), C.c_void_p(p_com.data_ptr()),
                                                   C.c_void_p(landing.data_ptr()) if want_landing else None, stream)
        _lib.check(rc, "g1mpc_walk_inputs_batch")
        return (x_ref, contact, c_hor, p_com, landing) if want_landing else (x_ref, contact, c_hor, p_com)

    def wbid_refs(self, u_opt, x_opt, foot_positions, wbid_mass):
        """WBID.setReference's derived references (wbid.py:244-297) for a batch -> [B,42] (layout: g1mpc_b200.h)."""
        B, N = u_opt.shape[0], self.N
        pu = self._chk(u_opt, (B, N, self.nu), "u_opt"); px = self._chk(x_opt, (B, N, NS), "x_opt")
        pf = self._chk(foot_positions, (B, 4, 3), "foot_positions")
        refs = torch.empty(B, 42, device=self.device, dtype=torch.float64)
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        with torch.cuda.device(self.device):
            rc = self._lib.g1mpc_wbid_refs_batch(self._h, B, pu, px, pf, float(wbid_mass), C.c_void_p(refs.data_ptr()), stream)
        _lib.check(rc, "g1mpc_wbid_refs_batch")
        return refs

    def closed_loop(self, x0, contacts_now, sim_time, ticks, x_ref=None, x_opt=None, record=True, warm_start=False):
        """`ticks` control periods of B walkers on the device (g1mpc_closed_loop_batch_ex).  x0, contacts_now,
        sim_time are advanced in place.  warm_start: every tick after the first starts from the previous tick's solution.
        Returns dict(traj_x[B,ticks,13], traj_u[B,ticks,12], status_any[B], iters_sum[B], u_opt, x_opt, x_ref)."""
        B, N = x0.shape[0], self.N
        px0 = self._chk(x0, (B, NS), "x0"); pc = self._chk(contacts_now, (B, 12), "contacts_now")
        pt = self._chk(sim_time, (B,), "sim_time")
        mk = lambda *s: torch.zeros(*s, device=self.device, dtype=torch.float64)
        x_ref = mk(B, N, NS) if x_ref is None else x_ref
        x_opt = mk(B, N, NS) if x_opt is None else x_opt
        pxr = self._chk(x_ref, (B, N, NS), "x_ref"); pxo = self._chk(x_opt, (B, N, NS), "x_opt")
        u_opt = mk(B, N, 12)
        traj_x = mk(B, ticks, NS) if record else None
        traj_u = mk(B, ticks, 12) if record else None
        status_any = torch.zeros(B, device=self.device, dtype=torch.int32)
        iters_sum = torch.zeros(B, device=self.device, dtype=torch.int32)
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        with torch.cuda.device(self.device), self._range("closed_loop_batch"):
            rc = self._lib.g1mpc_closed_loop_batch_ex(self._h, B, int(ticks), int(bool(warm_start)), px0, pc, pt, pxr, pxo,
                                                      C.c_void_p(u_opt.data_ptr()),
                                                      C.c_void_p(traj_x.data_ptr()) if record else None,
                                                      C.c_void_p(traj_u.data_ptr()) if record else None,
                                                      C.c_void_p(status_any.data_ptr()), C.c_void_p(iters_sum.data_ptr()), stream)
        _lib.check(rc, "g1mpc_closed_loop_batch_ex")
        return dict(traj_x=traj_x, traj_u=traj_u, status_any=status_any, iters_sum=iters_sum, u_opt=u_opt, x_opt=x_opt, x_ref=x_ref)

    # -- host API (the reference-facing call: numpy in, numpy out, copies inside) -----------------
    def update_host(self, x0, x_ref, contact, c_horizon, p_com, x_current=None, one_rollout=False, x_opt=None):
        f = lambda a, shape: np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(shape))
        x0 = np.asarray(x0, dtype=np.float64)
        B = 1 if x0.size == NS else x0.reshape(-1, NS).shape[0]
        N, nc, nu = self.N, self.nc, self.nu
        x0 = f(x0, (B, NS)); x_ref = f(x_ref, (B, N, NS)); contact = f(contact, (B, N, nc))
        c_horizon = f(c_horizon, (B, N, nu)); p_com = f(p_com, (B, N, 3))
        xc = None if x_current is None else f(x_current, (B, NS))
        u_opt = np.empty((B, N, nu)); status = np.empty(B, np.int32); iters = np.empty(B, np.int32)
        x_opt = np.zeros((B, N, NS)) if x_opt is None else f(x_opt, (B, N, NS)).copy()
        resid = np.empty((B, 4))
        p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
        with self._range("update_host"):
            rc = self._lib.g1mpc_update_host(self._h, B, p(x0), p(x_ref), p(contact), p(c_horizon), p(p_com), p(xc),
                                             int(one_rollout), p(u_opt), p(x_opt), p(status), p(iters), p(resid))
        _lib.check(rc, "g1mpc_update_host")
        return u_opt, x_opt, dict(status=status, iters=iters, resid=resid)

    HOST_SLOTS = 3      # G1MPC_HOST_SLOTS

    def pinned(self, *shape, dtype=np.float64):
        """numpy array in page-locked host memory (held by a torch tensor): the host API copies such buffers by DMA
        straight from / to where they lie instead of staging them (include/g1mpc_b200.h, g1mpc_update_host)."""
        t = torch.empty(*shape, dtype={np.float64: torch.float64, np.int32: torch.int32}[dtype], pin_memory=True)
        a = t.numpy()
        self._pinned_keep.append(t)
        return a

    def update_host_begin(self, slot, x0, x_ref, contact, c_horizon, p_com, x_current=None, one_rollout=False, out=None):
        """First half of update_host on `slot` (g1mpc_update_host_begin): enqueue and return.  The inputs must be
        C-contiguous float64 arrays of the batch shapes and stay untouched until update_host_end(slot).  `out`: dict
        of preallocated arrays u_opt, x_opt (in/out), status, iters, resid (e.g. from out_buffers); a fresh one if None."""
        B, N, nc, nu = x0.shape[0], self.N, self.nc, self.nu
        if out is None:
            out = self.out_buffers(B, pinned=False)
        for a, shape in ((x0, (B, NS)), (x_ref, (B, N, NS)), (contact, (B, N, nc)), (c_horizon, (B, N, nu)), (p_com, (B, N, 3)),
                         (out["u_opt"], (B, N, nu)), (out["x_opt"], (B, N, NS)), (out["resid"], (B, 4))):
            if a.dtype != np.float64 or not a.flags.c_contiguous or a.shape != shape:
                raise ValueError("update_host_begin needs C-contiguous float64 arrays of the batch shapes")
        if x_current is not None and (x_current.dtype != np.float64 or not x_current.flags.c_contiguous or x_current.shape != (B, NS)):
            raise ValueError("update_host_begin: x_current must be a C-contiguous float64 [B,13] array")
        p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
        rc = self._lib.g1mpc_update_host_begin(self._h, int(slot), B, p(x0), p(x_ref), p(contact), p(c_horizon), p(p_com),
                                               p(x_current), int(one_rollout), p(out["u_opt"]), p(out["x_opt"]),
                                               p(out["status"]), p(out["iters"]), p(out["resid"]))
        _lib.check(rc, "g1mpc_update_host_begin")
        self._inflight[int(slot)] = (out, (x0, x_ref, contact, c_horizon, p_com, x_current))     # keep the buffers alive
        return out

    def update_host_end(self, slot):
        """Second half: wait for `slot`; returns (u_opt, x_opt, info) like update_host."""
        _lib.check(self._lib.g1mpc_update_host_end(self._h, int(slot)), "g1mpc_update_host_end")
        out, _ = self._inflight.pop(int(slot))
        return out["u_opt"], out["x_opt"], dict(status=out["status"], iters=out["iters"], resid=out["resid"])

    def out_buffers(self, B, pinned=True):
        mk = (lambda *sh, dtype=np.float64: self.pinned(*sh, dtype=dtype)) if pinned else (lambda *sh, dtype=np.float64: np.zeros(sh, dtype))
        o = dict(u_opt=mk(B, self.N, self.nu), x_opt=mk(B, self.N, NS), status=mk(B, dtype=np.int32), iters=mk(B, dtype=np.int32),
                 resid=mk(B, 4))
        o["x_opt"][...] = 0.0
        return o


class MPC:
    """Drop-in for the reference `MPC` class (g1_mpc/scripts/mpc.py:10): same constructor signature, public
    attributes and step methods; all arithmetic of the path runs on the GPU through the C-ABI (batch of one).

    * `update()` / `solve_qp()` run the fused solve kernel on the current inputs (x0, x_ref_hor and the arguments).
    * The reference leaves every intermediate current after `update()` (mpc.py:360-377): here `update()` and every
      `calculate_*`/`set_*` method take a snapshot of the inputs and mark the intermediates stale; the first read of
      `A_continuous, A_discrete, B_continuous(_hor), B_discrete(_hor), Aqp, Bqp, Ac, Hessian, gradient,
      lower/upper_bounds_horizon, rotation_z_T, Q, R, psi, r` refreshes ALL of them from ONE emit-kernel launch on that
      snapshot (a caller that never reads them never pays for them).  Call sequences such as
      g1_mpc/scripts/main.py:290-310 and g1_mujoco_sim/src/run_simulation.py:96-106 work unchanged.
    * `compute_rollout()` rolls out the CURRENT `self.u_opt` (mpc.py:379-386) with g1mpc_rollout_batch -- no solve; the
      reference's debug prints of the full rollout (mpc.py:389-391) are not reproduced.
    * `cost_scale`, `INERTIA_BODY`, `Q_WEIGHTS`, `R_WEIGHTS`, `NUM_CONTACTS` may be overridden before `init_matrices()`
      exactly like on the reference object.  Q_WEIGHTS must be diagonal and R_WEIGHTS a uniform diagonal
      (what both reference variants ship, mpc.py:47-51, mpc.cpp:90-98): anything else raises instead of being dropped.
    * `warm_start` (default False = the reference's fresh solver per call): True starts each update() from the previous
      solution shifted by one stage -- what setWarmStart(true) (mpc.cpp:386) / warm_start=True (mpc.py:342) ask for."""

    _EMIT = {"A_discrete": "A", "B_discrete_hor": "Bd", "Aqp": "Aqp", "Bqp": "Bqp", "Hessian": "H", "gradient": "g",
             "Ac": "Acon", "lower_bounds_horizon": "l", "upper_bounds_horizon": "u"}
    _DERIVED = ("A_continuous", "B_continuous_hor", "B_continuous", "B_discrete", "rotation_z_T", "Q", "R", "r")

    def __init__(self, mu=0.3, fz_min=10., fz_max=666., dt=0.04, HORIZON_LENGTH=15, *, preset="python_default",
                 device=None):
        base = get_preset(preset)
        self.g = GRAVITY
        self.ROBOT_MASS = ROBOT_MASS
        self.dt = dt
        self.NUM_CONTACTS = 4
        self.HORIZON_LENGTH = HORIZON_LENGTH
        self.NUM_STATES = 13
        self.NUM_CONTROLS = 3 * self.NUM_CONTACTS
        self.NUM_BOUNDS = 5
        self.INERTIA_BODY = np.asarray(base["inertia_body"], float).reshape(3, 3)
        self.x = np.zeros((self.NUM_STATES, 1))
        self.u = np.zeros((self.NUM_CONTROLS, 1))
        self.Q_WEIGHTS = np.diag(np.asarray(base["q_weights"], float))
        self.R_WEIGHTS = np.diag(np.full(self.NUM_CONTROLS, base["r_weight"]))
        self.cost_scale = base["cost_scale"]
        self.mu, self.fz_min, self.fz_max = mu, fz_min, fz_max
        self.warm_start = False
        self._device = device
        self._solver = None
        self._vals, self._snap, self._stale = {}, None, True
        self._warm_on = False

    @classmethod
    def cpp_node(cls, device=None):
        """The constants of the shipped C++ demo (mpc.hpp:22, mpc.cpp:25-27,90-98, cost x2)."""
        p = get_preset("cpp_node")
        return cls(p["mu"], p["fz_min"], p["fz_max"], p["dt"], p["horizon"], preset="cpp_node", device=device)

    # mpc.py:80-107
    def init_matrices(self):
        N, ns, nu = self.HORIZON_LENGTH, self.NUM_STATES, 3 * self.NUM_CONTACTS
        self.NUM_CONTROLS = nu
        z = np.zeros
        self._vals = dict(A_continuous=z((ns, ns)), A_discrete=z((ns, ns)), Aqp=z((ns * N, ns)), B_continuous=z((ns, nu)),
                          B_continuous_hor=z((ns * N, nu)), B_discrete=z((ns, nu)), B_discrete_hor=z((ns * N, nu)),
                          Bqp=z((ns * N, nu * N)), Q=z((ns * N, ns * N)), R=z((nu * N, nu * N)),
                          Ac=z((5 * self.NUM_CONTACTS * N, nu * N)), Hessian=z((nu * N, nu * N)), gradient=z((nu * N, 1)),
                          rotation_z_T=z((3, 3)), lower_bounds_horizon=z(5 * self.NUM_CONTACTS * N),
                          upper_bounds_horizon=z(5 * self.NUM_CONTACTS * N), r=z(3))
        self._snap, self._stale = None, False               # as after the reference's init_matrices: all zeros
        self.c_horizon, self.p_com_horizon = z((N, self.NUM_CONTACTS * 3)), z((N, 3))
        self.x0, self.x_ref_hor = z((ns, 1)), z((N, ns))
        self.u_ref_hor, self.u_opt, self.x_opt, self.u_opt0 = z((N, nu)), z((N, nu)), z((N, ns)), z((nu, 1))
        self._contact = np.ones((N, self.NUM_CONTACTS))
        self.psi = 0.0
        qw = np.asarray(self.Q_WEIGHTS, float)
        if qw.shape != (ns, ns) or np.count_nonzero(qw - np.diag(np.diag(qw))):
            raise ValueError("Q_WEIGHTS must be a diagonal 13x13 matrix (both reference variants ship a diagonal, mpc.py:47-50, "
                             "mpc.cpp:90-96): off-diagonal state weights are not supported by the stage-wise solver")
        rw = np.asarray(self.R_WEIGHTS, float)
        rdiag = np.diag(rw)
        if rw.shape != (nu, nu) or np.count_nonzero(rw - np.diag(rdiag)) or not np.allclose(rdiag, rdiag[0]):
            raise ValueError("R_WEIGHTS must be a uniform diagonal (the reference ships 0.001 I, mpc.py:51)")
        params = dict(horizon=N, num_contacts=self.NUM_CONTACTS, mu=self.mu, fz_min=self.fz_min, fz_max=self.fz_max,
                      dt=self.dt, mass=self.ROBOT_MASS, gravity=self.g, inertia_body=np.asarray(self.INERTIA_BODY).ravel(),
                      q_weights=np.diag(qw), r_weight=float(rdiag[0]), cost_scale=self.cost_scale)
        if self._solver is not None:
            self._solver.close()
        self._solver = BatchedMPC(params, device=self._device)      # raises with the supported range for a shape that does not fit
        self._warm_on = False
        # persistent host buffers of the single-problem call (no allocation per update)
        self._hb = dict(x0=z((1, ns)), xr=z((1, N, ns)), ct=z((1, N, self.NUM_CONTACTS)), ch=z((1, N, nu)), pc=z((1, N, 3)),
                        xc=z((1, ns)), u=z((1, N, nu)), xo=z((1, N, ns)), res=z((1, 4)), st=np.zeros(1, np.int32), it=np.zeros(1, np.int32))
        self._hp = {k: v.ctypes.data_as(C.c_void_p) for k, v in self._hb.items()}

    # ---- internals -------------------------------------------------------------------------------
    def _need(self):
        if self._solver is None:
            self.init_matrices()
        return self._solver

    def _snapshot(self):
        """Inputs the intermediates belong to (taken by update() and by every calculate_*/set_* method)."""
        N, nc = self.HORIZON_LENGTH, self.NUM_CONTACTS
        f = lambda a, shape: np.array(a, dtype=np.float64).reshape(shape)
        self._snap = (f(self.x0, (1, 13)), f(self.x_ref_hor, (1, N, 13)), f(self._contact, (1, N, nc)),
                      f(self.c_horizon, (1, N, 3 * nc)), f(self.p_com_horizon, (1, N, 3)))
        self._stale = True

    def _refresh(self):
        """One emit launch on the snapshot -> every intermediate the reference keeps as an attribute."""
        s = self._need()
        if self._snap is None:
            self._snapshot()
        dev, N, nc, dt = s.device, self.HORIZON_LENGTH, self.NUM_CONTACTS, self.dt
        tens = [torch.as_tensor(a).to(dev) for a in self._snap]
        out = {k: v[0].cpu().numpy() for k, v in s.emit_qp(*tens).items()}
        v = self._vals
        for attr, key in self._EMIT.items():
            v[attr] = out[key]
        from scipy import sparse
        v["Ac"] = sparse.csr_matrix(out["Acon"])                          # mpc.py:256 (block_diag, csr)
        v["gradient"] = out["g"].reshape(-1, 1)                           # mpc.py:317 (column vector)
        A = out["A"]
        v["A_continuous"] = (A - np.eye(13)) / dt                         # mpc.py:139-156
        v["rotation_z_T"] = (A[0:3, 6:9] / dt).T                          # A(0:3,6:9) = dt rotation_z_T' (mpc.py:142,155)
        v["B_continuous_hor"] = out["Bd"] / dt                            # mpc.py:188-192
        v["B_continuous"] = v["B_continuous_hor"][-self.NUM_STATES:].copy()
        v["B_discrete"] = out["Bd"][-self.NUM_STATES:].copy()
        v["Q"] = np.kron(np.eye(N), np.asarray(self.Q_WEIGHTS, float))    # mpc.py:130-138
        v["R"] = np.kron(np.eye(N), np.asarray(self.R_WEIGHTS, float))
        ch, pc = self._snap[3][0], self._snap[4][0]
        v["r"] = ch[N - 1, 3 * (nc - 1):3 * nc] - pc[N - 1]               # the last lever arm of the double loop (mpc.py:176)
        self._stale = False

    def __getattr__(self, name):
        # only reached when normal lookup fails: the lazily refreshed intermediates live in _vals
        if name in MPC._EMIT or name in MPC._DERIVED:
            vals = self.__dict__.get("_vals")
            if vals is None or name not in vals:
                raise AttributeError(f"{name}: call init_matrices() first (as on the reference object)")
            if self.__dict__.get("_stale"):
                self._refresh()
            return self._vals[name]
        raise AttributeError(name)

    def __setattr__(self, name, value):
        if name in MPC._EMIT or name in MPC._DERIVED:
            self.__dict__.setdefault("_vals", {})[name] = value           # a caller may overwrite an intermediate, as on the reference
        else:
            object.__setattr__(self, name, value)

    # ---- the reference's step methods (mpc.py:109-321) -------------------------------------------
    def extract_psi(self):
        xr = np.asarray(self.x_ref_hor, float)
        self.psi = float(xr[:, 2].sum() / xr.shape[0])          # = np.mean (same pairwise sum), without its call overhead
        return self.psi

    def _contact_array(self, contact):
        """contact[step][leg] of the reference (lists of arrays accepted, mpc.py:264) as a float [N, nc] array."""
        N, nc = self.HORIZON_LENGTH, self.NUM_CONTACTS
        if isinstance(contact, np.ndarray) and contact.shape == (N, nc):
            return contact.astype(np.float64)
        return np.array([[float(contact[i][j]) for j in range(nc)] for i in range(N)])

    def vector_to_skew_symmetric_matrix(self, v):
        return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]])

    def rotation_matrix_T(self):
        self._snapshot()

    def set_Q(self):
        self._snapshot()

    def set_R(self):
        self._snapshot()

    def calculate_A_continuous(self):
        self._snapshot()

    def calculate_A_discrete(self):
        self._snapshot()

    def calculate_B_continuous(self, c_horizon, p_com_horizon):
        self.c_horizon = np.array(c_horizon, float)
        self.p_com_horizon = np.array(p_com_horizon, float)
        self._snapshot()

    def calculate_B_discrete(self):
        self._snapshot()

    def calculate_Aqp(self):
        self._snapshot()

    def calculate_Bqp(self):
        self._snapshot()

    def calculate_Ac(self):
        self._snapshot()

    def calculate_bounds(self, contact):
        self._contact = self._contact_array(contact)
        self._snapshot()

    def calculate_hessian(self):
        self._snapshot()

    def calculate_gradient(self):
        self.x_ref_hor_vec = np.asarray(self.x_ref_hor, float).reshape((self.NUM_STATES * self.HORIZON_LENGTH, 1))
        self.x0 = np.asarray(self.x0, float).reshape((self.NUM_STATES, 1))
        self._snapshot()

    class _Result:
        pass

    def _solve(self, x_current, one_rollout, do_rollout):
        s = self._need()
        hb, hp = self._hb, self._hp
        hb["x0"][...] = np.asarray(self.x0, float).reshape(1, 13)
        hb["xr"][0] = self.x_ref_hor
        hb["ct"][0] = self._contact
        hb["ch"][0] = self.c_horizon
        hb["pc"][0] = self.p_com_horizon
        if x_current is not None:
            hb["xc"][0] = x_current
        hb["xo"][0] = self.x_opt if do_rollout else 0.0
        if bool(self.warm_start) != self._warm_on:
            s.set_host_warm_start(bool(self.warm_start), 1)
            self._warm_on = bool(self.warm_start)
        rc = s._lib.g1mpc_update_host(s._h, 1, hp["x0"], hp["xr"], hp["ct"], hp["ch"], hp["pc"],
                                      hp["xc"] if x_current is not None else None, int(one_rollout), hp["u"], hp["xo"],
                                      hp["st"], hp["it"], hp["res"])
        if rc:
            _lib.check(rc, "g1mpc_update_host")
        self.u_opt = hb["u"][0].copy()
        self.u_opt0 = self.u_opt[0].reshape(-1, 1)
        if do_rollout:
            self.x_opt = hb["xo"][0].copy()
        r = MPC._Result()
        r.x = self.u_opt.ravel().copy()
        r.info = dict(status=int(hb["st"][0]), iters=int(hb["it"][0]), resid=hb["res"][0].copy())
        self.qp_info = r.info
        return r

    def solve_qp(self):
        """mpc.py:323-358 -- returns an object with `.x` like the osqp result the reference returns; sets u_opt only."""
        return self._solve(None, True, False)

    def update(self, contact, c_horizon, p_com_horizon, x_current=None, one_rollout=False):
        """mpc.py:360-377 -- one fused kernel launch; returns (u_opt[N,nu], x_opt[N,13])."""
        self.c_horizon = np.array(c_horizon, float)
        self.p_com_horizon = np.array(p_com_horizon, float)
        self._contact = self._contact_array(contact)
        self.extract_psi()
        xc = None if x_current is None else np.asarray(x_current, float).reshape(self.NUM_STATES)
        self._solve(xc, one_rollout, True)
        hb = self._hb                                      # the intermediates of THIS call, refreshed on first read
        self._snap = (hb["x0"].copy(), hb["xr"].copy(), hb["ct"].copy(), hb["ch"].copy(), hb["pc"].copy())
        self._stale = True
        return self.u_opt, self.x_opt

    def compute_rollout(self, x_current=None, only_first_step=False):
        """mpc.py:379-391 -- rollout of the CURRENT self.u_opt from x_current (or from x_opt[0]) with the A, B_i of the last
        update()/calculate_* snapshot (g1mpc_rollout_batch; no solve)."""
        s = self._need()
        if self._snap is None:
            self._snapshot()
        N, nu = self.HORIZON_LENGTH, self.NUM_CONTROLS
        f = lambda a, shape: np.ascontiguousarray(np.array(a, dtype=np.float64).reshape(shape))
        u, xo = f(self.u_opt, (1, N, nu)), f(self.x_opt, (1, N, 13))
        xc = None if x_current is None else f(x_current, (1, 13))
        p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
        rc = s._lib.g1mpc_rollout_host(s._h, 1, p(self._snap[1]), p(self._snap[3]), p(self._snap[4]), p(u), p(xc),
                                       int(bool(only_first_step)), p(xo))
        _lib.check(rc, "g1mpc_rollout_host")
        self.x_opt = xo[0]
