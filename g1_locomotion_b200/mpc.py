"""Host-side mirror of the reference MPC interface over the B200 C-ABI.

`BatchedMPC`  -- the batched solver: update_batch() on CUDA float64 tensors (device pointers go
                 straight into g1mpc_solve_batch), update_host() on numpy arrays (H2D/D2H inside the call).
`MPC`         -- drop-in for the reference Python class g1_mpc/scripts/mpc.py:10-391 (same constructor,
                 attributes and step-method names; one problem = batch of 1).  The C++ class
                 g1_mpc::MPC (mpc.hpp:15-199) maps onto the same calls.
There is no CPU path: every numeric result comes from the CUDA kernels.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .presets import get_preset, GRAVITY, ROBOT_MASS

NS = 13


def _c_params(p, gap_tol=0.0, max_iter=0):
    cp = _lib.G1mpcParams()
    cp.horizon, cp.num_contacts = int(p["horizon"]), int(p["num_contacts"])
    cp.mu, cp.fz_min, cp.fz_max, cp.dt = float(p["mu"]), float(p["fz_min"]), float(p["fz_max"]), float(p["dt"])
    cp.mass, cp.gravity = float(p["mass"]), float(p.get("gravity", GRAVITY))
    cp.inertia_body[:] = [float(v) for v in np.asarray(p["inertia_body"], float).ravel()]
    cp.q_weights[:] = [float(v) for v in np.asarray(p["q_weights"], float).ravel()]
    cp.r_weight, cp.cost_scale = float(p["r_weight"]), float(p["cost_scale"])
    cp.gap_tol, cp.max_iter, cp.reserved = float(gap_tol), int(max_iter), 0
    return cp


class BatchedMPC:
    """Batch of independent SRBD-MPC problems sharing one parameter set, on one CUDA device."""

    def __init__(self, preset="cpp_node", device=None, gap_tol=0.0, max_iter=0, **overrides):
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
        cp = _c_params(self.params, gap_tol, max_iter)
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
        return dict(smem_bytes=a.value, threads=b.value, ctas_per_sm=c.value, problems_per_warp=ppw.value)

    # -- device API ------------------------------------------------------------------------------
    def update_batch(self, x0, x_ref, contact, c_horizon, p_com, x_current=None, *, one_rollout=False,
                     u_opt=None, x_opt=None, want_resid=True, want_multipliers=False):
        """MPC.update over a batch of CUDA tensors (mpc.py:360-377 / mpc.cpp:448-473).

        Returns (u_opt[B,N,nu], x_opt[B,N,13], info) with info = {status[B], iters[B], resid[B,4], y[B,5ncN]?}.
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
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        with torch.cuda.device(self.device):
            rc = self._lib.g1mpc_solve_batch(self._h, B, px0, pxr, pct, pch, ppc, pxc, int(one_rollout), pu, pxo,
                                             C.c_void_p(status.data_ptr()), C.c_void_p(iters.data_ptr()),
                                             C.c_void_p(resid.data_ptr()) if want_resid else None,
                                             C.c_void_p(y.data_ptr()) if want_multipliers else None, stream)
        _lib.check(rc, "g1mpc_solve_batch")
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
        with torch.cuda.device(self.device):
            rc = self._lib.g1mpc_emit_qp_batch(self._h, B, px0, pxr, pct, pch, ppc, ptr("A"), ptr("Bd"), ptr("Aqp"),
                                               ptr("Bqp"), ptr("H"), ptr("g"), ptr("Acon"), ptr("l"), ptr("u"), stream)
        _lib.check(rc, "g1mpc_emit_qp_batch")
        return out

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

    def closed_loop(self, x0, contacts_now, sim_time, ticks, x_ref=None, x_opt=None, record=True):
        """`ticks` control periods of B walkers on the device (g1mpc_closed_loop_batch).  x0, contacts_now,
        sim_time are advanced in place.  Returns dict(traj_x[B,ticks,13], traj_u[B,ticks,12], status_any[B], u_opt, x_opt, x_ref)."""
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
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        with torch.cuda.device(self.device):
            rc = self._lib.g1mpc_closed_loop_batch(self._h, B, int(ticks), px0, pc, pt, pxr, pxo, C.c_void_p(u_opt.data_ptr()),
                                                   C.c_void_p(traj_x.data_ptr()) if record else None,
                                                   C.c_void_p(traj_u.data_ptr()) if record else None,
                                                   C.c_void_p(status_any.data_ptr()), stream)
        _lib.check(rc, "g1mpc_closed_loop_batch")
        return dict(traj_x=traj_x, traj_u=traj_u, status_any=status_any, u_opt=u_opt, x_opt=x_opt, x_ref=x_ref)

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
    attributes and step methods; all arithmetic runs on the GPU through the C-ABI (batch of one).

    The reference recomputes each intermediate in its own method (mpc.py:109-321).  Here every
    `calculate_*`/`set_*` method refreshes its attribute(s) from one emit-kernel launch on the current
    inputs, `solve_qp()`/`update()` run the fused solve kernel, so call sequences such as
    g1_mpc/scripts/main.py:290-310 and g1_mujoco_sim/src/run_simulation.py:96-106 work unchanged.
    `cost_scale`, `INERTIA_BODY`, `Q_WEIGHTS`, `R_WEIGHTS`, `NUM_CONTACTS` may be overridden before
    `init_matrices()` exactly like on the reference object."""

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
        self._device = device
        self._solver = None
        self.psi = 0.0

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
        self.A_continuous, self.A_discrete, self.Aqp = z((ns, ns)), z((ns, ns)), z((ns * N, ns))
        self.B_continuous, self.B_continuous_hor = z((ns, nu)), z((ns * N, nu))
        self.B_discrete, self.B_discrete_hor, self.Bqp = z((ns, nu)), z((ns * N, nu)), z((ns * N, nu * N))
        self.Q, self.R = z((ns * N, ns * N)), z((nu * N, nu * N))
        self.c_horizon, self.p_com_horizon = z((N, self.NUM_CONTACTS * 3)), z((N, 3))
        self.Ac = z((5 * self.NUM_CONTACTS * N, nu * N))
        self.Hessian, self.gradient = z((nu * N, nu * N)), z((nu * N, 1))
        self.rotation_z_T = z((3, 3))
        self.x0, self.x_ref_hor = z((ns, 1)), z((N, ns))
        self.u_ref_hor, self.u_opt, self.x_opt, self.u_opt0 = z((N, nu)), z((N, nu)), z((N, ns)), z((nu, 1))
        self._contact = np.ones((N, self.NUM_CONTACTS))
        self.lower_bounds_horizon = z(5 * self.NUM_CONTACTS * N)
        self.upper_bounds_horizon = z(5 * self.NUM_CONTACTS * N)
        rdiag = np.diag(np.asarray(self.R_WEIGHTS, float))
        if not np.allclose(rdiag, rdiag[0]):
            raise ValueError("R_WEIGHTS must be a uniform diagonal (the reference ships 0.001 I, mpc.py:51)")
        params = dict(horizon=N, num_contacts=self.NUM_CONTACTS, mu=self.mu, fz_min=self.fz_min, fz_max=self.fz_max,
                      dt=self.dt, mass=self.ROBOT_MASS, gravity=self.g, inertia_body=np.asarray(self.INERTIA_BODY).ravel(),
                      q_weights=np.diag(np.asarray(self.Q_WEIGHTS, float)), r_weight=float(rdiag[0]),
                      cost_scale=self.cost_scale)
        if self._solver is not None:
            self._solver.close()
        self._solver = BatchedMPC(params, device=self._device)

    # ---- internals -------------------------------------------------------------------------------
    def _need(self):
        if self._solver is None:
            self.init_matrices()
        return self._solver

    def _emit(self, which):
        s = self._need()
        dev = s.device
        t = lambda a, shape: torch.as_tensor(np.asarray(a, float).reshape(shape), dtype=torch.float64).contiguous().to(dev)
        N, nc = self.HORIZON_LENGTH, self.NUM_CONTACTS
        out = s.emit_qp(t(self.x0, (1, 13)), t(self.x_ref_hor, (1, N, 13)), t(self._contact, (1, N, nc)),
                        t(self.c_horizon, (1, N, 3 * nc)), t(self.p_com_horizon, (1, N, 3)), which=which)
        return {k: v[0].cpu().numpy() for k, v in out.items()}

    # ---- the reference's step methods (mpc.py:109-321) -------------------------------------------
    def extract_psi(self):
        self.psi = float(np.mean(np.asarray(self.x_ref_hor, float)[:, 2]))
        return self.psi

    def vector_to_skew_symmetric_matrix(self, v):
        return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]])

    def rotation_matrix_T(self):
        A = self._emit(("A",))["A"]                      # A(0:3,6:9) = dt * rotation_z_T' (mpc.py:142,155)
        self.rotation_z_T = (A[0:3, 6:9] / self.dt).T

    def set_Q(self):
        N, ns = self.HORIZON_LENGTH, self.NUM_STATES
        self.Q = np.kron(np.eye(N), np.asarray(self.Q_WEIGHTS, float))

    def set_R(self):
        self.R = np.kron(np.eye(self.HORIZON_LENGTH), np.asarray(self.R_WEIGHTS, float))

    def calculate_A_continuous(self):
        A = self._emit(("A",))["A"]
        self.A_continuous = (A - np.eye(13)) / self.dt

    def calculate_A_discrete(self):
        self.A_discrete = self._emit(("A",))["A"]

    def calculate_B_continuous(self, c_horizon, p_com_horizon):
        self.c_horizon = np.array(c_horizon, float)
        self.p_com_horizon = np.array(p_com_horizon, float)
        Bd = self._emit(("Bd",))["Bd"]
        self.B_continuous_hor = Bd / self.dt
        self.B_continuous = self.B_continuous_hor[-self.NUM_STATES:].copy()

    def calculate_B_discrete(self):
        self.B_discrete_hor = self._emit(("Bd",))["Bd"]
        self.B_discrete = self.B_discrete_hor[-self.NUM_STATES:].copy()

    def calculate_Aqp(self):
        self.Aqp = self._emit(("Aqp",))["Aqp"]

    def calculate_Bqp(self):
        self.Bqp = self._emit(("Bqp",))["Bqp"]

    def calculate_Ac(self):
        from scipy import sparse
        self.Ac = sparse.csr_matrix(self._emit(("Acon",))["Acon"])

    def calculate_bounds(self, contact):
        self._contact = np.array([[float(contact[i][j]) for j in range(self.NUM_CONTACTS)]
                                  for i in range(self.HORIZON_LENGTH)])
        o = self._emit(("l", "u"))
        self.lower_bounds_horizon, self.upper_bounds_horizon = o["l"], o["u"]

    def calculate_hessian(self):
        self.Hessian = self._emit(("H",))["H"]

    def calculate_gradient(self):
        self.x_ref_hor_vec = np.asarray(self.x_ref_hor, float).reshape((self.NUM_STATES * self.HORIZON_LENGTH, 1))
        self.x0 = np.asarray(self.x0, float).reshape((self.NUM_STATES, 1))
        self.gradient = self._emit(("g",))["g"].reshape(-1, 1)

    class _Result:
        pass

    def _solve(self, x_current, one_rollout, do_rollout):
        s = self._need()
        xo = self.x_opt if do_rollout else np.zeros_like(self.x_opt)
        u, x, info = s.update_host(self.x0, self.x_ref_hor, self._contact, self.c_horizon, self.p_com_horizon,
                                   x_current=x_current, one_rollout=one_rollout, x_opt=xo)
        self.u_opt = u[0]
        self.u_opt0 = self.u_opt[0].reshape(-1, 1)
        if do_rollout:
            self.x_opt = x[0]
        r = MPC._Result()
        r.x = self.u_opt.ravel().copy()
        r.info = {k: v[0] for k, v in info.items()}
        self.qp_info = r.info
        return r

    def solve_qp(self):
        """mpc.py:323-358 -- returns an object with `.x` like the osqp result the reference returns."""
        return self._solve(None, True, False)

    def update(self, contact, c_horizon, p_com_horizon, x_current=None, one_rollout=False):
        """mpc.py:360-377 -- one fused kernel launch; returns (u_opt[N,nu], x_opt[N,13])."""
        self.c_horizon = np.array(c_horizon, float)
        self.p_com_horizon = np.array(p_com_horizon, float)
        self._contact = np.array([[float(contact[i][j]) for j in range(self.NUM_CONTACTS)]
                                  for i in range(self.HORIZON_LENGTH)])
        self.extract_psi()
        xc = None if x_current is None else np.asarray(x_current, float).reshape(self.NUM_STATES)
        self._solve(xc, one_rollout, True)
        return self.u_opt, self.x_opt

    def compute_rollout(self, x_current=None, only_first_step=False):
        """mpc.py:379-391 -- rollout of the current u_opt (re-solves on the GPU; the QP is deterministic)."""
        xc = None if x_current is None else np.asarray(x_current, float).reshape(self.NUM_STATES)
        self._solve(xc, only_first_step, True)
