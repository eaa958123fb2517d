"""numpy model of the solve kernel's algorithm (g1_locomotion_b200/csrc/g1mpc_kernels.cu), one problem at a time.

TEST INFRASTRUCTURE: lets the CPU suite check the *mathematics* the CUDA kernel relies on (wrench-space
factorisation of the reference Hessian, stage-wise Riccati solution of the interior-point Newton systems)
against the oracle without a GPU.  Never imported by the product.

Reference maths: g1_mpc/src/mpc.cpp:153-446.
"""
import numpy as np


def skew(v):
    return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0.0]])


class Problem:
    """Wrench-space form of one MPC problem: H = sr I + W' M W, g = W' h  (kernel header comment)."""

    def __init__(self, p, x0, xref, contact, c_hor, p_com):
        N, nc = p["horizon"], p["num_contacts"]
        self.N, self.nc, self.nu = N, nc, 3 * nc
        dt, s = p["dt"], p["cost_scale"]
        q = np.asarray(p["q_weights"], float)
        self.mu, self.sr, self.bv, self.dt, self.s, self.q = p["mu"], s * p["r_weight"], dt / p["mass"], dt, s, q
        psi = float(np.mean(xref[:, 2]))                                  # mpc.cpp:153-162
        c, sn = np.cos(psi), np.sin(psi)
        T = np.array([[c, sn, 0], [-sn, c, 0], [0, 0, 1.0]])              # mpc.cpp:175-177
        self.Rz = T.T
        Iwi = np.linalg.inv(T @ np.asarray(p["inertia_body"], float).reshape(3, 3) @ T.T)   # mpc.cpp:220-221
        self.Bw = np.zeros((N, nc, 3, 3))
        for k in range(N):
            for j in range(nc):
                self.Bw[k, j] = dt * Iwi @ skew(c_hor[k, 3 * j:3 * j + 3] - p_com[k])      # mpc.cpp:229-237,251
        self.lo = p["fz_min"] * contact                                   # mpc.cpp:344
        self.hi = np.maximum(p["fz_max"] * contact, self.lo)              # mpc.cpp:350
        self.stance = self.hi > self.lo
        self.weight = p["mass"] * abs(p.get("gravity", -9.80665))
        # stage dynamics on the 12 weighted states y = (theta, p, omega, v):  y+ = Abar y + Ebar w (+ gravity)
        self.Phi = np.zeros((6, 6))
        self.Phi[0:3, 0:3] = dt * self.Rz                                 # mpc.cpp:199,210 (quirk Q2)
        self.Phi[3:6, 3:6] = dt * np.eye(3)
        self.Abar = np.eye(12)
        self.Abar[0:6, 6:12] = self.Phi
        # free response error e_i = A^(i+1) x0 - xref_i (12 weighted states)  mpc.cpp:256-270,369-375
        E = np.zeros((N, 12))
        g0 = x0[12]
        for i in range(N):
            kk = i + 1
            E[i, 0:3] = x0[0:3] + kk * dt * self.Rz @ x0[6:9] - xref[i, 0:3]
            E[i, 3:6] = x0[3:6] + kk * dt * x0[9:12] - xref[i, 3:6]
            E[i, 5] += 0.5 * kk * (kk - 1) * dt * dt * g0
            E[i, 6:9] = x0[6:9] - xref[i, 6:9]
            E[i, 9:12] = x0[9:12] - xref[i, 9:12]
            E[i, 11] += kk * dt * g0
        self.E = E

    def W(self, k):
        Wk = np.zeros((6, self.nu))
        for j in range(self.nc):
            Wk[0:3, 3 * j:3 * j + 3] = self.Bw[k, j]
            Wk[3:6, 3 * j:3 * j + 3] = self.bv * np.eye(3)
        return Wk

    def Wfull(self):
        N, nu = self.N, self.nu
        Wf = np.zeros((6 * N, nu * N))
        for k in range(N):
            Wf[6 * k:6 * k + 6, nu * k:nu * k + nu] = self.W(k)
        return Wf

    def G(self):
        """wrench sequence -> stacked states y_1..y_N (12N x 6N)."""
        N = self.N
        G = np.zeros((12 * N, 6 * N))
        for j in range(N):
            col = np.zeros((12, 6)); col[6:12] = np.eye(6)
            for i in range(j, N):
                G[12 * i:12 * i + 12, 6 * j:6 * j + 6] = col
                col = self.Abar @ col
        return G

    def M_h(self):
        G = self.G()
        Qd = np.tile(self.s * self.q[:12], self.N)
        return G.T @ (Qd[:, None] * G), G.T @ (Qd * self.E.ravel())

    def dense_H_g(self):
        M, h = self.M_h()
        Wf = self.Wfull()
        return self.sr * np.eye(self.nu * self.N) + Wf.T @ M @ Wf, Wf.T @ h


# ------------------------------------------------------------------------------------------------------
# stage-wise (Riccati) solution of  (M^-1 + blockdiag(S_k)) y = t   -- the interior-point Newton system in
# wrench space.  y_k is the wrench part of the co-state; S_k = W_k Dblk_k^-1 W_k' may be singular.
# ------------------------------------------------------------------------------------------------------
class RiccatiFactor:
    def __init__(self, prob, S):
        """S: [N,6,6].  Backward sweep; stores per stage k (control stage 0..N-1) what the solves need."""
        N = prob.N
        self.prob, self.S = prob, S
        sQ = np.diag(prob.s * prob.q[:12])
        P = sQ.copy()                                   # P_N
        self.L, self.C, self.P2, self.P22 = [None] * N, [None] * N, [None] * N, [None] * N
        A = prob.Abar
        for k in range(N - 1, -1, -1):
            P22 = P[6:12, 6:12]
            L = np.linalg.cholesky(P22)
            K = np.eye(6) + L.T @ S[k] @ L
            C = np.linalg.cholesky(K)
            self.L[k], self.C[k], self.P2[k], self.P22[k] = L, C, P[6:12, :].copy(), P22.copy()
            if k == 0:
                break
            Z = np.linalg.solve(L, P[6:12, :])          # 6x12 ; Z[:,6:12] = L'
            Y = np.linalg.solve(C, Z)
            Pt = Y.T @ Y
            Pt[0:6, 0:6] += P[0:6, 0:6] - Z[:, 0:6].T @ Z[:, 0:6]
            P = sQ + A.T @ Pt @ A

    def _zeta(self, k, rhs6):
        """(I + P22 S)^-1 rhs6 = L K^-1 L^-1 rhs6."""
        L, C = self.L[k], self.C[k]
        a = np.linalg.solve(L, rhs6)
        a = np.linalg.solve(C.T, np.linalg.solve(C, a))
        return L @ a

    def solve(self, t):
        """t: [N,6] -> y [N,6] with (M^-1 + S) y = t."""
        N, A = self.prob.N, self.prob.Abar
        p = np.zeros(12)
        ps = [None] * N
        for k in range(N - 1, -1, -1):                  # backward: linear term of the value function
            ps[k] = p.copy()                            # p_{k+1}
            z0 = self._zeta(k, p[6:12] + self.P22[k] @ t[k])
            dw0 = t[k] - self.S[k] @ z0
            p = A.T @ (p + self.P2[k].T @ dw0)
        y = np.zeros((N, 6))
        dy = np.zeros(12)
        for k in range(N):                              # forward
            Ady = A @ dy
            eta = self.P2[k] @ Ady + ps[k][6:12]
            y[k] = self._zeta(k, eta + self.P22[k] @ t[k])
            dw = t[k] - self.S[k] @ y[k]
            dy = Ady
            dy[6:12] += dw
        return y


# ------------------------------------------------------------------------------------------------------
# Mehrotra predictor-corrector interior point, same steps / same closed forms as the kernel
# ------------------------------------------------------------------------------------------------------
def _contact_blocks(D, sr, mu):
    """3x3 inverse of Dblk = sr I + Cb' D Cb (as 6 symmetric entries) and X = Dblk^-1 Cb' (6 rows x 3),
    in the cancellation-free closed forms of the kernel.  Cb rows: fx-mu fz<=0, -fx-mu fz<=0, fy.., -fy.., fz<=hi, -fz<=-lo."""
    a, bq = sr + D[0] + D[1], sr + D[2] + D[3]
    e, f = -mu * (D[0] - D[1]), -mu * (D[2] - D[3])
    d2 = sr + D[4] + D[5] + mu * mu * ((sr * (D[0] + D[1]) + 4.0 * D[0] * D[1]) / a + (sr * (D[2] + D[3]) + 4.0 * D[2] * D[3]) / bq)
    l20, l21, id2 = e / a, f / bq, 1.0 / d2
    Dinv = np.array([[1.0 / a + l20 * l20 * id2, l20 * l21 * id2, -l20 * id2],
                     [l20 * l21 * id2, 1.0 / bq + l21 * l21 * id2, -l21 * id2],
                     [-l20 * id2, -l21 * id2, id2]])
    ia, ib = 1.0 / a, 1.0 / bq
    iad, ibd = ia * id2, ib * id2
    X = np.zeros((6, 3))
    x2 = -mu * (sr + 2.0 * D[1]) * iad; X[0] = [(1.0 - e * x2) * ia, -f * x2 * ib, x2]
    x2 = -mu * (sr + 2.0 * D[0]) * iad; X[1] = [(-1.0 - e * x2) * ia, -f * x2 * ib, x2]
    x2 = -mu * (sr + 2.0 * D[3]) * ibd; X[2] = [-e * x2 * ia, (1.0 - f * x2) * ib, x2]
    x2 = -mu * (sr + 2.0 * D[2]) * ibd; X[3] = [-e * x2 * ia, (-1.0 - f * x2) * ib, x2]
    X[4] = [-e * id2 * ia, -f * id2 * ib, id2]
    X[5] = -X[4]
    return Dinv, X


def warm_point(prob, u_prev, margin, shift=1):
    """Start point from a previous solution (the kernel's warm start): u_prev shifted by `shift` stages (the last
    stage repeated), pulled `margin` newtons inside every row of the friction pyramid / fz box so that the slacks
    s = b - C u are consistent with u AND at least `margin`; contacts that were in swing in the shifted solution (or
    whose box is narrower than 2 margin) get the cold start value."""
    N, nc, mu = prob.N, prob.nc, prob.mu
    u = np.zeros((N, nc, 3)); s = np.ones((N, nc, 6))
    for k in range(N):
        ns = int(prob.stance[k].sum())
        kp = min(k + shift, N - 1)
        for j in range(nc):
            if not prob.stance[k, j]:
                continue
            lo, hi = prob.lo[k, j], prob.hi[k, j]
            f = np.array(u_prev[kp, j], float)
            m = min(margin, 0.1 * (hi - lo))
            if f[2] <= 0.0:                                   # was a swing contact: cold value
                span = hi - lo
                f = np.array([0.0, 0.0, min(max(prob.weight / ns, lo + 0.1 * span), hi - 0.1 * span)])
            f[2] = min(max(f[2], lo + m), hi - m)
            lim = max(mu * f[2] - m, 0.0)
            f[0] = min(max(f[0], -lim), lim); f[1] = min(max(f[1], -lim), lim)
            u[k, j] = f
            s[k, j] = [mu * f[2] - f[0], mu * f[2] + f[0], mu * f[2] - f[1], mu * f[2] + f[1], hi - f[2], f[2] - lo]
    return u, s


def solve_ipm(prob, gap_tol=1e-10, max_iter=40, linear="riccati", warm=None, trace=None):
    """Returns u[N,nc,3], iters, gap.  `linear`: "stage" (what the kernel runs), "riccati" (textbook stage-wise),
    "dense" (6N x 6N Woodbury matrix) or "stage_s" (the stage form with S_k Gamma formed as a product: what the kernel ran before Theta_k).
    `warm` = dict(u=previous solution [N,nc,3], margin=, mu0=, shift=): start from warm_point() with lambda = mu0 / s."""
    N, nc, mu, sr, bv = prob.N, prob.nc, prob.mu, prob.sr, prob.bv
    M, h = prob.M_h()
    st = prob.stance
    Ct = lambda q: np.array([q[0] - q[1], q[2] - q[3], -mu * (q[0] + q[1] + q[2] + q[3]) + q[4] - q[5]])
    u = np.zeros((N, nc, 3)); s = np.ones((N, nc, 6)); lam = np.zeros((N, nc, 6))
    for k in range(N):
        ns = int(st[k].sum())
        for j in range(nc):
            if st[k, j]:
                span = prob.hi[k, j] - prob.lo[k, j]
                fz = min(max(prob.weight / ns, prob.lo[k, j] + 0.1 * span), prob.hi[k, j] - 0.1 * span)
                u[k, j, 2] = fz
                s[k, j] = [mu * fz] * 4 + [prob.hi[k, j] - fz, fz - prob.lo[k, j]]
    m_ineq = 6.0 * max(int(st.sum()), 1)
    Wk = [prob.W(k) for k in range(N)]

    def grad(u):
        w = np.array([Wk[k] @ u[k].ravel() for k in range(N)])
        v = (M @ w.ravel() + h).reshape(N, 6)
        return np.array([sr * u[k].ravel() + Wk[k].T @ v[k] for k in range(N)]).reshape(N, nc, 3)

    if warm is not None:
        u, s = warm_point(prob, warm["u"], warm["margin"], warm.get("shift", 1))
    gf = grad(u)
    mu0 = max(1e-2, 0.1 * np.abs(gf[st]).max()) if st.any() else 0.0
    if warm is not None:
        mu0 = warm["mu0"]
    lam[st] = mu0 / s[st]
    gap_start = mu0
    it = 0
    exact = True            # the kernel carries grad f along with the step while the gap is large (srbd_solve_warp.cuh)
    while True:
        if exact:
            gf = grad(u)
        gap = float((s[st] * lam[st]).sum() / m_ineq) if st.any() else 0.0
        if trace is not None:
            trace.append((it, gap))
        if warm is not None and it == warm.get("fallback_it", 4) and gap > warm.get("fallback_ratio", 0.5) * gap_start:
            # Safeguard of the warm start (mirrors the kernel): a start point whose multipliers are far off stalls close to
            # the boundary (steps of a few per cent for tens of iterations).  If the gap has not halved after 4 iterations
            # the problem is restarted from the cold point; the iterations spent are counted.
            u2, it2, gap2 = solve_ipm(prob, gap_tol, max_iter - it, linear, None, trace)
            return u2, it + it2, gap2
        if gap <= gap_tol or not st.any() or it >= max_iter:
            if not exact:
                exact = True
                continue
            return u, it, gap
        Dinv = np.zeros((N, nc, 3, 3)); X = np.zeros((N, nc, 6, 3)); S = np.zeros((N, 6, 6))
        for k in range(N):
            for j in range(nc):
                if st[k, j]:
                    Dinv[k, j], X[k, j] = _contact_blocks(lam[k, j] / s[k, j], sr, mu)
                    Wc = Wk[k][:, 3 * j:3 * j + 3]
                    S[k] += Wc @ Dinv[k, j] @ Wc.T
        if linear == "riccati":
            fac = RiccatiFactor(prob, S)
            lin_solve = fac.solve
        elif linear in ("stage", "stage_s"):
            fac = StageFactor(prob, S, theta=linear == "stage")
            lin_solve = fac.solve
        else:
            Sd = np.zeros((6 * N, 6 * N))
            for k in range(N):
                Sd[6 * k:6 * k + 6, 6 * k:6 * k + 6] = S[k]
            Yd = np.linalg.inv(M) + Sd
            lin_solve = lambda t: np.linalg.solve(Yd, t.ravel()).reshape(N, 6)
        ds = np.zeros_like(s); dl = np.zeros_like(s); du = np.zeros_like(u)
        sigma_gap = 0.0
        for pas in range(2):
            q = np.zeros_like(s)
            if pas == 1:
                q[st] = (min(1.0, a_aff / 0.3) * ds[st] * dl[st] - sigma_gap) / s[st]      # (safeguard: see the kernel)
            rhs = np.zeros_like(u); tw = np.zeros((N, 6))
            for k in range(N):
                for j in range(nc):
                    if st[k, j]:
                        rhs[k, j] = -gf[k, j] + Ct(q[k, j])
                        tw[k] += Wk[k][:, 3 * j:3 * j + 3] @ (Dinv[k, j] @ rhs[k, j])
            y = lin_solve(tw)
            amin = 1e300
            wy = np.zeros_like(u)
            for k in range(N):
                for j in range(nc):
                    if st[k, j]:
                        wy[k, j] = Wk[k][:, 3 * j:3 * j + 3].T @ y[k]
                        rt = rhs[k, j] - wy[k, j]
                        du[k, j] = Dinv[k, j] @ rt
                        ds[k, j] = -(X[k, j] @ rt)
                        dl[k, j] = -lam[k, j] - q[k, j] - (lam[k, j] / s[k, j]) * ds[k, j]
                        for i in range(6):
                            if ds[k, j, i] < 0: amin = min(amin, -s[k, j, i] / ds[k, j, i])
                            if dl[k, j, i] < 0: amin = min(amin, -lam[k, j, i] / dl[k, j, i])
            if pas == 0:
                a_aff = min(1.0, amin)
                g_aff = float(((s[st] + a_aff * ds[st]) * (lam[st] + a_aff * dl[st])).sum() / m_ineq)
                sig = (g_aff / gap) ** 3
                sigma_gap = sig * gap
            else:
                alpha = min(1.0, min(0.99999, max(0.995, 1.0 - (1.0 - a_aff) ** 2)) * amin)
        u[st] += alpha * du[st]; lam[st] += alpha * dl[st]; s[st] += alpha * ds[st]
        gf = gf.copy()
        gf[st] += alpha * (sr * du[st] + wy[st])        # y = M W du: the carried gradient
        exact = gap * (1.0 - alpha * (1.0 - sig)) < 1e7 * gap_tol
        it += 1


# ------------------------------------------------------------------------------------------------------
# The form the kernel runs: LDL' based stage factorisation with explicit Gamma_k = (I + P22 S_k)^-1 and
# Psi_k = P12 S_k Gamma_k (closed-loop blocks), so that every sweep step is one small mat-vec.
# NB (measured here, see test_ipm_model_cpu.py): recovering y as M*dw by an adjoint sweep instead of from the
# value function loses ~1e-4 relative on the cpp_node preset (cond(H) 3e8) and stalls the interior point.
# ------------------------------------------------------------------------------------------------------
def _ldl(Amat):
    n = Amat.shape[0]
    A = Amat.copy(); L = np.eye(n); d = np.zeros(n)
    for j in range(n):
        d[j] = A[j, j]
        L[j + 1:, j] = A[j + 1:, j] / d[j]
        A[j + 1:, j + 1:] -= np.outer(L[j + 1:, j], A[j + 1:, j])
    return L, d


class StageFactor:
    """factor_stages of the kernel: Pi = P22^-1, T = Pi + S_k = L D L', Gamma = T^-1 Pi, A21 = Gamma P21, A22 = T^-1;
    S_k times a column is never formed as a product with S_k: S Gamma = Pi (I - Gamma) =: Theta_k (symmetric) and
    S A21 = Pi (P21 - A21).  (S (Gamma x) loses eps |S| |P22| relative: cond(H) 4e11 stalls on it,
    tests/golden/ipm_illconditioned_problems.npz; `theta=False` keeps that form for the test that documents it.)"""

    def __init__(self, prob, S, theta=True):
        N = prob.N
        self.prob, self.S, self.theta = prob, S, theta
        sq = prob.s * prob.q[:12]
        Phi = prob.Phi
        P11, P12, P22 = np.diag(sq[:6]), np.zeros((6, 6)), np.diag(sq[6:])
        self.Gam, self.Psi, self.P12, self.P22, self.Theta = [None] * N, [None] * N, [None] * N, [None] * N, [None] * N
        self.ok = True
        for k in range(N - 1, -1, -1):
            Pi = np.linalg.inv(P22)
            Pi = 0.5 * (Pi + Pi.T)
            L, d = _ldl(Pi + S[k])
            if np.any(np.linalg.eigvalsh(P22) <= 0) or np.any(d <= 0):
                self.ok = False
            tinv = lambda col: np.linalg.solve(L.T, np.linalg.solve(L, col) / d)
            Gam = np.column_stack([tinv(Pi[:, b]) for b in range(6)])
            A21 = np.column_stack([tinv(Pi @ P12.T[:, a]) for a in range(6)])
            A22 = np.column_stack([tinv(np.eye(6)[:, b]) for b in range(6)])
            if theta:
                Theta = np.tril(Pi @ (np.eye(6) - Gam))              # the kernel keeps the lower triangle
                Theta = Theta + np.tril(Theta, -1).T
                Psi = P12 @ (Pi @ (np.eye(6) - Gam))
                A11 = P11 - P12 @ (Pi @ (P12.T - A21))
            else:
                Theta = None
                Psi = P12 @ (S[k] @ Gam)
                A11 = P11 - P12 @ (S[k] @ A21)
            self.Gam[k], self.Psi[k], self.P12[k], self.P22[k], self.Theta[k] = Gam, Psi, P12, P22, Theta
            A12 = A21.T
            M12 = A12 + A11 @ Phi
            P22n = np.diag(sq[6:]) + A22 + Phi.T @ M12 + A21 @ Phi
            P11, P12, P22 = np.diag(sq[:6]) + A11, M12, 0.5 * (P22n + P22n.T)

    def solve(self, t):
        """(M^-1 + S) y = t  ->  y [N,6]."""
        pr = self.prob
        N, Phi = pr.N, pr.Phi
        gt = np.array([self.Gam[k].T @ t[k] for k in range(N)])          # parallel over stages
        c1 = np.array([self.P12[k] @ gt[k] for k in range(N)])
        c2 = np.array([self.P22[k] @ gt[k] for k in range(N)])
        p1 = np.zeros(6); p2 = np.zeros(6)
        p2s = np.zeros((N, 6)); b = np.zeros((N, 6))
        for k in range(N - 1, -1, -1):                   # backward sweep: one 12x6 mat-vec on the chain
            p2s[k] = p2
            g = self.Gam[k] @ p2
            b[k] = gt[k] - (self.Theta[k] @ p2 if self.theta else self.S[k] @ g)        # (off the chain)
            p1 = p1 - self.Psi[k] @ p2 + c1[k]
            p2 = Phi.T @ p1 + g + c2[k]
        y1 = np.zeros(6); y2 = np.zeros(6)
        y = np.zeros((N, 6))
        for k in range(N):                               # forward sweep: one 6x12 mat-vec on the chain
            y1 = y1 + Phi @ y2
            y2 = b[k] + self.Gam[k].T @ y2 - self.Psi[k].T @ y1
            y[k] = self.P12[k].T @ y1 + self.P22[k] @ y2 + p2s[k]       # (off the chain)
        return y
