/*
 * oracle/srbd_mpc_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Plain-C, FP64, single-threaded-per-problem CPU restatement of the reference hot path
 * g1_mpc::MPC::updateMPC  (reference: g1_mpc/src/mpc.cpp:448-473; Python twin
 * g1_mpc/scripts/mpc.py:360-377).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may call into this file.
 *
 * Parity status
 *   build half  (A, B_i, Aqp, Bqp, Ac, l, u, H, g, rollout): PINNED against golden vectors produced
 *               by the reference's own g1_mpc/scripts/mpc.py imported unmodified in the build
 *               container (oracle/gen_golden.py -> tests/golden/, checked by tests/test_oracle_golden.py).
 *   solve half  (OSQP): the arithmetic lives in the third-party OSQP library (C lib through
 *               OsqpEigen at mpc.cpp:380-414, pip `osqp` at mpc.py:341-348).  Its source is NOT under
 *               /root/reference and its version is not pinned by the reference (find_package(osqp),
 *               g1_mpc/CMakeLists.txt:19) -> "parity unpinned".  orc_osqp_solve() restates the
 *               published algorithm (Stellato et al., "OSQP: an operator splitting solver for
 *               quadratic programs", Math. Prog. Comp. 2020; constants as in OSQP 0.6.x) with the
 *               settings the reference call sites leave at default.  orc_qp_exact() returns the
 *               unique minimiser x* of the same QP (active-set solve + KKT certificate), which is
 *               what the "GRFs within 1e-3 N" criterion is checked against.  x* itself is cross-checked
 *               against an independent third-party solver, HiGHS' convex-QP code as bundled with scipy
 *               (tests/highs_qp.py, tests/test_oracle_qp.py): same optimum to 1e-11 relative, same
 *               active rows, x* never the worse point.
 *
 * Storage: every matrix is dense, row-major, double.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define NS 13 /* states: roll,pitch,yaw, px,py,pz, wx,wy,wz, vx,vy,vz, g   (mpc.py:30) */

typedef struct {
    int32_t N;       /* horizon_length_ (mpc.hpp:19,22) */
    int32_t nc;      /* num_contacts_ (mpc.cpp:12)      */
    double mu, fz_min, fz_max, dt; /* ctor arguments (mpc.hpp:22 / mpc.py:12) */
    double mass;     /* robot_mass_ (mpc.hpp:136)       */
    double Ib[9];    /* inertia_body_ row-major (mpc.cpp:25-27 / mpc.py:24-28) */
    double q[NS];    /* diag(Q_weights_) (mpc.cpp:90-96 / mpc.py:47-50)       */
    double r;        /* diag(R_weights_) entry (mpc.cpp:98 / mpc.py:51)       */
    double scale;    /* 2 for the C++ variant (mpc.cpp:361,375), 1 for Python (mpc.py:301,317) */
} orc_params;

typedef struct {
    double rho, sigma, alpha, eps_abs, eps_rel;
    int32_t max_iter, check_termination, scaling, adaptive_rho, adaptive_rho_interval;
    double adaptive_rho_tolerance;
} orc_osqp_settings;

typedef struct {
    int32_t iters, status /* 1 solved, 2 max_iter */, rho_updates;
    double prim_res, dual_res, rho_final;
} orc_osqp_info;

/* ------------------------------------------------------------------------------------------ */
/* small dense helpers                                                                        */
/* ------------------------------------------------------------------------------------------ */
static void matmul(const double *A, const double *B, double *C, int m, int k, int n)
{ /* C[m,n] = A[m,k] B[k,n] */
    for (int i = 0; i < m; i++) {
        double *c = C + (size_t)i * n;
        for (int j = 0; j < n; j++) c[j] = 0.0;
        for (int p = 0; p < k; p++) {
            const double a = A[(size_t)i * k + p];
            if (a == 0.0) continue;
            const double *b = B + (size_t)p * n;
            for (int j = 0; j < n; j++) c[j] += a * b[j];
        }
    }
}

static void inv3(const double *M, double *R)
{
    const double a = M[0], b = M[1], c = M[2], d = M[3], e = M[4], f = M[5], g = M[6], h = M[7], i = M[8];
    const double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
    const double id = 1.0 / det;
    R[0] = (e * i - f * h) * id; R[1] = (c * h - b * i) * id; R[2] = (b * f - c * e) * id;
    R[3] = (f * g - d * i) * id; R[4] = (a * i - c * g) * id; R[5] = (c * d - a * f) * id;
    R[6] = (d * h - e * g) * id; R[7] = (b * g - a * h) * id; R[8] = (a * e - b * d) * id;
}

/* dot product with four independent partial sums (fixed order: reproducible, and the compiler can keep four
 * FMA chains in flight / use one AVX2 register -- a single running sum is latency bound at 4 cycles per term) */
static inline double dot4(const double *a, const double *b, int n)
{
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int k = 0;
    for (; k + 4 <= n; k += 4) { s0 += a[k] * b[k]; s1 += a[k + 1] * b[k + 1]; s2 += a[k + 2] * b[k + 2]; s3 += a[k + 3] * b[k + 3]; }
    for (; k < n; k++) s0 += a[k] * b[k];
    return (s0 + s1) + (s2 + s3);
}

/* in-place lower Cholesky of SPD n x n (row-major, full storage; the strict upper triangle receives L^T so that both
 * triangular solves below stream contiguous rows); returns 0 on success */
static int chol(double *A, int n)
{
    for (int j = 0; j < n; j++) {
        double d = A[(size_t)j * n + j] - dot4(A + (size_t)j * n, A + (size_t)j * n, j);
        if (!(d > 0.0)) return -1;
        d = sqrt(d);
        A[(size_t)j * n + j] = d;
        const double id = 1.0 / d;
        for (int i = j + 1; i < n; i++) {
            const double v = (A[(size_t)i * n + j] - dot4(A + (size_t)i * n, A + (size_t)j * n, j)) * id;
            A[(size_t)i * n + j] = v;
        }
    }
    for (int j = 0; j < n; j++) for (int i = j + 1; i < n; i++) A[(size_t)j * n + i] = A[(size_t)i * n + j];
    return 0;
}
static void chol_solve(const double *L, int n, double *b)
{
    /* forward: L y = b, column sweep over the rows of L^T (stored in the upper triangle) */
    for (int k = 0; k < n; k++) {
        const double yk = b[k] / L[(size_t)k * n + k];
        b[k] = yk;
        const double *lt = L + (size_t)k * n;
        for (int i = k + 1; i < n; i++) b[i] -= lt[i] * yk;
    }
    /* backward: L^T x = y, column sweep over the rows of L */
    for (int i = n - 1; i >= 0; i--) {
        const double xi = b[i] / L[(size_t)i * n + i];
        b[i] = xi;
        const double *lr = L + (size_t)i * n;
        for (int k = 0; k < i; k++) b[k] -= lr[k] * xi;
    }
}

/* ------------------------------------------------------------------------------------------ */
/* build half: follows the reference step by step, as coded                                   */
/* ------------------------------------------------------------------------------------------ */

/* Sizes: A[13*13], Bd[13N*nu], Aqp[13N*13], Bqp[13N*nuN], H[nuN*nuN], g[nuN], Acon[5ncN*nuN],
 * l,u[5ncN].  Any output pointer may be NULL except those needed downstream are computed in
 * scratch.  Returns 0. */
int orc_build(const orc_params *p, const double *x0, const double *xref, const double *contact,
              const double *chor, const double *pcom, double *A_out, double *Bd_out, double *Aqp_out,
              double *Bqp_out, double *H_out, double *g_out, double *Acon_out, double *l_out,
              double *u_out)
{
    const int N = p->N, nc = p->nc, nu = 3 * nc, n = nu * N, nx = NS * N, m = 5 * nc * N;
    double *Bd = (double *)calloc((size_t)nx * nu, sizeof(double));
    double *Aqp = (double *)calloc((size_t)nx * NS, sizeof(double));
    double *Bqp = (double *)calloc((size_t)nx * n, sizeof(double));

    /* extractPsi  mpc.cpp:153-162 */
    double yaw_sum = 0.0;
    for (int i = 0; i < N; i++) yaw_sum += xref[i * NS + 2];
    const double psi = yaw_sum / N;
    /* calculateRotationMatrixT  mpc.cpp:173-178 */
    const double T[9] = {cos(psi), sin(psi), 0, -sin(psi), cos(psi), 0, 0, 0, 1};
    /* calculateAContinuous / calculateADiscrete  mpc.cpp:196-211 */
    double Ac[NS * NS], A[NS * NS];
    memset(Ac, 0, sizeof Ac);
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) Ac[r * NS + 6 + c] = T[c * 3 + r]; /* rotation_z_T_.transpose() */
    for (int r = 0; r < 3; r++) Ac[(3 + r) * NS + 9 + r] = 1.0;
    Ac[11 * NS + 12] = 1.0;
    for (int i = 0; i < NS * NS; i++) A[i] = Ac[i] * p->dt;
    for (int i = 0; i < NS; i++) A[i * NS + i] += 1.0;
    /* calculateBContinuous / calculateBDiscrete  mpc.cpp:213-254 */
    double TI[9], Iw[9], Iwi[9], Tt[9];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) Tt[r * 3 + c] = T[c * 3 + r];
    matmul(T, p->Ib, TI, 3, 3, 3);
    matmul(TI, Tt, Iw, 3, 3, 3);
    inv3(Iw, Iwi);
    for (int i = 0; i < N; i++) {
        for (int j = 0; j < nc; j++) {
            double r[3];
            for (int a = 0; a < 3; a++) r[a] = chor[i * nu + 3 * j + a] - pcom[i * 3 + a];
            const double S[9] = {0, -r[2], r[1], r[2], 0, -r[0], -r[1], r[0], 0}; /* mpc.cpp:164-171 */
            double IS[9];
            matmul(Iwi, S, IS, 3, 3, 3);
            for (int a = 0; a < 3; a++)
                for (int b = 0; b < 3; b++) {
                    Bd[(size_t)(i * NS + 6 + a) * nu + 3 * j + b] = IS[a * 3 + b] * p->dt;
                    Bd[(size_t)(i * NS + 9 + a) * nu + 3 * j + b] = ((a == b) ? 1.0 / p->mass : 0.0) * p->dt;
                }
        }
    }
    /* calculateAqp  mpc.cpp:256-270: block i = block(i-1) * A */
    memcpy(Aqp, A, sizeof A);
    for (int i = 1; i < N; i++) matmul(Aqp + (size_t)(i - 1) * NS * NS, A, Aqp + (size_t)i * NS * NS, NS, NS, NS);
    /* calculateBqp  mpc.cpp:272-291 */
    double *blk = (double *)malloc(sizeof(double) * NS * nu);
    for (int i = 0; i < N; i++)
        for (int j = 0; j <= i; j++) {
            const double *Bj = Bd + (size_t)j * NS * nu;
            const double *src = Bj;
            if (i != j) {
                matmul(Aqp + (size_t)(i - j - 1) * NS * NS, Bj, blk, NS, NS, nu);
                src = blk;
            }
            for (int r = 0; r < NS; r++)
                memcpy(Bqp + (size_t)(i * NS + r) * n + (size_t)j * nu, src + (size_t)r * nu, sizeof(double) * nu);
        }
    free(blk);
    /* calculateAc  mpc.cpp:293-326 (dense view) */
    if (Acon_out) {
        memset(Acon_out, 0, sizeof(double) * (size_t)m * n);
        const double G[15] = {1, 0, p->mu, 1, 0, -p->mu, 0, 1, p->mu, 0, 1, -p->mu, 0, 0, 1};
        for (int b = 0; b < nc * N; b++)
            for (int r = 0; r < 5; r++)
                for (int c = 0; c < 3; c++) Acon_out[(size_t)(5 * b + r) * n + 3 * b + c] = G[r * 3 + c];
    }
    /* calculateBounds  mpc.cpp:328-357 */
    if (l_out && u_out)
        for (int i = 0; i < N; i++)
            for (int j = 0; j < nc; j++) {
                const double c = contact[i * nc + j];
                double *l = l_out + (size_t)(i * nc + j) * 5, *u = u_out + (size_t)(i * nc + j) * 5;
                l[0] = 0; l[1] = -INFINITY; l[2] = 0; l[3] = -INFINITY; l[4] = p->fz_min * c;
                u[0] = INFINITY; u[1] = 0; u[2] = INFINITY; u[3] = 0; u[4] = p->fz_max * c;
            }
    /* calculateHessian  mpc.cpp:359-364: H = s (Bqp^T Q Bqp + R), with Q the dense block-diagonal
     * matrix of setQ (mpc.cpp:180-186).  The product is evaluated as (Bqp^T Q) Bqp like Eigen does;
     * Q is diagonal so (Bqp^T Q) is a column scaling -- the zero off-diagonal terms the reference
     * multiplies through contribute exact zeros. */
    if (H_out || g_out) {
        double *BtQ = (double *)malloc(sizeof(double) * (size_t)n * nx); /* n x nx */
        for (int c = 0; c < n; c++)
            for (int r = 0; r < nx; r++) BtQ[(size_t)c * nx + r] = Bqp[(size_t)r * n + c] * p->q[r % NS];
        if (H_out) {
            matmul(BtQ, Bqp, H_out, n, nx, n);
            for (size_t i = 0; i < (size_t)n * n; i++) H_out[i] *= p->scale;
            for (int i = 0; i < n; i++) H_out[(size_t)i * n + i] += p->scale * p->r;
        }
        /* calculateGradient  mpc.cpp:366-376 */
        if (g_out) {
            double *e = (double *)malloc(sizeof(double) * nx);
            for (int r = 0; r < nx; r++) {
                double s = 0.0;
                for (int c = 0; c < NS; c++) s += Aqp[(size_t)r * NS + c] * x0[c];
                e[r] = s - xref[r]; /* row-major flatten of x_ref_hor */
            }
            for (int c = 0; c < n; c++) {
                double s = 0.0;
                for (int r = 0; r < nx; r++) s += BtQ[(size_t)c * nx + r] * e[r];
                g_out[c] = p->scale * s;
            }
            free(e);
        }
        free(BtQ);
    }
    if (A_out) memcpy(A_out, A, sizeof A);
    if (Bd_out) memcpy(Bd_out, Bd, sizeof(double) * (size_t)nx * nu);
    if (Aqp_out) memcpy(Aqp_out, Aqp, sizeof(double) * (size_t)nx * NS);
    if (Bqp_out) memcpy(Bqp_out, Bqp, sizeof(double) * (size_t)nx * n);
    free(Bd); free(Aqp); free(Bqp);
    return 0;
}

/* computeRollout  mpc.cpp:427-446.  x_opt is in/out (rows that are not written keep their value,
 * as the member x_opt_ does in the reference). */
void orc_rollout(const orc_params *p, const double *A, const double *Bd, const double *u_opt,
                 const double *x_current, int only_first_step, double *x_opt)
{
    const int N = p->N, nu = 3 * p->nc;
    if (x_current) memcpy(x_opt, x_current, sizeof(double) * NS);
    for (int i = 1; i < N; i++) {
        for (int r = 0; r < NS; r++) {
            double s = 0.0;
            for (int c = 0; c < NS; c++) s += A[r * NS + c] * x_opt[(i - 1) * NS + c];
            for (int c = 0; c < nu; c++) s += Bd[(size_t)((i - 1) * NS + r) * nu + c] * u_opt[(i - 1) * nu + c];
            x_opt[i * NS + r] = s;
        }
        if (only_first_step) return;
    }
}

/* ------------------------------------------------------------------------------------------ */
/* OSQP restatement (see header: third-party, unpinned)                                        */
/* ------------------------------------------------------------------------------------------ */
#define OSQP_INFTY 1e30
#define MIN_SCALING 1e-4
#define MAX_SCALING 1e4
#define RHO_MIN 1e-6
#define RHO_MAX 1e6
#define RHO_EQ_OVER_RHO_INEQ 1e3
#define RHO_TOL 1e-4

void orc_osqp_default_settings(orc_osqp_settings *s)
{
    s->rho = 0.1; s->sigma = 1e-6; s->alpha = 1.6; s->eps_abs = 1e-3; s->eps_rel = 1e-3;
    s->max_iter = 4000; s->check_termination = 25; s->scaling = 10; s->adaptive_rho = 1;
    s->adaptive_rho_interval = 100; /* OSQP 0.6.x picks this from wall-clock timing (non-reproducible);
                                       100 = 4 x check_termination is the library's non-profiling value */
    s->adaptive_rho_tolerance = 5.0;
}

static double limit_scaling(double v) { v = v < MIN_SCALING ? 1.0 : v; return v > MAX_SCALING ? MAX_SCALING : v; }
static double vinf(const double *v, int n) { double m = 0; for (int i = 0; i < n; i++) { double a = fabs(v[i]); if (a > m) m = a; } return m; }

typedef struct {
    int n, m;
    double *P, *q, *l, *u;         /* scaled copies */
    int *Ap, *Aj; double *Ax;      /* A in CSR (row pointers, column indices, scaled values): OSQP keeps A sparse; only the
                                      stored entries are touched, in column order, so sums equal the dense ones */
    double *D, *E, c;              /* scaling */
    double *rho_vec, *K;           /* K = chol(P + sigma I + A^T rho A) */
} osqp_work;

static int osqp_factor(osqp_work *w, double rho, double sigma)
{
    const int n = w->n, m = w->m;
    for (int i = 0; i < m; i++) {
        if (w->l[i] < -OSQP_INFTY * MIN_SCALING && w->u[i] > OSQP_INFTY * MIN_SCALING) w->rho_vec[i] = RHO_MIN;
        else if (w->u[i] - w->l[i] < RHO_TOL) w->rho_vec[i] = RHO_EQ_OVER_RHO_INEQ * rho;
        else w->rho_vec[i] = rho;
    }
    memcpy(w->K, w->P, sizeof(double) * (size_t)n * n);
    for (int i = 0; i < n; i++) w->K[(size_t)i * n + i] += sigma;
    for (int r = 0; r < m; r++) {
        const double rr = w->rho_vec[r];
        for (int ii = w->Ap[r]; ii < w->Ap[r + 1]; ii++) {
            const double ai = rr * w->Ax[ii];
            for (int jj = w->Ap[r]; jj < w->Ap[r + 1]; jj++) w->K[(size_t)w->Aj[ii] * n + w->Aj[jj]] += ai * w->Ax[jj];
        }
    }
    return chol(w->K, n);
}

/* min 1/2 x'Px + q'x  s.t. l <= Ax <= u ; P dense symmetric n x n, A dense m x n.  Cold start. */
int orc_osqp_solve(int n, int m, const double *P_in, const double *q_in, const double *A_in,
                   const double *l_in, const double *u_in, const orc_osqp_settings *st, double *x_out,
                   double *y_out, orc_osqp_info *info)
{
    osqp_work w; w.n = n; w.m = m;
    w.P = (double *)malloc(sizeof(double) * (size_t)n * n); memcpy(w.P, P_in, sizeof(double) * (size_t)n * n);
    {   /* CSR of the non-zero entries of A (Ac is block diagonal with 5x3 blocks, mpc.cpp:296-322: 9 non-zeros per block) */
        int nnz = 0;
        for (size_t i = 0; i < (size_t)m * n; i++) nnz += A_in[i] != 0.0;
        w.Ap = (int *)malloc(sizeof(int) * (m + 1)); w.Aj = (int *)malloc(sizeof(int) * (nnz + 1)); w.Ax = (double *)malloc(sizeof(double) * (nnz + 1));
        int k = 0;
        for (int i = 0; i < m; i++) {
            w.Ap[i] = k;
            for (int j = 0; j < n; j++) if (A_in[(size_t)i * n + j] != 0.0) { w.Aj[k] = j; w.Ax[k] = A_in[(size_t)i * n + j]; k++; }
        }
        w.Ap[m] = k;
    }
    w.q = (double *)malloc(sizeof(double) * n); memcpy(w.q, q_in, sizeof(double) * n);
    w.l = (double *)malloc(sizeof(double) * m); w.u = (double *)malloc(sizeof(double) * m);
    for (int i = 0; i < m; i++) { /* bounds are clipped to +-OSQP_INFTY at setup */
        w.l[i] = l_in[i] < -OSQP_INFTY ? -OSQP_INFTY : l_in[i];
        w.u[i] = u_in[i] > OSQP_INFTY ? OSQP_INFTY : u_in[i];
    }
    w.D = (double *)malloc(sizeof(double) * n); w.E = (double *)malloc(sizeof(double) * m); w.c = 1.0;
    for (int i = 0; i < n; i++) w.D[i] = 1.0;
    for (int i = 0; i < m; i++) w.E[i] = 1.0;
    w.rho_vec = (double *)malloc(sizeof(double) * m);
    w.K = (double *)malloc(sizeof(double) * (size_t)n * n);
    double *Dt = (double *)malloc(sizeof(double) * n), *Et = (double *)malloc(sizeof(double) * m);

    /* Ruiz equilibration with cost scaling */
    for (int it = 0; it < st->scaling; it++) {
        for (int j = 0; j < n; j++) Dt[j] = 0.0;
        for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) { double a = fabs(w.P[(size_t)i * n + j]); if (a > Dt[j]) Dt[j] = a; }
        for (int i = 0; i < m; i++) {
            double e = 0.0;
            for (int k = w.Ap[i]; k < w.Ap[i + 1]; k++) { double a = fabs(w.Ax[k]); if (a > Dt[w.Aj[k]]) Dt[w.Aj[k]] = a; if (a > e) e = a; }
            Et[i] = e;
        }
        for (int j = 0; j < n; j++) Dt[j] = 1.0 / sqrt(limit_scaling(Dt[j]));
        for (int i = 0; i < m; i++) Et[i] = 1.0 / sqrt(limit_scaling(Et[i]));
        for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) w.P[(size_t)i * n + j] *= Dt[i] * Dt[j];
        for (int i = 0; i < m; i++) for (int k = w.Ap[i]; k < w.Ap[i + 1]; k++) w.Ax[k] *= Et[i] * Dt[w.Aj[k]];
        for (int j = 0; j < n; j++) { w.q[j] *= Dt[j]; w.D[j] *= Dt[j]; }
        for (int i = 0; i < m; i++) w.E[i] *= Et[i];
        /* cost normalisation */
        double mean = 0.0;
        for (int j = 0; j < n; j++) { double cmax = 0; for (int i = 0; i < n; i++) { double a = fabs(w.P[(size_t)i * n + j]); if (a > cmax) cmax = a; } mean += cmax; }
        mean /= n;
        mean = limit_scaling(mean);
        double qn = limit_scaling(vinf(w.q, n));
        double ct = mean > qn ? mean : qn;
        ct = 1.0 / limit_scaling(ct);
        for (size_t i = 0; i < (size_t)n * n; i++) w.P[i] *= ct;
        for (int j = 0; j < n; j++) w.q[j] *= ct;
        w.c *= ct;
    }
    for (int i = 0; i < m; i++) { w.l[i] *= w.E[i]; w.u[i] *= w.E[i]; }
    free(Dt); free(Et);

    double rho = st->rho;
    int rho_updates = 0;
    int rc = osqp_factor(&w, rho, st->sigma);
    double *x = (double *)calloc(n, sizeof(double)), *z = (double *)calloc(m, sizeof(double)), *y = (double *)calloc(m, sizeof(double));
    double *xt = (double *)malloc(sizeof(double) * n), *zt = (double *)malloc(sizeof(double) * m);
    double *Ax = (double *)malloc(sizeof(double) * m), *Px = (double *)malloc(sizeof(double) * n), *Aty = (double *)malloc(sizeof(double) * n);
    int iter = 0, status = 2;
    double prim_res = 0, dual_res = 0;
    for (iter = 1; rc == 0 && iter <= st->max_iter; iter++) {
        /* x_tilde: (P + sigma I + A' rho A) xt = sigma x - q + A'(rho z - y) */
        for (int j = 0; j < n; j++) xt[j] = st->sigma * x[j] - w.q[j];
        for (int i = 0; i < m; i++) {
            const double t = w.rho_vec[i] * z[i] - y[i];
            if (t == 0.0) continue;
            for (int k = w.Ap[i]; k < w.Ap[i + 1]; k++) xt[w.Aj[k]] += w.Ax[k] * t;
        }
        chol_solve(w.K, n, xt);
        for (int i = 0; i < m; i++) { double s = 0; for (int k = w.Ap[i]; k < w.Ap[i + 1]; k++) s += w.Ax[k] * xt[w.Aj[k]]; zt[i] = s; }
        for (int j = 0; j < n; j++) x[j] = st->alpha * xt[j] + (1.0 - st->alpha) * x[j];
        for (int i = 0; i < m; i++) {
            const double zh = st->alpha * zt[i] + (1.0 - st->alpha) * z[i];
            double zn = zh + y[i] / w.rho_vec[i];
            zn = zn < w.l[i] ? w.l[i] : (zn > w.u[i] ? w.u[i] : zn);
            y[i] += w.rho_vec[i] * (zh - zn);
            z[i] = zn;
        }
        const int check = (st->check_termination && iter % st->check_termination == 0);
        const int adapt = (st->adaptive_rho && st->adaptive_rho_interval && iter % st->adaptive_rho_interval == 0);
        if (!check && !adapt && iter != st->max_iter) continue;
        /* residuals */
        for (int i = 0; i < m; i++) { double s = 0; for (int k = w.Ap[i]; k < w.Ap[i + 1]; k++) s += w.Ax[k] * x[w.Aj[k]]; Ax[i] = s; }
        for (int i = 0; i < n; i++) { double s = 0; const double *pr = w.P + (size_t)i * n; for (int j = 0; j < n; j++) s += pr[j] * x[j]; Px[i] = s; }
        for (int j = 0; j < n; j++) Aty[j] = 0.0;
        for (int i = 0; i < m; i++) { if (y[i] == 0.0) continue; for (int k = w.Ap[i]; k < w.Ap[i + 1]; k++) Aty[w.Aj[k]] += w.Ax[k] * y[i]; }
        if (check || iter == st->max_iter) { /* unscaled termination (scaled_termination = 0) */
            double pr = 0, nAx = 0, nz = 0, dr = 0, nq = 0, nAty = 0, nPx = 0;
            for (int i = 0; i < m; i++) {
                const double ei = 1.0 / w.E[i];
                double a = fabs(ei * (Ax[i] - z[i])); if (a > pr) pr = a;
                a = fabs(ei * Ax[i]); if (a > nAx) nAx = a;
                a = fabs(ei * z[i]); if (a > nz) nz = a;
            }
            for (int j = 0; j < n; j++) {
                const double di = 1.0 / w.D[j];
                double a = fabs(di * (Px[j] + w.q[j] + Aty[j])); if (a > dr) dr = a;
                a = fabs(di * w.q[j]); if (a > nq) nq = a;
                a = fabs(di * Aty[j]); if (a > nAty) nAty = a;
                a = fabs(di * Px[j]); if (a > nPx) nPx = a;
            }
            const double ci = 1.0 / w.c;
            prim_res = pr; dual_res = ci * dr;
            const double eps_p = st->eps_abs + st->eps_rel * (nAx > nz ? nAx : nz);
            double md = nq > nAty ? nq : nAty; md = md > nPx ? md : nPx;
            const double eps_d = st->eps_abs + st->eps_rel * ci * md;
            if (prim_res <= eps_p && dual_res <= eps_d) { status = 1; break; }
        }
        if (adapt) { /* compute_rho_estimate on scaled quantities */
            double pr = 0, dr = 0, nAx = vinf(Ax, m), nz = vinf(z, m);
            for (int i = 0; i < m; i++) { double a = fabs(Ax[i] - z[i]); if (a > pr) pr = a; }
            for (int j = 0; j < n; j++) { double a = fabs(Px[j] + w.q[j] + Aty[j]); if (a > dr) dr = a; }
            double md = vinf(w.q, n); double t = vinf(Aty, n); md = md > t ? md : t; t = vinf(Px, n); md = md > t ? md : t;
            pr /= ((nAx > nz ? nAx : nz) + 1e-10);
            dr /= (md + 1e-10);
            double rho_new = rho * sqrt(pr / (dr + 1e-10));
            rho_new = rho_new < RHO_MIN ? RHO_MIN : (rho_new > RHO_MAX ? RHO_MAX : rho_new);
            if (rho_new > rho * st->adaptive_rho_tolerance || rho_new < rho / st->adaptive_rho_tolerance) {
                rho = rho_new; rho_updates++;
                rc = osqp_factor(&w, rho, st->sigma);
            }
        }
    }
    if (iter > st->max_iter) iter = st->max_iter;
    for (int j = 0; j < n; j++) x_out[j] = w.D[j] * x[j];
    if (y_out) for (int i = 0; i < m; i++) y_out[i] = w.E[i] * y[i] / w.c;
    if (info) { info->iters = iter; info->status = rc ? -1 : status; info->rho_updates = rho_updates; info->prim_res = prim_res; info->dual_res = dual_res; info->rho_final = rho; }
    free(w.P); free(w.Ap); free(w.Aj); free(w.Ax); free(w.q); free(w.l); free(w.u); free(w.D); free(w.E); free(w.rho_vec); free(w.K);
    free(x); free(z); free(y); free(xt); free(zt); free(Ax); free(Px); free(Aty);
    return rc;
}

/* ------------------------------------------------------------------------------------------ */
/* exact minimiser x* of the reference QP (ground truth for the 1e-3 N criterion)              */
/* ------------------------------------------------------------------------------------------ */

/* Gaussian elimination with partial pivoting, in place; returns 0 on success */
static int gesv(double *M, double *b, int n)
{
    for (int k = 0; k < n; k++) {
        int piv = k; double best = fabs(M[(size_t)k * n + k]);
        for (int i = k + 1; i < n; i++) { double a = fabs(M[(size_t)i * n + k]); if (a > best) { best = a; piv = i; } }
        if (best == 0.0) return -1;
        if (piv != k) {
            for (int j = 0; j < n; j++) { double t = M[(size_t)k * n + j]; M[(size_t)k * n + j] = M[(size_t)piv * n + j]; M[(size_t)piv * n + j] = t; }
            double t = b[k]; b[k] = b[piv]; b[piv] = t;
        }
        const double ip = 1.0 / M[(size_t)k * n + k];
        for (int i = k + 1; i < n; i++) {
            const double f = M[(size_t)i * n + k] * ip;
            if (f == 0.0) continue;
            M[(size_t)i * n + k] = f; /* keep multiplier (unused later, rhs handled here) */
            for (int j = k + 1; j < n; j++) M[(size_t)i * n + j] -= f * M[(size_t)k * n + j];
            b[i] -= f * b[k];
        }
    }
    for (int i = n - 1; i >= 0; i--) {
        double s = b[i];
        for (int j = i + 1; j < n; j++) s -= M[(size_t)i * n + j] * b[j];
        b[i] = s / M[(size_t)i * n + i];
    }
    return 0;
}

/*
 * Specialised to the reference's constraint structure (mpc.cpp:293-357): 5 rows per contact,
 * [fx+mu fz >= 0, fx-mu fz <= 0, fy+mu fz >= 0, fy-mu fz <= 0, lo <= fz <= hi].
 * Contacts with lo == hi == 0 (swing, contact flag 0) force f = 0 and are eliminated first.
 * Primal active-set iteration started from a tight-tolerance ADMM solution; each step solves the
 * equality-constrained KKT system with long-double iterative refinement.
 * cert[0] = ||Hx+g+A'y||inf, cert[1] = max primal violation, cert[2] = max wrong-sign multiplier,
 * cert[3] = number of active-set changes.  y uses OSQP's sign convention (y<0 at lower bound).
 * Returns 0 if the KKT certificate holds to (1e-7 * (1+||g||inf), 1e-9, 1e-9).
 */
int orc_qp_exact(const orc_params *p, const double *H, const double *g, const double *contact,
                 double *x_out, double *y_out, double *cert)
{
    const int N = p->N, nc = p->nc, ncon = N * nc, n = 3 * ncon, m = 5 * ncon;
    const double mu = p->mu;
    /* start: ADMM at tight tolerance on the full problem */
    double *Acon = (double *)calloc((size_t)m * n, sizeof(double));
    double *l = (double *)malloc(sizeof(double) * m), *u = (double *)malloc(sizeof(double) * m);
    const double G[15] = {1, 0, mu, 1, 0, -mu, 0, 1, mu, 0, 1, -mu, 0, 0, 1};
    for (int b = 0; b < ncon; b++) {
        for (int r = 0; r < 5; r++) for (int c = 0; c < 3; c++) Acon[(size_t)(5 * b + r) * n + 3 * b + c] = G[r * 3 + c];
        l[5 * b + 0] = 0; l[5 * b + 1] = -INFINITY; l[5 * b + 2] = 0; l[5 * b + 3] = -INFINITY; l[5 * b + 4] = p->fz_min * contact[b];
        u[5 * b + 0] = INFINITY; u[5 * b + 1] = 0; u[5 * b + 2] = INFINITY; u[5 * b + 3] = 0; u[5 * b + 4] = p->fz_max * contact[b];
    }
    orc_osqp_settings st; orc_osqp_default_settings(&st);
    st.eps_abs = 1e-7; st.eps_rel = 1e-7; st.max_iter = 20000; st.adaptive_rho_interval = 50;
    double *x = (double *)malloc(sizeof(double) * n), *y = (double *)malloc(sizeof(double) * m);
    orc_osqp_info info;
    orc_osqp_solve(n, m, H, g, Acon, l, u, &st, x, y, &info);

    /* active set: state[row] = -1 lower active, +1 upper active, 0 inactive ; fixed[b]=1 swing */
    int *state = (int *)calloc(m, sizeof(int));
    int *fixed = (int *)calloc(ncon, sizeof(int));
    for (int b = 0; b < ncon; b++) {
        if (u[5 * b + 4] - l[5 * b + 4] <= 0.0 && l[5 * b + 4] == 0.0) { fixed[b] = 1; continue; }
        for (int r = 0; r < 5; r++) {
            const int i = 5 * b + r;
            double ax = 0; for (int c = 0; c < 3; c++) ax += G[r * 3 + c] * x[3 * b + c];
            if (y[i] < -1e-9 && fabs(ax - l[i]) < 1e-3) state[i] = -1;
            else if (y[i] > 1e-9 && fabs(ax - u[i]) < 1e-3) state[i] = +1;
        }
    }
    int *vidx = (int *)malloc(sizeof(int) * n); int nf = 0;
    for (int b = 0; b < ncon; b++) if (!fixed[b]) for (int c = 0; c < 3; c++) vidx[nf++] = 3 * b + c;
    int changes = 0, ok = 0;
    double *K = NULL, *rhs = NULL, *sol = NULL; int *arow = (int *)malloc(sizeof(int) * m);
    for (int outer = 0; outer < 400; outer++) {
        int na = 0;
        for (int i = 0; i < m; i++) if (!fixed[i / 5] && state[i]) arow[na++] = i;
        const int dim = nf + na;
        K = (double *)realloc(K, sizeof(double) * (size_t)dim * dim);
        rhs = (double *)realloc(rhs, sizeof(double) * dim);
        sol = (double *)realloc(sol, sizeof(double) * dim);
        long double *acc = (long double *)calloc(dim, sizeof(long double));
        double *Kc = (double *)malloc(sizeof(double) * (size_t)dim * dim);
        memset(Kc, 0, sizeof(double) * (size_t)dim * dim);
        for (int a = 0; a < nf; a++) for (int b2 = 0; b2 < nf; b2++) Kc[(size_t)a * dim + b2] = H[(size_t)vidx[a] * n + vidx[b2]];
        for (int k = 0; k < na; k++) {
            const int i = arow[k];
            for (int a = 0; a < nf; a++) { const double v = Acon[(size_t)i * n + vidx[a]]; Kc[(size_t)(nf + k) * dim + a] = v; Kc[(size_t)a * dim + nf + k] = v; }
        }
        double *b0 = (double *)malloc(sizeof(double) * dim);
        for (int a = 0; a < nf; a++) b0[a] = -g[vidx[a]];
        for (int k = 0; k < na; k++) b0[nf + k] = state[arow[k]] < 0 ? l[arow[k]] : u[arow[k]];
        memset(sol, 0, sizeof(double) * dim);
        for (int a = 0; a < dim; a++) acc[a] = 0.0L;
        int fail = 0;
        for (int ref = 0; ref < 4 && !fail; ref++) { /* iterative refinement, residual in long double */
            for (int a = 0; a < dim; a++) {
                long double s = b0[a];
                for (int b2 = 0; b2 < dim; b2++) s -= (long double)Kc[(size_t)a * dim + b2] * acc[b2];
                rhs[a] = (double)s;
            }
            memcpy(K, Kc, sizeof(double) * (size_t)dim * dim);
            if (gesv(K, rhs, dim)) { fail = 1; break; }
            for (int a = 0; a < dim; a++) acc[a] += rhs[a];
        }
        for (int a = 0; a < dim; a++) sol[a] = (double)acc[a];
        free(acc); free(Kc); free(b0);
        if (fail) break;
        memset(x, 0, sizeof(double) * n); memset(y, 0, sizeof(double) * m);
        for (int a = 0; a < nf; a++) x[vidx[a]] = sol[a];
        for (int k = 0; k < na; k++) y[arow[k]] = sol[nf + k];
        /* most violated inactive row */
        int worst = -1, wside = 0; double wv = 1e-9;
        for (int i = 0; i < m; i++) {
            if (fixed[i / 5] || state[i]) continue;
            const int b = i / 5, r = i % 5;
            double ax = 0; for (int c = 0; c < 3; c++) ax += G[r * 3 + c] * x[3 * b + c];
            if (l[i] - ax > wv) { wv = l[i] - ax; worst = i; wside = -1; }
            if (ax - u[i] > wv) { wv = ax - u[i]; worst = i; wside = +1; }
        }
        if (worst >= 0) { state[worst] = wside; changes++; continue; }
        /* most wrong-signed multiplier */
        double wm = 1e-9; worst = -1;
        for (int k = 0; k < na; k++) {
            const int i = arow[k];
            const double bad = state[i] < 0 ? y[i] : -y[i]; /* lower-active needs y<=0, upper-active y>=0 */
            if (bad > wm) { wm = bad; worst = i; }
        }
        if (worst >= 0) { state[worst] = 0; changes++; continue; }
        ok = 1; break;
    }
    /* certificate on the full problem */
    double rd = 0, pv = 0, ws = 0;
    for (int j = 0; j < n; j++) {
        double s = g[j];
        for (int k = 0; k < n; k++) s += H[(size_t)j * n + k] * x[k];
        for (int r = 0; r < 5; r++) { const int i = 5 * (j / 3) + r; s += Acon[(size_t)i * n + j] * y[i]; }
        if (!fixed[j / 3] && fabs(s) > rd) rd = fabs(s);
        if (fixed[j / 3]) { /* swing: multipliers absorb the gradient; put it on the rows for reporting */ }
    }
    for (int i = 0; i < m; i++) {
        const int b = i / 5, r = i % 5;
        double ax = 0; for (int c = 0; c < 3; c++) ax += G[r * 3 + c] * x[3 * b + c];
        if (l[i] - ax > pv) pv = l[i] - ax;
        if (ax - u[i] > pv) pv = ax - u[i];
        if (!fixed[b]) {
            const double tol_act = 1e-7;
            if (y[i] < 0 && fabs(ax - l[i]) > tol_act && -y[i] > ws) ws = -y[i];
            if (y[i] > 0 && fabs(ax - u[i]) > tol_act && y[i] > ws) ws = y[i];
        }
    }
    /* swing contacts: a valid multiplier set always exists (the 5 rows positively span R^3 at f=0);
     * fill one so that the full stationarity residual can be checked by the caller:
     *   grad = (Hx+g) restricted to the contact;  need G' y = -grad with y1<=0,y2>=0,y3<=0,y4>=0, y5 free */
    for (int b = 0; b < ncon; b++) {
        if (!fixed[b]) continue;
        double gr[3];
        for (int c = 0; c < 3; c++) { double s = g[3 * b + c]; for (int k = 0; k < n; k++) s += H[(size_t)(3 * b + c) * n + k] * x[k]; gr[c] = s; }
        double *yy = y + 5 * b;
        yy[0] = yy[1] = yy[2] = yy[3] = 0;
        if (-gr[0] <= 0) yy[0] = -gr[0]; else yy[1] = -gr[0];
        if (-gr[1] <= 0) yy[2] = -gr[1]; else yy[3] = -gr[1];
        yy[4] = -gr[2] - mu * (yy[0] - yy[1] + yy[2] - yy[3]);
    }
    memcpy(x_out, x, sizeof(double) * n);
    if (y_out) memcpy(y_out, y, sizeof(double) * m);
    if (cert) { cert[0] = rd; cert[1] = pv; cert[2] = ws; cert[3] = changes; }
    const double gin = vinf(g, n);
    const int good = ok && rd <= 1e-7 * (1.0 + gin) && pv <= 1e-9 && ws <= 1e-9;
    free(Acon); free(l); free(u); free(x); free(y); free(state); free(fixed); free(vidx); free(arow);
    free(K); free(rhs); free(sol);
    return good ? 0 : 1;
}

/* ------------------------------------------------------------------------------------------ */
/* whole path: MPC::updateMPC (mpc.cpp:448-473) over a batch, one problem per OpenMP thread    */
/* ------------------------------------------------------------------------------------------ */
int orc_update_batch(const orc_params *p, int B, const double *x0, const double *xref,
                     const double *contact, const double *chor, const double *pcom,
                     const double *x_current /* may be NULL */, int one_rollout,
                     const orc_osqp_settings *st, double *u_opt, double *x_opt, int32_t *iters,
                     int nthreads)
{
    const int N = p->N, nc = p->nc, nu = 3 * nc, n = nu * N, nx = NS * N, m = 5 * nc * N;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
    int err = 0;
#pragma omp parallel
    {
        double *A = (double *)malloc(sizeof(double) * NS * NS), *Bd = (double *)malloc(sizeof(double) * (size_t)nx * nu);
        double *H = (double *)malloc(sizeof(double) * (size_t)n * n), *g = (double *)malloc(sizeof(double) * n);
        double *Ac = (double *)malloc(sizeof(double) * (size_t)m * n), *l = (double *)malloc(sizeof(double) * m), *u = (double *)malloc(sizeof(double) * m);
#pragma omp for schedule(dynamic, 1)
        for (int b = 0; b < B; b++) {
            orc_build(p, x0 + (size_t)b * NS, xref + (size_t)b * N * NS, contact + (size_t)b * N * nc,
                      chor + (size_t)b * N * nu, pcom + (size_t)b * N * 3, A, Bd, NULL, NULL, H, g, Ac, l, u);
            orc_osqp_info info;
            if (orc_osqp_solve(n, m, H, g, Ac, l, u, st, u_opt + (size_t)b * n, NULL, &info)) {
#pragma omp atomic write
                err = 1;
            }
            if (iters) iters[b] = info.iters;
            orc_rollout(p, A, Bd, u_opt + (size_t)b * n, x_current ? x_current + (size_t)b * NS : NULL,
                        one_rollout, x_opt + (size_t)b * N * NS);
        }
        free(A); free(Bd); free(H); free(g); free(Ac); free(l); free(u);
    }
    return err;
}

int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
