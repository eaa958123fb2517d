// g1mpc_kernels.cu -- B200 (sm_100a) batched SRBD-MPC: linearise -> condensed build -> constraints ->
// QP solve -> rollout, one WARP per problem, everything resident in shared memory, FP64.
//
// Reference path replaced: g1_mpc::MPC::updateMPC (g1_mpc/src/mpc.cpp:448-473) and every step it calls
// (mpc.cpp:153-446); Python twin g1_mpc/scripts/mpc.py:109-391.  This is not a translation:
//
//  * solve kernel (srbd_solve_warp.cuh): the condensed Hessian is never formed.  The reference Hessian
//    (mpc.cpp:361) factors as H = s r I + W' M W, g = W' h with W the per-step force->wrench map of
//    mpc.cpp:213-254 and M the tracking cost seen through the (nilpotent-Ac, mpc.cpp:196-211) dynamics; the
//    interior-point Newton systems are solved stage by stage (Riccati), O(N) work and memory.
//  * emit kernel (below): materialises the reference's dense Aqp/Bqp/H/g/Ac/l/u (mpc.cpp:256-376) in closed
//    form for parity tests and for callers that read those attributes; HBM-write bound.
//  * host side: C-ABI of include/g1mpc_b200.h.
#include "../../include/g1mpc_b200.h"

#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include <string>
#include <vector>

namespace {

constexpr int NS = G1MPC_NUM_STATES;
constexpr int kWarpThreads = 32;   // one warp per problem

thread_local std::string g_last_error;

struct DevParams {
    double mu, fz_min, fz_max, dt, bv, weight, sr, scale, r_weight, gap_tol, dual_tol, warm_margin, warm_mu0;
    double Ib[9];
    double q[13];
    int max_iter;
};

struct BatchPtrs {
    const double *x0, *x_ref, *contact, *c_hor, *p_com, *x_current;
    double *u_opt, *x_opt, *resid, *y_opt;
    int32_t *status, *iters;
    int one_rollout;
    const double *u_warm;       // [B,N,3nc] a previous solution to start from, or null (cold start)
    int warm_shift;             // stages the horizon has moved since that solution (closed loop: 1)
    double *workspace;          // global-workspace variant of the kernel: one slot of WLay::oEnd doubles per resident warp
};

// ------------------------------------------------------------------------------------------------
// linearisation shared by the solve and emit kernels
// ------------------------------------------------------------------------------------------------
struct Lin {            // per-problem linearisation constants (registers, computed redundantly per thread)
    double c, s;        // cos/sin of the mean reference yaw (mpc.cpp:153-178)
    double Iwi[9];      // inverse world inertia (mpc.cpp:220-221)
};

// mpc.cpp:153-162,173-178,220-221
__device__ __forceinline__ void linearise(const DevParams &P, const double *xref, int N, Lin &L)
{
    double yaw = 0.0;
    for (int i = 0; i < N; i++) yaw += xref[i * NS + 2];
    const double psi = yaw / N;
    sincos(psi, &L.s, &L.c);
    const double c = L.c, s = L.s;
    // T = [[c,s,0],[-s,c,0],[0,0,1]];  Iw = T Ib T'
    const double *I = P.Ib;
    double TI[9];
#pragma unroll
    for (int j = 0; j < 3; j++) {
        TI[0 + j] = c * I[0 + j] + s * I[3 + j];
        TI[3 + j] = -s * I[0 + j] + c * I[3 + j];
        TI[6 + j] = I[6 + j];
    }
    double Iw[9];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        Iw[i * 3 + 0] = TI[i * 3 + 0] * c + TI[i * 3 + 1] * s;
        Iw[i * 3 + 1] = -TI[i * 3 + 0] * s + TI[i * 3 + 1] * c;
        Iw[i * 3 + 2] = TI[i * 3 + 2];
    }
    const double a = Iw[0], b = Iw[1], cc = Iw[2], d = Iw[3], e = Iw[4], f = Iw[5], g = Iw[6], h = Iw[7], k = Iw[8];
    const double det = a * (e * k - f * h) - b * (d * k - f * g) + cc * (d * h - e * g);
    const double id = 1.0 / det;
    L.Iwi[0] = (e * k - f * h) * id; L.Iwi[1] = (cc * h - b * k) * id; L.Iwi[2] = (b * f - cc * e) * id;
    L.Iwi[3] = (f * g - d * k) * id; L.Iwi[4] = (a * k - cc * g) * id; L.Iwi[5] = (cc * d - a * f) * id;
    L.Iwi[6] = (d * h - e * g) * id; L.Iwi[7] = (b * g - a * h) * id; L.Iwi[8] = (a * e - b * d) * id;
}

// B_i(6:9, 3j:3j+3) = dt * Iw^-1 * skew(c_ij - p_i)      (mpc.cpp:164-171,229-237,251)
__device__ __forceinline__ void contact_Bw(const DevParams &P, const Lin &L, const double *cpt, const double *pc, double *Bw)
{
    const double r0 = cpt[0] - pc[0], r1 = cpt[1] - pc[1], r2 = cpt[2] - pc[2];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const double i0 = L.Iwi[a * 3 + 0], i1 = L.Iwi[a * 3 + 1], i2 = L.Iwi[a * 3 + 2];
        Bw[a * 3 + 0] = (i1 * r2 - i2 * r1) * P.dt;
        Bw[a * 3 + 1] = (i2 * r0 - i0 * r2) * P.dt;
        Bw[a * 3 + 2] = (i0 * r1 - i1 * r0) * P.dt;
    }
}

__device__ __forceinline__ double S2_entry(int a, int b, int N)
{   // sum_{i=max(a,b)}^{N-1} (i-a)(i-b)
    const int m = a > b ? a : b;
    double acc = 0.0;
    for (int i = m; i < N; i++) acc += (double)((i - a) * (i - b));
    return acc;
}

#include "srbd_solve_warp.cuh"
#include "node_adaptors.cuh"

// ------------------------------------------------------------------------------------------------
// emit kernel: materialise the reference's dense intermediates (mpc.cpp:196-376) -- HBM-write bound
// ------------------------------------------------------------------------------------------------
struct EmitPtrs {
    const double *x0, *x_ref, *contact, *c_hor, *p_com;
    double *A, *Bd, *Aqp, *Bqp, *H, *g, *Acon, *l, *u;
};

__global__ void __launch_bounds__(256) srbd_mpc_emit_kernel(const DevParams P, const EmitPtrs io, long long B, int N, int NC)
{
    extern __shared__ double sm[];
    const int NU = 3 * NC, NV = NU * N, NCN = NC * N, NX = NS * N, N6 = 6 * N, M5 = 5 * NCN;
    double *xref = sm;                 // [N*13]
    double *x0 = xref + N * NS;        // [16]
    double *Bw = x0 + 16;              // [NCN*9]   dt Iw^-1 skew(r)
    double *RB = Bw + NCN * 9;         // [NCN*9]   Rz * Bw
    double *E = RB + NCN * 9;          // [N*12]
    double *hv = E + N * 12;           // [N6]
    const int tid = threadIdx.x, nt = blockDim.x;
    const long long b = blockIdx.x;
    if (b >= B) return;
    for (int i = tid; i < N * NS; i += nt) xref[i] = io.x_ref[b * (N * NS) + i];
    for (int i = tid; i < NS; i += nt) x0[i] = io.x0[b * NS + i];
    __syncthreads();
    Lin lin;
    linearise(P, xref, N, lin);
    const double cz = lin.c, sz = lin.s, dt = P.dt, bv = P.bv;
    for (int c = tid; c < NCN; c += nt) {
        const int k = c / NC;
        double bw[9];
        contact_Bw(P, lin, io.c_hor + b * (NCN * 3) + 3 * c, io.p_com + b * (N * 3) + 3 * k, bw);
#pragma unroll
        for (int q = 0; q < 3; q++) {
            Bw[9 * c + q] = bw[q]; Bw[9 * c + 3 + q] = bw[3 + q]; Bw[9 * c + 6 + q] = bw[6 + q];
            RB[9 * c + q] = cz * bw[q] - sz * bw[3 + q];
            RB[9 * c + 3 + q] = sz * bw[q] + cz * bw[3 + q];
            RB[9 * c + 6 + q] = bw[6 + q];
        }
    }
    for (int i = tid; i < N; i += nt) {
        const double kk = (double)(i + 1), g0 = x0[12];
        double *Ei = E + 12 * i;
        const double *xr = xref + i * NS;
        const double wx = x0[6], wy = x0[7], wz = x0[8];
        Ei[0] = x0[0] + kk * dt * (cz * wx - sz * wy) - xr[0];
        Ei[1] = x0[1] + kk * dt * (sz * wx + cz * wy) - xr[1];
        Ei[2] = x0[2] + kk * dt * wz - xr[2];
        Ei[3] = x0[3] + kk * dt * x0[9] - xr[3];
        Ei[4] = x0[4] + kk * dt * x0[10] - xr[4];
        Ei[5] = x0[5] + kk * dt * x0[11] + (0.5 * kk * (kk - 1.0)) * dt * dt * g0 - xr[5];
        Ei[6] = wx - xr[6]; Ei[7] = wy - xr[7]; Ei[8] = wz - xr[8];
        Ei[9] = x0[9] - xr[9]; Ei[10] = x0[10] - xr[10]; Ei[11] = x0[11] + kk * dt * g0 - xr[11];
    }
    __syncthreads();
    for (int e = tid; e < N6; e += nt) {
        const int k = e / 6, al = e % 6;
        double acc = 0.0;
        for (int i = k; i < N; i++) {
            const double *Ei = E + 12 * i;
            const double lev = (double)(i - k) * dt;
            double a;
            if (al == 0) a = lev * (cz * P.q[0] * Ei[0] + sz * P.q[1] * Ei[1]) + P.q[6] * Ei[6];
            else if (al == 1) a = lev * (-sz * P.q[0] * Ei[0] + cz * P.q[1] * Ei[1]) + P.q[7] * Ei[7];
            else if (al == 2) a = lev * P.q[2] * Ei[2] + P.q[8] * Ei[8];
            else a = lev * P.q[al] * Ei[al] + P.q[6 + al] * Ei[6 + al];
            acc += a;
        }
        hv[e] = P.scale * acc;
    }
    __syncthreads();

    // A_discrete = I + dt Ac   (mpc.cpp:196-211)
    auto Apow = [&](int kpow, int r, int c) -> double {   // entry (r,c) of A^kpow
        double v = (r == c) ? 1.0 : 0.0;
        const double kd = (double)kpow * dt;
        if (r < 3 && c >= 6 && c < 9) {
            const int cc = c - 6;
            const double Rz = (r == 0) ? (cc == 0 ? cz : (cc == 1 ? -sz : 0.0))
                            : (r == 1) ? (cc == 0 ? sz : (cc == 1 ? cz : 0.0)) : (cc == 2 ? 1.0 : 0.0);
            v += kd * Rz;
        }
        if (r >= 3 && r < 6 && c == r + 6) v += kd;
        if (r == 11 && c == 12) v += kd;
        if (r == 5 && c == 12) v += 0.5 * (double)kpow * (double)(kpow - 1) * dt * dt;
        return v;
    };
    if (io.A) for (int e = tid; e < NS * NS; e += nt) io.A[b * (NS * NS) + e] = Apow(1, e / NS, e % NS);
    if (io.Aqp) for (int e = tid; e < NX * NS; e += nt) { const int row = e / NS, c = e % NS; io.Aqp[b * ((long long)NX * NS) + e] = Apow(row / NS + 1, row % NS, c); }
    // B_discrete_hor   (mpc.cpp:213-254)
    auto Bent = [&](int kap, int j, int r, int col) -> double {   // (A^kap B_j)(r, col)
        const int c = j * NC + col / 3, q = col % 3;
        if (r < 3) return (double)kap * dt * RB[9 * c + 3 * r + q];
        if (r < 6) return (r - 3 == q) ? (double)kap * dt * bv : 0.0;
        if (r < 9) return Bw[9 * c + 3 * (r - 6) + q];
        if (r < 12) return (r - 9 == q) ? bv : 0.0;
        return 0.0;
    };
    if (io.Bd) for (int e = tid; e < NX * NU; e += nt) { const int row = e / NU, col = e % NU; io.Bd[b * ((long long)NX * NU) + e] = Bent(0, row / NS, row % NS, col); }
    // Bqp   (mpc.cpp:272-291)
    if (io.Bqp) {
        double *out = io.Bqp + b * ((long long)NX * NV);
        for (long long e = tid; e < (long long)NX * NV; e += nt) {
            const int row = (int)(e / NV), col = (int)(e % NV), i = row / NS, j = col / NU;
            out[e] = (j <= i) ? Bent(i - j, j, row % NS, col % NU) : 0.0;
        }
    }
    // H = s (Bqp' Q Bqp + R)   (mpc.cpp:359-364), closed form per entry
    if (io.H) {
        double *out = io.H + b * ((long long)NV * NV);
        const double sc = P.scale, d2 = dt * dt;
        for (long long e = tid; e < (long long)NV * NV; e += nt) {
            const int I = (int)(e / NV), J = (int)(e % NV);
            const int a = I / NU, ca = a * NC + (I % NU) / 3, p = I % 3;
            const int bq = J / NU, cb = bq * NC + (J % NU) / 3, q = J % 3;
            const int mx = a > bq ? a : bq;
            const double s0 = (double)(N - mx), s2 = S2_entry(a, bq, N);
            double th = 0.0, om = 0.0;
#pragma unroll
            for (int d = 0; d < 3; d++) {
                th += P.q[d] * RB[9 * ca + 3 * d + p] * RB[9 * cb + 3 * d + q];
                om += P.q[6 + d] * Bw[9 * ca + 3 * d + p] * Bw[9 * cb + 3 * d + q];
            }
            double v = s2 * d2 * th + s0 * om;
            if (p == q) v += (s2 * d2 * P.q[3 + p] + s0 * P.q[9 + p]) * bv * bv;
            v *= sc;
            if (I == J) v += sc * P.r_weight;
            out[e] = v;
        }
    }
    if (io.g) {
        for (int e = tid; e < NV; e += nt) {
            const int k = e / NU, c = k * NC + (e % NU) / 3, p = e % 3;
            io.g[b * NV + e] = Bw[9 * c + p] * hv[6 * k] + Bw[9 * c + 3 + p] * hv[6 * k + 1] + Bw[9 * c + 6 + p] * hv[6 * k + 2] + bv * hv[6 * k + 3 + p];
        }
    }
    // Ac (dense) and bounds   (mpc.cpp:293-357)
    if (io.Acon) {
        double *out = io.Acon + b * ((long long)M5 * NV);
        for (long long e = tid; e < (long long)M5 * NV; e += nt) {
            const int row = (int)(e / NV), col = (int)(e % NV);
            double v = 0.0;
            if (row / 5 == col / 3) {
                const int r = row % 5, q = col % 3;
                if (r < 2) v = q == 0 ? 1.0 : (q == 2 ? (r == 0 ? P.mu : -P.mu) : 0.0);
                else if (r < 4) v = q == 1 ? 1.0 : (q == 2 ? (r == 2 ? P.mu : -P.mu) : 0.0);
                else v = q == 2 ? 1.0 : 0.0;
            }
            out[e] = v;
        }
    }
    if (io.l && io.u) {
        const double inf = __longlong_as_double(0x7ff0000000000000LL);
        for (int c = tid; c < NCN; c += nt) {
            const double cf = io.contact[b * NCN + c];
            double *l = io.l + b * M5 + 5 * c, *u = io.u + b * M5 + 5 * c;
            l[0] = 0.0; l[1] = -inf; l[2] = 0.0; l[3] = -inf; l[4] = P.fz_min * cf;
            u[0] = inf; u[1] = 0.0; u[2] = inf; u[3] = 0.0; u[4] = P.fz_max * cf;
        }
    }
}

// computeRollout (mpc.cpp:427-446) of a given u_opt: one thread per problem, the recurrence as the reference runs it
// (x_i = A x_{i-1} + B_{i-1} u_{i-1} with A = I + dt Ac, mpc.cpp:196-211, and the B_i of mpc.cpp:213-254).
__global__ void __launch_bounds__(128) srbd_rollout_kernel(const DevParams P, const double *x_ref, const double *c_hor, const double *p_com,
                                                            const double *u_opt, const double *x_current, int one_rollout, double *x_opt,
                                                            long long B, int N, int NC)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    Lin lin;
    linearise(P, x_ref + b * (N * NS), N, lin);
    double *xo = x_opt + b * (N * NS);
    double x[NS];
    for (int r = 0; r < NS; r++) x[r] = x_current ? x_current[b * NS + r] : xo[r];
    if (x_current) for (int r = 0; r < NS; r++) xo[r] = x[r];                  // mpc.cpp:430-433
    const int last = one_rollout ? 1 : N - 1;
    for (int i = 1; i <= last; i++) {                                           // mpc.cpp:435-445
        double tq[3] = {0.0, 0.0, 0.0}, fs[3] = {0.0, 0.0, 0.0};
        for (int j = 0; j < NC; j++) {
            const long long c = (b * N + (i - 1)) * NC + j;
            double bw[9];
            contact_Bw(P, lin, c_hor + 3 * c, p_com + 3 * (b * N + (i - 1)), bw);
            const double *f = u_opt + 3 * c;
#pragma unroll
            for (int a = 0; a < 3; a++) { tq[a] += bw[3 * a] * f[0] + bw[3 * a + 1] * f[1] + bw[3 * a + 2] * f[2]; fs[a] += f[a]; }
        }
        double y[NS];
        y[0] = x[0] + P.dt * (lin.c * x[6] - lin.s * x[7]);                     // Ac(0:3,6:9) = Rz(psi)  (quirk Q2)
        y[1] = x[1] + P.dt * (lin.s * x[6] + lin.c * x[7]);
        y[2] = x[2] + P.dt * x[8];
        y[3] = x[3] + P.dt * x[9]; y[4] = x[4] + P.dt * x[10]; y[5] = x[5] + P.dt * x[11];
        y[6] = x[6] + tq[0]; y[7] = x[7] + tq[1]; y[8] = x[8] + tq[2];
        y[9] = x[9] + P.bv * fs[0]; y[10] = x[10] + P.bv * fs[1]; y[11] = x[11] + P.dt * x[12] + P.bv * fs[2];
        y[12] = x[12];
        for (int r = 0; r < NS; r++) { x[r] = y[r]; xo[i * NS + r] = y[r]; }
    }
}

// FP64 FMA peak microbenchmark
__global__ void fp64_peak_kernel(double *out, int iters)
{
    double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double m = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; i++) {
        a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
        a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
#define CUDA_TRY(expr)                                                                           \
    do {                                                                                         \
        cudaError_t e_ = (expr);                                                                 \
        if (e_ != cudaSuccess) {                                                                 \
            g_last_error = std::string(#expr) + ": " + cudaGetErrorString(e_);                   \
            return G1MPC_ERR_CUDA;                                                               \
        }                                                                                        \
    } while (0)

// One entry per compiled (horizon, num_contacts): the kernel's address (attributes, occupancy) and a launcher with a
// shape-independent signature (the kernels take their layout object by value, so their own signatures differ).
typedef void (*launch_fn)(unsigned grid, unsigned block, size_t smem, cudaStream_t st, const DevParams &, const BatchPtrs &, long long B,
                          const v2::DLay &);
template <class LAY, int HL, int WSG>
void launch_static(unsigned grid, unsigned block, size_t smem, cudaStream_t st, const DevParams &P, const BatchPtrs &io, long long B,
                   const v2::DLay &)
{
    v2::srbd_mpc_solve_warp_kernel<LAY, HL, WSG><<<grid, block, smem, st>>>(P, io, B, LAY());
}
// any other shape: the same kernel source with the layout in registers
template <int WSG, int HL = 32>
void launch_dynamic(unsigned grid, unsigned block, size_t smem, cudaStream_t st, const DevParams &P, const BatchPtrs &io, long long B,
                    const v2::DLay &L)
{
    v2::srbd_mpc_solve_warp_kernel<v2::DLay, HL, WSG><<<grid, block, smem, st>>>(P, io, B, L);
}
template <class LAY, int HL>
void launch_lat(unsigned grid, unsigned block, size_t smem, cudaStream_t st, const DevParams &P, const BatchPtrs &io, long long B,
                const v2::DLay &)
{
    v2::srbd_mpc_solve_lat_kernel<LAY, HL><<<grid, block, smem, st>>>(P, io, B, LAY());
}
typedef void (*launch_refill_fn)(unsigned grid, size_t smem, cudaStream_t st, const DevParams &, const BatchPtrs &, long long B,
                                 unsigned long long *next);
template <class LAY>
void launch_refill(unsigned grid, size_t smem, cudaStream_t st, const DevParams &P, const BatchPtrs &io, long long B, unsigned long long *next)
{
    v2::srbd_mpc_solve_refill_kernel<LAY, 16><<<grid, 32, smem, st>>>(P, io, B, LAY(), next);
}
struct KernelEntry {
    int N, NC;
    const void *fn; launch_fn launch; size_t smem;      // workspace in shared memory; smem = bytes per problem
    const void *fn2; launch_fn launch2;                 // ... two problems per warp (or null)
    const void *fng; launch_fn launchg;                 // workspace in global memory, persistent grid (or null)
    const void *fnl; launch_fn launchl;                 // partner-warp mode: one problem per two-warp CTA (or null)
    const void *fnl2; launch_fn launchl2;               // ... two problems per two-warp CTA (or null)
    const void *fnh; launch_fn launchh;                 // hybrid workspace: per-stage region in shared, per-contact region in global memory (or null)
    const void *fnh2; launch_fn launchh2;               // ... two problems per warp (or null)
    const void *fnr; launch_refill_fn launchr;          // two problems per warp with refill, persistent grid (or null)
};
// the two-problems-per-warp layout is compiled for the shapes whose stages fit the 16 lanes of a half-warp in one round
template <int N, int NC>
constexpr bool two_per_warp() { return N <= 16 && NC == 4 && 2 * v2::WLay<N, NC>::bytes <= 227 * 1024; }
// the global-workspace variant is compiled for the shapes of which fewer than four fit the shared memory of an SM
template <int N, int NC>
constexpr bool global_ws() { return 4 * v2::WLay<N, NC>::bytes > 227 * 1024; }
// the hybrid variant (per-stage region in shared memory, per-contact region in global memory) is compiled for the same shapes
template <int N, int NC>
constexpr bool hybrid_ws() { return global_ws<N, NC>(); }
// ... and with two problems per warp (the stages of a problem must fit the 16 lanes of a half-warp in one round)
template <int N, int NC>
constexpr bool hybrid2() { return N >= 10 && N <= 16 && !(N == 10 && NC == 4); }      // (bench shape: measured, loses 25 %)
template <int N, int NC>
KernelEntry make_entry()
{
    using LAY = v2::WLay<N, NC>;
    KernelEntry e{N, NC, (const void *)v2::srbd_mpc_solve_warp_kernel<LAY, 32, false>, launch_static<LAY, 32, false>, LAY::bytes,
                  nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    if constexpr (two_per_warp<N, NC>()) {
        e.fn2 = (const void *)v2::srbd_mpc_solve_warp_kernel<LAY, 16, false>;
        e.launch2 = launch_static<LAY, 16, false>;
        if constexpr (N == 10) {         // (the bench shape)
            e.fnr = (const void *)v2::srbd_mpc_solve_refill_kernel<LAY, 16>;
            e.launchr = launch_refill<LAY>;
        }
    }
    if constexpr (two_per_warp<N, NC>() && N <= 15) {                   // (one named barrier per stage: N <= 15)
        e.fnl = (const void *)v2::srbd_mpc_solve_lat_kernel<LAY, 32>;   // the short-horizon shapes: where latency is the metric
        e.launchl = launch_lat<LAY, 32>;
        if constexpr (N == 10) {        // (layout 4, measured and never chosen automatically: kept for the bench shape only)
            e.fnl2 = (const void *)v2::srbd_mpc_solve_lat_kernel<LAY, 16>;
            e.launchl2 = launch_lat<LAY, 16>;
        }
    }
    if constexpr (global_ws<N, NC>()) {
        e.fng = (const void *)v2::srbd_mpc_solve_warp_kernel<LAY, 32, 1>;
        e.launchg = launch_static<LAY, 32, 1>;
    }
    if constexpr (hybrid_ws<N, NC>()) {
        e.fnh = (const void *)v2::srbd_mpc_solve_warp_kernel<LAY, 32, 2>;
        e.launchh = launch_static<LAY, 32, 2>;
    }
    if constexpr (hybrid2<N, NC>()) {
        e.fnh2 = (const void *)v2::srbd_mpc_solve_warp_kernel<LAY, 16, 2>;
        e.launchh2 = launch_static<LAY, 16, 2>;
    }
    return e;
}

#ifdef G1MPC_DEV_SHAPE   /* development builds: only the bench shape */
#define G1MPC_SHAPES(X) X(10, 4)
#else
#define G1MPC_SHAPES(X) X(10, 4) X(15, 4) X(20, 4) X(40, 4) X(10, 8) X(12, 8) X(20, 8) X(40, 8) X(5, 4) X(8, 4)
#endif
#define X_ENTRY(n, c) make_entry<n, c>(),
const KernelEntry kKernels[] = {G1MPC_SHAPES(X_ENTRY)};
#undef X_ENTRY

} // namespace

// ------------------------------------------------------------------------------------------------
// Scratch of the no-allocation entry point (g1mpc_solve_batch_static_ws, below): the global-workspace variant of the solve
// kernel needs one working-set slot per resident warp; here the slots lie in this file-scope array instead of a cudaMalloc'ed
// block, so the call allocates nothing.  Every GPU that loads the library holds the array's G1MPC_STATIC_WS_BYTES until the
// process exits.  In HOST code the name `g_static_ws` is a host-side stand-in, not a device address: the address the kernel
// gets is the one cudaGetSymbolAddress returns, looked up once per device with that device current (static_ws_setup, called
// from g1mpc_create).
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) double g_static_ws[G1MPC_STATIC_WS_BYTES / sizeof(double)];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_sws_mutex;
bool g_sws_done[kMaxDevices];
double *g_sws_ptr[kMaxDevices];         // DEVICE address of g_static_ws on device d (null: look-up failed, see g_sws_err)
cudaError_t g_sws_err[kMaxDevices];

// once per device; `device` must be the current device
double *static_ws_setup(int device, cudaError_t *err)
{
    if (device < 0 || device >= kMaxDevices) { *err = cudaErrorInvalidDevice; return nullptr; }
    std::lock_guard<std::mutex> lock(g_sws_mutex);
    if (!g_sws_done[device]) {
        void *ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_ws);
        g_sws_err[device] = e;
        g_sws_ptr[device] = e == cudaSuccess ? (double *)ptr : nullptr;
        g_sws_done[device] = true;
    }
    *err = g_sws_err[device];
    return g_sws_ptr[device];
}
} // namespace

struct g1mpc_handle {
    g1mpc_params params;
    DevParams dev;
    int device;
    const KernelEntry *kernel;
    KernelEntry dyn_entry;                 // used when no compiled shape matches: run-time layout, one problem per warp
    v2::DLay dlay;
    // global-workspace variant (long horizons): workspace of ws_warps slots, launches serialised on it by ws_event
    bool use_global = false, fits_smem = true;
    long long global_above = 0;             // launches of more problems than this use the global workspace (auto mode)
    double *d_ws = nullptr;
    int ws_warps = 0, ws_ctas_per_sm = 0;
    // hybrid workspace (stance list, per-stage region and scratch in shared memory, per-contact region in a slot of d_ws)
    bool use_hybrid = false;
    size_t hy_smem = 0;
    int hy_warps = 0, hy_ctas_per_sm = 0, hy_ppw = 1;
    const void *hy_fn = nullptr; launch_fn hy_launch = nullptr;
    bool hy_whole_only = false;             // not for the chunks of a host-API batch (they would take turns on the one workspace)
    long long hy_from = 0, hy_upto = 0, hy_also_from = 0;     // launches of hy_from < B <= hy_upto problems, or of B >= hy_also_from, use it
    cudaEvent_t ws_event = nullptr;
    int warps_per_block = 1;
    int ctas_per_sm = 0;
    // two problems per warp pay off once a scheduler would otherwise hold two warps: above 4 problems per SM
    // (G1MPC_PPW=1|2 in the environment forces one layout -- tests compare the two bit for bit)
    long long two_above = 0, lat_upto = 0;      // launches of at most lat_upto problems run the latency-mode kernel
    int lat2 = 0;                               // partner-warp kernel with two problems per warp: 0 off, 1 in place of the plain one, 2 always
    int force_ppw = 0;
    // two problems per warp with refill: a counter per launch in flight (round robin), persistent grid of refill_ctas CTAs
    static constexpr int kRefillCounters = 64;
    unsigned long long *d_refill = nullptr;
    int refill_k = 0, refill_ctas = 0;
    long long refill_above = 0x7fffffffffffLL;
    int host_warm = 0, host_warm_shift = 1;     // g1mpc_set_host_warm_start
    // host-API staging: kSlots batches in flight, each cut into up to kPipe chunks with a stream of their own
    static constexpr int kPipe = 4, kSlots = G1MPC_HOST_SLOTS;
    static constexpr int64_t kZeroCopyMax = 4;  // host-API batches up to this size run without copies (see g1mpc_update_host_begin)
    struct Off { size_t x0, xr, ct, ch, pc, xc, xo, u, res, st; };
    struct Slot {
        cudaStream_t pipe[kPipe] = {nullptr, nullptr, nullptr, nullptr};
        char *h_stage = nullptr, *d_stage = nullptr;
        size_t stage_bytes = 0, stride = 0, cap = 0, nbs[kPipe] = {0, 0, 0, 0};
        Off off[kPipe];
        int nchunk = 0;
        bool pending = false, staged_out = false;
        int64_t prev_B = -1;                    // batch size of the solution still lying in d_stage (host warm start)
        bool prev_zero_copy = false;            // ... or in h_stage (zero-copy path)
        double *u_opt = nullptr, *x_opt = nullptr, *resid = nullptr;     // caller's output buffers (unpacked in _end)
        int32_t *status = nullptr, *iters = nullptr;
    } slot[kSlots];
    // closed-loop workspace
    char *d_loop = nullptr;
    size_t loop_bytes = 0;    // no-allocation entry point (g1mpc_solve_batch_static_ws): device address of the static scratch array on this handle's
    // device, the global-workspace kernel it runs and that kernel's resident warps per SM -- all looked up in g1mpc_create
    double *sws_ptr = nullptr;
    cudaError_t sws_err = cudaSuccess;
    launch_fn sws_launch = nullptr;
    int sws_ctas_per_sm = 0, sws_sms = 0;
};

extern "C" {

const char *g1mpc_last_error(void) { return g_last_error.c_str(); }
int g1mpc_abi_version(void) { return G1MPC_ABI_VERSION; }

int g1mpc_create(const g1mpc_params *p, int device, g1mpc_handle **out)
{
    if (!p || !out) { g_last_error = "null argument"; return G1MPC_ERR_INVALID_ARGUMENT; }
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        g_last_error = "no CUDA device: g1mpc_b200 has no CPU fallback";
        return G1MPC_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= ndev) { g_last_error = "bad device index"; return G1MPC_ERR_INVALID_ARGUMENT; }
    if (p->horizon < 2 || p->num_contacts < 1 || !(p->dt > 0) || !(p->mass > 0) || !(p->mu > 0) ||
        !(p->fz_max > p->fz_min) || !(p->fz_min >= 0) || !(p->r_weight > 0) || !(p->cost_scale > 0)) {
        g_last_error = "invalid MPC parameters";
        return G1MPC_ERR_INVALID_ARGUMENT;
    }
    for (int i = 0; i < 12; i++) if (!(p->q_weights[i] > 0)) { g_last_error = "q_weights[0..11] must be positive"; return G1MPC_ERR_INVALID_ARGUMENT; }
    if (p->q_weights[12] != 0.0) {       // the gravity state is a constant, its weight is zero in both presets (mpc.cpp:96, mpc.py:50)
        g_last_error = "q_weights[12] (gravity state) must be 0: the state is constant and is not part of the tracked cost";
        return G1MPC_ERR_INVALID_ARGUMENT;
    }
    if (p->layout < 0 || p->layout > 5) { g_last_error = "layout must be 0 (auto), 1, 2, 3, 4 or 5"; return G1MPC_ERR_INVALID_ARGUMENT; }
    const KernelEntry *ke = nullptr;
    for (const auto &k : kKernels) if (k.N == p->horizon && k.NC == p->num_contacts) ke = &k;
    const v2::DLay dl(p->horizon, p->num_contacts);
    if (!dl.valid()) {
        g_last_error = "unsupported shape: needs horizon >= 2 and 1 <= num_contacts <= 13 (the step vectors alias the stage storage)";
        return G1MPC_ERR_UNSUPPORTED_SHAPE;
    }
    CUDA_TRY(cudaSetDevice(device));
    int max_smem = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    if (p->workspace < 0 || p->workspace > 3) { g_last_error = "workspace must be 0 (auto), 1 (shared), 2 (global) or 3 (hybrid)"; return G1MPC_ERR_INVALID_ARGUMENT; }
    const bool fits_smem = dl.bytes <= (size_t)max_smem;
    if (!fits_smem && p->workspace == 1) {
        g_last_error = "shape exceeds shared memory: (horizon, num_contacts) = (" + std::to_string(p->horizon) + ", " + std::to_string(p->num_contacts) +
                       ") needs " + std::to_string(dl.bytes) + " B per problem, the device allows " + std::to_string(max_smem) +
                       " B per CTA; use workspace = 0 or 2 (global-memory workspace)";
        return G1MPC_ERR_UNSUPPORTED_SHAPE;
    }
    g1mpc_handle *h = new g1mpc_handle();
    h->params = *p; h->device = device; h->dlay = dl;
    if (!ke) {
        // no compile-time instantiation for this shape: the run-time-layout instantiation of the same kernel
        h->dyn_entry = KernelEntry{p->horizon, p->num_contacts, (const void *)v2::srbd_mpc_solve_warp_kernel<v2::DLay, 32, false>, launch_dynamic<false>,
                                   dl.bytes, nullptr, nullptr, (const void *)v2::srbd_mpc_solve_warp_kernel<v2::DLay, 32, true>, launch_dynamic<true>,
                                   nullptr, nullptr, nullptr, nullptr, (const void *)v2::srbd_mpc_solve_warp_kernel<v2::DLay, 32, 2>, launch_dynamic<2>, nullptr, nullptr, nullptr, nullptr};
        // two problems per warp where the stages of a problem fit the 16 lanes of a half-warp in one round (as for the compiled shapes)
        if (p->horizon <= 16 && p->num_contacts == 4 && 2 * dl.bytes <= (size_t)max_smem) {
            h->dyn_entry.fn2 = (const void *)v2::srbd_mpc_solve_warp_kernel<v2::DLay, 16, 0>;
            h->dyn_entry.launch2 = launch_dynamic<0, 16>;
        }
        ke = &h->dyn_entry;
    }
    // workspace in global memory: forced, needed (does not fit shared memory), or -- auto -- where fewer than four problems fit an SM
    h->use_global = ke->fng && p->workspace != 3 && (p->workspace == 2 || !fits_smem || (p->workspace == 0 && 4 * dl.bytes > (size_t)max_smem));
    h->fits_smem = fits_smem;
    if (p->workspace == 2 && !ke->fng) { g_last_error = "no global-workspace kernel compiled for this shape"; delete h; return G1MPC_ERR_UNSUPPORTED_SHAPE; }
    h->kernel = ke;
    if (!(h->params.gap_tol > 0)) h->params.gap_tol = 1e-10;
    if (h->params.max_iter <= 0) h->params.max_iter = 40;
    if (!(h->params.dual_tol > 0)) h->params.dual_tol = 1e-6;
    if (!(h->params.warm_margin > 0)) h->params.warm_margin = 0.01 * p->mass * fabs(p->gravity);
    if (!(h->params.warm_mu0 > 0)) h->params.warm_mu0 = 0.1;
    DevParams &d = h->dev;
    d.mu = p->mu; d.fz_min = p->fz_min; d.fz_max = p->fz_max; d.dt = p->dt;
    d.bv = (1.0 / p->mass) * p->dt;   // (I/m)*dt as in mpc.cpp:238,251
    d.weight = p->mass * fabs(p->gravity);
    d.scale = p->cost_scale; d.r_weight = p->r_weight; d.sr = p->cost_scale * p->r_weight;
    d.gap_tol = h->params.gap_tol; d.max_iter = h->params.max_iter;
    d.dual_tol = h->params.dual_tol; d.warm_margin = h->params.warm_margin; d.warm_mu0 = h->params.warm_mu0;
    memcpy(d.Ib, p->inertia_body, sizeof d.Ib);
    memcpy(d.q, p->q_weights, sizeof d.q);
    cudaError_t e = cudaSuccess;
    if (fits_smem) {
        e = cudaFuncSetAttribute(ke->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(ke->smem * h->warps_per_block));
        if (e != cudaSuccess) { g_last_error = std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e); g1mpc_destroy(h); return G1MPC_ERR_CUDA; }
    }
    if (h->use_global) {
        int sms = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        cudaFuncSetAttribute(ke->fng, cudaFuncAttributePreferredSharedMemoryCarveout, 0);       // all of the on-chip SRAM as L1
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&h->ws_ctas_per_sm, ke->fng, kWarpThreads, 0);
        if (h->ws_ctas_per_sm < 1) h->ws_ctas_per_sm = 1;
        h->ws_warps = sms * h->ws_ctas_per_sm;
        e = cudaMalloc((void **)&h->d_ws, (size_t)h->ws_warps * dl.bytes);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ws_event, cudaEventDisableTiming);
        if (e != cudaSuccess) { g_last_error = std::string("global workspace: ") + cudaGetErrorString(e); g1mpc_destroy(h); return G1MPC_ERR_CUDA; }
    }
    cudaFuncSetAttribute((const void *)srbd_mpc_emit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    if (fits_smem) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&h->ctas_per_sm, ke->fn, kWarpThreads * h->warps_per_block, ke->smem * h->warps_per_block);
    if (h->use_global && fits_smem && p->workspace == 0) {
        // auto: a launch that the shared-memory kernel holds in one wave stays on chip (measured equal or a few per cent
        // faster there, profiles/r2_c_global_workspace.txt); larger ones gain from the eight warps per SM of the global variant
        int sms = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        h->global_above = (long long)sms * h->ctas_per_sm;
    }
    if (ke->fn2) {
        e = cudaFuncSetAttribute(ke->fn2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * ke->smem * h->warps_per_block));
        if (e != cudaSuccess) { g_last_error = std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e); g1mpc_destroy(h); return G1MPC_ERR_CUDA; }
    }
    if (ke->fnr && (p->layout == 0 || p->layout == 5)) {
        int sms = 0, n = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        e = cudaFuncSetAttribute(ke->fnr, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * ke->smem));
        if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, ke->fnr, kWarpThreads, 2 * ke->smem);
        if (e == cudaSuccess) e = cudaMalloc((void **)&h->d_refill, sizeof(unsigned long long) * g1mpc_handle::kRefillCounters);
        if (e != cudaSuccess) { g_last_error = std::string("refill kernel: ") + cudaGetErrorString(e); g1mpc_destroy(h); return G1MPC_ERR_CUDA; }
        h->refill_ctas = sms * (n > 0 ? n : 1);
        h->refill_above = p->layout == 5 ? -1 : 0x7fffffffffffLL;
    }
    {
        int sms = 0, ctas2 = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        h->two_above = 4LL * sms;
        // ... and only where it does not cost residency: two-problem CTAs must hold as many problems per SM as one-problem CTAs do
        // (N = 15: 2 x 84 KB against 5 x 42 KB per SM -> one problem per warp stays)
        if (ke->fn2) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas2, ke->fn2, kWarpThreads * h->warps_per_block, 2 * ke->smem * h->warps_per_block);
        if (2 * ctas2 < h->ctas_per_sm) h->two_above = 0x7fffffffffffLL;
        h->force_ppw = (p->layout == 1 || p->layout == 2) ? p->layout : 0;
        if (ke->fnl2 && p->layout == 4) {       // (never chosen automatically: measured slower than the plain two-per-warp layout)
            e = cudaFuncSetAttribute(ke->fnl2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * ke->smem));
            if (e != cudaSuccess) { g_last_error = std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e); g1mpc_destroy(h); return G1MPC_ERR_CUDA; }
            h->lat2 = 2;
        }
        // latency mode: two warps per problem while every warp still has a scheduler to itself (2 problems per SM); layout = 3 forces
        // it for any batch, layout = 1 / 2 switch it off
        if (ke->fnl && (p->layout == 0 || p->layout == 3)) {
            e = cudaFuncSetAttribute(ke->fnl, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ke->smem);
            if (e != cudaSuccess) { g_last_error = std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e); g1mpc_destroy(h); return G1MPC_ERR_CUDA; }
            h->lat_upto = p->layout == 3 ? 0x7fffffffLL : 2LL * sms;
        }
    }
    // Two problems per warp from the hybrid workspace: forced by workspace = 3 with layout = 2; automatic on the short-horizon shapes
    // for which two problems per warp do NOT pay from shared memory alone, because two working sets per CTA leave half the
    // schedulers without a warp (N = 15 / nc = 4: 2 x 42 KB; nc = 8: 2 x 40-48 KB) -- with only the stage factors on chip four to
    // six such CTAs fit (profiles/r2_g_hybrid_workspace.txt: +10 to +20 % from 1,024 problems on; on the bench shape, where the
    // shared-memory two-per-warp layout holds eight problems per SM, it loses 25 % and is not used).
    const bool smem_two = ke->fn2 && h->two_above < 0x7fffffffffffLL;
    const bool hy2_auto = p->workspace == 0 && p->layout == 0 && ke->fnh2 && !h->use_global && !smem_two;
    const bool hy2 = (p->workspace == 3 && p->layout == 2 && ke->fnh2) || hy2_auto;
    h->hy_fn = hy2 ? ke->fnh2 : ke->fnh; h->hy_launch = hy2 ? ke->launchh2 : ke->launchh; h->hy_ppw = hy2 ? 2 : 1;
    if (h->hy_fn && (p->workspace == 3 || hy2_auto || (p->workspace == 0 && h->use_global))) {
        // hybrid: stance list + per-stage region + scratch (oSt .. oEnd) in shared memory, per-contact region (0 .. oSt) in a global slot
        // per resident warp.  workspace = 3: always.  Auto, on the shapes that have the global-workspace variant (fewer than four problems
        // fit the shared memory of an SM) -- measured, profiles/r2_g_hybrid_workspace.txt: a launch that the shared-memory kernel does NOT
        // hold in one wave but this one does runs 20-40 % faster here than from the global workspace; beyond one wave it wins (10-15 %)
        // only where it keeps at least five warps per SM, and from two waves on (a second wave of a few problems is a long tail).
        const size_t stage_bytes = sizeof(double) * (size_t)(dl.oEnd - dl.oSt) * h->hy_ppw;
        int sms = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        if (stage_bytes <= (size_t)max_smem) {
            e = cudaFuncSetAttribute(h->hy_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stage_bytes);
            if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&h->hy_ctas_per_sm, h->hy_fn, kWarpThreads, stage_bytes);
            if (h->hy_ctas_per_sm < 1) h->hy_ctas_per_sm = 1;
            h->hy_warps = sms * h->hy_ctas_per_sm;
            const size_t need = (size_t)h->hy_warps * h->hy_ppw * sizeof(double) * (size_t)dl.oSt;
            if (e == cudaSuccess && !h->d_ws) {         // (a handle with the global workspace has the larger block already)
                e = cudaMalloc((void **)&h->d_ws, need);
                if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ws_event, cudaEventDisableTiming);
            }
            if (e != cudaSuccess) { g_last_error = std::string("hybrid workspace: ") + cudaGetErrorString(e); g1mpc_destroy(h); return G1MPC_ERR_CUDA; }
            h->use_hybrid = true; h->hy_smem = stage_bytes;
            if (p->workspace == 3) { h->hy_from = -1; h->hy_upto = 0x7fffffffffffLL; h->hy_also_from = 0; }
            else if (hy2_auto) { h->hy_from = 4LL * sms; h->hy_upto = 0x7fffffffffffLL; h->hy_also_from = 0x7fffffffffffLL; h->hy_whole_only = true; }
            else {
                int shared_ctas = 0;
                if (fits_smem) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&shared_ctas, ke->fn, kWarpThreads, ke->smem);
                h->hy_from = (long long)sms * shared_ctas;                       // (= global_above: up to there the shared-memory kernel runs)
                h->hy_upto = h->hy_ctas_per_sm > shared_ctas ? (long long)h->hy_warps : h->hy_from;
                h->hy_also_from = h->hy_ctas_per_sm >= 5 ? 2LL * h->hy_warps : 0x7fffffffffffLL;
            }
        } else if (p->workspace == 3) {
            g_last_error = "hybrid workspace: the per-stage region of this shape exceeds shared memory";
            g1mpc_destroy(h); return G1MPC_ERR_UNSUPPORTED_SHAPE;
        }
    } else if (p->workspace == 3) {
        g_last_error = "no hybrid-workspace kernel compiled for this shape";
        g1mpc_destroy(h); return G1MPC_ERR_UNSUPPORTED_SHAPE;
    }
    for (int sl = 0; sl < g1mpc_handle::kSlots; sl++)
        for (int i = 0; i < g1mpc_handle::kPipe; i++) {
            e = cudaStreamCreateWithFlags(&h->slot[sl].pipe[i], cudaStreamNonBlocking);
            if (e != cudaSuccess) { g_last_error = "cudaStreamCreate failed"; g1mpc_destroy(h); return G1MPC_ERR_CUDA; }
        }
    {
        // set-up of the no-allocation entry point: address of the static scratch array on this device, and the kernel variant
        // with the workspace in global memory (the compile-time instantiation where the shape has one, else the run-time layout)
        h->sws_ptr = static_ws_setup(device, &h->sws_err);
        const void *fn = h->kernel->fng ? h->kernel->fng : (const void *)v2::srbd_mpc_solve_warp_kernel<v2::DLay, 32, true>;
        h->sws_launch = h->kernel->fng ? h->kernel->launchg : launch_dynamic<true>;
        cudaDeviceGetAttribute(&h->sws_sms, cudaDevAttrMultiProcessorCount, device);
        cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, 0);             // (no shared memory: all of the SRAM as L1)
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&h->sws_ctas_per_sm, fn, kWarpThreads, 0);
        if (h->sws_ctas_per_sm < 1) h->sws_ctas_per_sm = 1;
    }
    *out = h;
    return G1MPC_OK;
}

void g1mpc_destroy(g1mpc_handle *h)
{
    if (!h) return;
    cudaSetDevice(h->device);
    for (int sl = 0; sl < g1mpc_handle::kSlots; sl++) {
        g1mpc_handle::Slot &S = h->slot[sl];
        for (int i = 0; i < g1mpc_handle::kPipe; i++) if (S.pipe[i]) { cudaStreamSynchronize(S.pipe[i]); cudaStreamDestroy(S.pipe[i]); }
        if (S.h_stage) cudaFreeHost(S.h_stage);
        if (S.d_stage) cudaFree(S.d_stage);
    }
    if (h->d_loop) cudaFree(h->d_loop);
    if (h->d_ws) cudaFree(h->d_ws);
    if (h->d_refill) cudaFree(h->d_refill);
    if (h->ws_event) cudaEventDestroy(h->ws_event);
    delete h;
}

int g1mpc_kernel_info(const g1mpc_handle *h, int32_t *smem_bytes, int32_t *threads, int32_t *ctas_per_sm)
{
    if (!h) return G1MPC_ERR_INVALID_ARGUMENT;
    if (smem_bytes) *smem_bytes = (h->use_hybrid && h->hy_from < 0) ? (int32_t)h->hy_smem : h->use_global ? 0 : (int32_t)(h->kernel->smem * h->warps_per_block);
    if (threads) *threads = kWarpThreads * h->warps_per_block;
    if (ctas_per_sm) *ctas_per_sm = (h->use_hybrid && h->hy_from < 0) ? h->hy_ctas_per_sm : h->use_global ? h->ws_ctas_per_sm : h->ctas_per_sm;
    return G1MPC_OK;
}

// `load` = number of problems that will be on the device together with this launch (the host API cuts a batch into chunks
// on several streams): it selects the lane layout, the results do not depend on it
static int solve_batch_impl(g1mpc_handle *h, int64_t B, int64_t load, const double *x0, const double *x_ref, const double *contact,
                            const double *c_horizon, const double *p_com, const double *x_current, int one_rollout,
                            double *u_opt, double *x_opt, int32_t *status, int32_t *iters, double *resid, double *y_opt,
                            void *stream, const double *u_warm = nullptr, int warm_shift = 0)
{
    if (!h || B < 0 || (B > 0 && (!x0 || !x_ref || !contact || !c_horizon || !p_com || !u_opt || !x_opt))) {
        g_last_error = "g1mpc_solve_batch: null argument"; return G1MPC_ERR_INVALID_ARGUMENT;
    }
    if (B == 0) return G1MPC_OK;
    if (B > 0x7fffffffLL) { g_last_error = "batch too large for one launch"; return G1MPC_ERR_INVALID_ARGUMENT; }
    CUDA_TRY(cudaSetDevice(h->device));
    if (warm_shift < 0) { g_last_error = "warm_shift must be >= 0"; return G1MPC_ERR_INVALID_ARGUMENT; }
    BatchPtrs io{x0, x_ref, contact, c_horizon, p_com, x_current, u_opt, x_opt, resid, y_opt, status, iters, one_rollout, u_warm, warm_shift, h->d_ws};
    if (h->use_hybrid && ((load > h->hy_from && load <= h->hy_upto) || load >= h->hy_also_from) && (!h->hy_whole_only || B == load)) {
        cudaStream_t st = (cudaStream_t)stream;
        CUDA_TRY(cudaStreamWaitEvent(st, h->ws_event, 0));
        const long long nw = (B + h->hy_ppw - 1) / h->hy_ppw, want = nw < h->hy_warps ? nw : h->hy_warps;
        h->hy_launch((unsigned)want, kWarpThreads, h->hy_smem, st, h->dev, io, (long long)B, h->dlay);
        CUDA_TRY(cudaGetLastError());
        CUDA_TRY(cudaEventRecord(h->ws_event, st));
        return G1MPC_OK;
    }
    if (h->use_global && load > h->global_above) {
        // persistent grid over the handle's workspace; launches on different streams take turns on it
        cudaStream_t st = (cudaStream_t)stream;
        CUDA_TRY(cudaStreamWaitEvent(st, h->ws_event, 0));
        const long long want = B < h->ws_warps ? B : h->ws_warps;
        h->kernel->launchg((unsigned)want, kWarpThreads, 0, st, h->dev, io, (long long)B, h->dlay);
        CUDA_TRY(cudaGetLastError());
        CUDA_TRY(cudaEventRecord(h->ws_event, st));
        return G1MPC_OK;
    }
    if (h->kernel->launchl && load <= h->lat_upto) {
        h->kernel->launchl((unsigned)B, 2 * kWarpThreads, h->kernel->smem, (cudaStream_t)stream, h->dev, io,
                           (long long)B, h->dlay);
        CUDA_TRY(cudaGetLastError());
        return G1MPC_OK;
    }
    if (h->kernel->launchr && load > h->refill_above && !u_warm) {
        cudaStream_t st = (cudaStream_t)stream;
        unsigned long long *cnt = h->d_refill + (h->refill_k++ % g1mpc_handle::kRefillCounters);
        CUDA_TRY(cudaMemsetAsync(cnt, 0, sizeof(unsigned long long), st));
        const long long nw = (B + 1) / 2, grid = nw < h->refill_ctas ? nw : h->refill_ctas;
        h->kernel->launchr((unsigned)grid, 2 * h->kernel->smem, st, h->dev, io, (long long)B, cnt);
        CUDA_TRY(cudaGetLastError());
        return G1MPC_OK;
    }
    const int wpb = h->warps_per_block;
    const bool two = h->kernel->fn2 && (h->force_ppw == 2 || (h->force_ppw != 1 && load > h->two_above));
    if (h->kernel->launchl2 && (h->lat2 == 2 || (h->lat2 == 1 && two))) {
        h->kernel->launchl2((unsigned)((B + 1) / 2), 2 * kWarpThreads, 2 * h->kernel->smem, (cudaStream_t)stream,
                            h->dev, io, (long long)B, h->dlay);
        CUDA_TRY(cudaGetLastError());
        return G1MPC_OK;
    }
    if (two)
        h->kernel->launch2((unsigned)((B + 2 * wpb - 1) / (2 * wpb)), kWarpThreads * wpb, 2 * h->kernel->smem * wpb, (cudaStream_t)stream, h->dev, io, (long long)B, h->dlay);
    else
        h->kernel->launch((unsigned)((B + wpb - 1) / wpb), kWarpThreads * wpb, h->kernel->smem * wpb, (cudaStream_t)stream, h->dev, io, (long long)B, h->dlay);
    CUDA_TRY(cudaGetLastError());
    return G1MPC_OK;
}

int g1mpc_solve_batch(g1mpc_handle *h, int64_t B, const double *x0, const double *x_ref, const double *contact,
                      const double *c_horizon, const double *p_com, const double *x_current, int one_rollout,
                      double *u_opt, double *x_opt, int32_t *status, int32_t *iters, double *resid, double *y_opt,
                      void *stream)
{
    return solve_batch_impl(h, B, B, x0, x_ref, contact, c_horizon, p_com, x_current, one_rollout, u_opt, x_opt, status, iters,
                            resid, y_opt, stream);
}

int g1mpc_solve_batch_warm(g1mpc_handle *h, int64_t B, const double *x0, const double *x_ref, const double *contact,
                           const double *c_horizon, const double *p_com, const double *x_current, int one_rollout,
                           const double *u_warm, int warm_shift, double *u_opt, double *x_opt, int32_t *status, int32_t *iters,
                           double *resid, double *y_opt, void *stream)
{
    return solve_batch_impl(h, B, B, x0, x_ref, contact, c_horizon, p_com, x_current, one_rollout, u_opt, x_opt, status, iters,
                            resid, y_opt, stream, u_warm, warm_shift);
}

// g1mpc_solve_batch for callers that cannot allocate (CUDA graph capture, tight serving loops): the global-workspace
// variant of the solve kernel on the file-scope scratch array g_static_ws.  No cudaMalloc / cudaFree / allocating library call
// here: the array's device address, the kernel and its occupancy were looked up in g1mpc_create.  Nothing in the scratch has
// to start at zero -- a problem writes every word of its slot before it reads it (the cudaMalloc'ed workspace of the existing
// path is never cleared either, and one warp reuses its slot for problem after problem) -- so the previous call's contents
// are simply overwritten.  All calls on one GPU share the array: calls that overlap on the GPU race.
int g1mpc_solve_batch_static_ws(g1mpc_handle *h, int64_t B, const double *x0, const double *x_ref, const double *contact,
                                const double *c_horizon, const double *p_com, const double *x_current, int one_rollout,
                                double *u_opt, double *x_opt, int32_t *status, int32_t *iters, double *resid, double *y_opt,
                                void *stream)
{
    if (!h || B < 0 || (B > 0 && (!x0 || !x_ref || !contact || !c_horizon || !p_com || !u_opt || !x_opt))) {
        g_last_error = "g1mpc_solve_batch_static_ws: null argument"; return G1MPC_ERR_INVALID_ARGUMENT;
    }
    if (B == 0) return G1MPC_OK;
    if (B > 0x7fffffffLL) { g_last_error = "batch too large for one launch"; return G1MPC_ERR_INVALID_ARGUMENT; }
    if (!h->sws_ptr) {
        g_last_error = std::string("g1mpc_solve_batch_static_ws: cudaGetSymbolAddress of the static scratch array failed at set-up: ") +
                       cudaGetErrorString(h->sws_err);
        return G1MPC_ERR_CUDA;
    }
    // one working-set slot per SM at least must fit the array (below that the persistent grid would leave SMs idle: use
    // g1mpc_solve_batch, which allocates what the shape needs)
    const size_t slot = h->dlay.bytes, fit = (size_t)G1MPC_STATIC_WS_BYTES / slot;
    if (fit < (size_t)h->sws_sms) {
        g_last_error = "g1mpc_solve_batch_static_ws: (horizon, num_contacts) = (" + std::to_string(h->params.horizon) + ", " +
                       std::to_string(h->params.num_contacts) + ") needs " + std::to_string(slot) + " B of scratch per problem in flight; the static " +
                       std::to_string((size_t)G1MPC_STATIC_WS_BYTES) + " B array must hold one per SM (" + std::to_string(h->sws_sms) +
                       "), i.e. at most " + std::to_string((size_t)G1MPC_STATIC_WS_BYTES / (size_t)h->sws_sms) +
                       " B each: use g1mpc_solve_batch for this shape";
        return G1MPC_ERR_UNSUPPORTED_SHAPE;
    }
    CUDA_TRY(cudaSetDevice(h->device));
    long long warps = (long long)h->sws_sms * h->sws_ctas_per_sm;      // persistent grid: warp w takes problems w, w + W, ...
    if (warps > (long long)fit) warps = (long long)fit;
    if (warps > B) warps = B;
    BatchPtrs io{x0, x_ref, contact, c_horizon, p_com, x_current, u_opt, x_opt, resid, y_opt, status, iters, one_rollout, nullptr, 0, h->sws_ptr};
    h->sws_launch((unsigned)warps, kWarpThreads, 0, (cudaStream_t)stream, h->dev, io, (long long)B, h->dlay);
    CUDA_TRY(cudaGetLastError());
    return G1MPC_OK;
}

int g1mpc_rollout_batch(g1mpc_handle *h, int64_t B, const double *x_ref, const double *c_horizon, const double *p_com,
                        const double *u_opt, const double *x_current, int one_rollout, double *x_opt, void *stream)
{
    if (!h || B < 0 || (B > 0 && (!x_ref || !c_horizon || !p_com || !u_opt || !x_opt))) {
        g_last_error = "g1mpc_rollout_batch: null argument"; return G1MPC_ERR_INVALID_ARGUMENT;
    }
    if (B == 0) return G1MPC_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    srbd_rollout_kernel<<<(unsigned)((B + 127) / 128), 128, 0, (cudaStream_t)stream>>>(h->dev, x_ref, c_horizon, p_com, u_opt, x_current,
                                                                                       one_rollout, x_opt, (long long)B, h->params.horizon,
                                                                                       h->params.num_contacts);
    CUDA_TRY(cudaGetLastError());
    return G1MPC_OK;
}

int g1mpc_rollout_host(g1mpc_handle *h, int64_t B, const double *x_ref, const double *c_horizon, const double *p_com,
                       const double *u_opt, const double *x_current, int one_rollout, double *x_opt)
{
    if (!h || B < 0 || (B > 0 && (!x_ref || !c_horizon || !p_com || !u_opt || !x_opt))) {
        g_last_error = "g1mpc_rollout_host: null argument"; return G1MPC_ERR_INVALID_ARGUMENT;
    }
    if (B == 0) return G1MPC_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    const size_t N = h->params.horizon, NU = 3 * (size_t)h->params.num_contacts;
    const size_t n_xr = B * N * NS, n_ch = B * N * NU, n_pc = B * N * 3, n_xc = (size_t)B * NS;
    const size_t need = sizeof(double) * (2 * n_xr + 2 * n_ch + n_pc + n_xc);
    if (need > h->loop_bytes) {                       // (shares the closed-loop workspace: neither call is re-entrant on one handle)
        if (h->d_loop) cudaFree(h->d_loop);
        h->d_loop = nullptr; h->loop_bytes = 0;
        CUDA_TRY(cudaMalloc((void **)&h->d_loop, need));
        h->loop_bytes = need;
    }
    double *d_xr = (double *)h->d_loop, *d_ch = d_xr + n_xr, *d_pc = d_ch + n_ch, *d_u = d_pc + n_pc, *d_xc = d_u + n_ch, *d_xo = d_xc + n_xc;
    cudaStream_t st = h->slot[0].pipe[0];
    CUDA_TRY(cudaMemcpyAsync(d_xr, x_ref, sizeof(double) * n_xr, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(d_ch, c_horizon, sizeof(double) * n_ch, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(d_pc, p_com, sizeof(double) * n_pc, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(d_u, u_opt, sizeof(double) * n_ch, cudaMemcpyHostToDevice, st));
    if (x_current) CUDA_TRY(cudaMemcpyAsync(d_xc, x_current, sizeof(double) * n_xc, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(d_xo, x_opt, sizeof(double) * n_xr, cudaMemcpyHostToDevice, st));
    int rc = g1mpc_rollout_batch(h, B, d_xr, d_ch, d_pc, d_u, x_current ? d_xc : nullptr, one_rollout, d_xo, st);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(x_opt, d_xo, sizeof(double) * n_xr, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return G1MPC_OK;
}

int g1mpc_set_host_warm_start(g1mpc_handle *h, int enable, int shift)
{
    if (!h || shift < 0) { g_last_error = "g1mpc_set_host_warm_start: bad argument"; return G1MPC_ERR_INVALID_ARGUMENT; }
    h->host_warm = enable ? 1 : 0;
    h->host_warm_shift = shift;
    for (auto &S : h->slot) S.prev_B = -1;
    return G1MPC_OK;
}

int g1mpc_layout_info(const g1mpc_handle *h, int64_t B, int32_t *problems_per_warp, int32_t *smem_bytes, int32_t *ctas_per_sm)
{
    if (!h) return G1MPC_ERR_INVALID_ARGUMENT;
    if (h->use_hybrid && ((B > h->hy_from && B <= h->hy_upto) || B >= h->hy_also_from)) {        // hybrid workspace: persistent grid, one problem per warp, the per-stage region in shared memory
        if (problems_per_warp) *problems_per_warp = h->hy_ppw == 2 ? -3 : -2;
        if (smem_bytes) *smem_bytes = (int32_t)h->hy_smem;
        if (ctas_per_sm) *ctas_per_sm = h->hy_ctas_per_sm;
        return G1MPC_OK;
    }
    if (h->use_global && B > h->global_above) {        // workspace in global memory: persistent grid, one problem per warp, no dynamic shared memory
        if (problems_per_warp) *problems_per_warp = 1;
        if (smem_bytes) *smem_bytes = 0;
        if (ctas_per_sm) *ctas_per_sm = h->ws_ctas_per_sm;
        return G1MPC_OK;
    }
    if (h->kernel->launchl && B <= h->lat_upto) {       // latency mode: one problem per CTA of two warps
        if (problems_per_warp) *problems_per_warp = 1;
        if (problems_per_warp) *problems_per_warp = -1;    // (marks the partner-warp mode: one problem per TWO warps)
        if (smem_bytes) *smem_bytes = (int32_t)h->kernel->smem;
        if (ctas_per_sm) {
            int n = 0;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, h->kernel->fnl, 2 * kWarpThreads, h->kernel->smem);
            *ctas_per_sm = n;
        }
        return G1MPC_OK;
    }
    const bool two = h->kernel->fn2 && (h->force_ppw == 2 || (h->force_ppw != 1 && B > h->two_above));
    const int ppw = two ? 2 : 1;
    if (problems_per_warp) *problems_per_warp = ppw;
    if (smem_bytes) *smem_bytes = (int32_t)(ppw * h->kernel->smem * h->warps_per_block);
    if (ctas_per_sm) {
        int n = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, two ? h->kernel->fn2 : h->kernel->fn, kWarpThreads * h->warps_per_block,
                                                      ppw * h->kernel->smem * h->warps_per_block);
        *ctas_per_sm = n;
    }
    return G1MPC_OK;
}

int g1mpc_emit_qp_batch(g1mpc_handle *h, int64_t B, const double *x0, const double *x_ref, const double *contact,
                        const double *c_horizon, const double *p_com, double *A, double *Bd, double *Aqp, double *Bqp,
                        double *H, double *g, double *Acon, double *l, double *u, void *stream)
{
    if (!h || B < 0 || (B > 0 && (!x0 || !x_ref || !contact || !c_horizon || !p_com))) {
        g_last_error = "g1mpc_emit_qp_batch: null argument"; return G1MPC_ERR_INVALID_ARGUMENT;
    }
    if ((l == nullptr) != (u == nullptr)) { g_last_error = "l and u must be given together"; return G1MPC_ERR_INVALID_ARGUMENT; }
    if (B == 0) return G1MPC_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    const int N = h->params.horizon, NC = h->params.num_contacts;
    const size_t smem = sizeof(double) * ((size_t)N * NS + 16 + 2 * (size_t)NC * N * 9 + (size_t)N * 12 + 6 * (size_t)N);
    if (smem > 96 * 1024) { g_last_error = "emit: shape too large"; return G1MPC_ERR_UNSUPPORTED_SHAPE; }
    EmitPtrs io{x0, x_ref, contact, c_horizon, p_com, A, Bd, Aqp, Bqp, H, g, Acon, l, u};
    srbd_mpc_emit_kernel<<<(unsigned)B, 256, smem, (cudaStream_t)stream>>>(h->dev, io, (long long)B, N, NC);
    CUDA_TRY(cudaGetLastError());
    return G1MPC_OK;
}

// true if cudaMemcpyAsync from/to p is a real asynchronous DMA (page-locked host memory: cudaMallocHost / cudaHostRegister)
static bool host_pinned(const void *p)
{
    if (!p) return true;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

int g1mpc_update_host_begin(g1mpc_handle *h, int slot, int64_t B, const double *x0, const double *x_ref, const double *contact,
                            const double *c_horizon, const double *p_com, const double *x_current, int one_rollout,
                            double *u_opt, double *x_opt, int32_t *status, int32_t *iters, double *resid)
{
    if (!h || slot < 0 || slot >= g1mpc_handle::kSlots || B < 0 ||
        (B > 0 && (!x0 || !x_ref || !contact || !c_horizon || !p_com || !u_opt || !x_opt))) {
        g_last_error = "g1mpc_update_host: null argument or slot out of range"; return G1MPC_ERR_INVALID_ARGUMENT;
    }
    g1mpc_handle::Slot &S = h->slot[slot];
    if (S.pending) { g_last_error = "g1mpc_update_host_begin: slot still in flight (call g1mpc_update_host_end first)"; return G1MPC_ERR_INVALID_ARGUMENT; }
    S.nchunk = 0;
    if (B == 0) { S.pending = true; return G1MPC_OK; }
    CUDA_TRY(cudaSetDevice(h->device));
    const size_t N = h->params.horizon, NC = h->params.num_contacts, NU = 3 * NC;
    // The batch is cut into up to kPipe chunks, each with its own stream, so that the copies of one chunk overlap the
    // solve of another.  Device block of a chunk:
    //   x0 | x_ref | contact | c_horizon | p_com | x_current | x_opt (in/out) | u_opt | resid | status | iters
    // Caller buffers that are page-locked are copied straight from / to where they lie (one DMA per array and chunk);
    // pageable ones go through the slot's pinned staging block (pack -> one H2D, one D2H -> unpack in _end).
    // x_opt only travels host->device when the reference would leave some of its rows untouched (no x_current: row 0;
    // one_rollout: rows 2..; mpc.cpp:427-446).
    const bool xopt_in = (x_current == nullptr) || one_rollout;
    const bool direct_in = host_pinned(x0) && host_pinned(x_ref) && host_pinned(contact) && host_pinned(c_horizon) &&
                           host_pinned(p_com) && host_pinned(x_current) && (!xopt_in || host_pinned(x_opt));
    const bool direct_out = host_pinned(u_opt) && host_pinned(x_opt) && host_pinned(status) && host_pinned(iters) && host_pinned(resid);
    const int nchunk = B >= 512 ? g1mpc_handle::kPipe : (B >= 128 ? 2 : 1);
    const size_t cap = ((size_t)B + nchunk - 1) / nchunk;
    const size_t d_in = NS + N * NS + N * NC + N * NU + N * 3 + NS, d_xo = N * NS, d_out = N * NU + 4;   // doubles per problem
    const size_t stride = ((sizeof(double) * (d_in + d_xo + d_out) + 2 * sizeof(int32_t)) * cap + 255) & ~(size_t)255;
    const size_t bytes = stride * nchunk;
    const bool warm = h->host_warm && S.prev_B == B && bytes <= S.stage_bytes;
    S.prev_B = B;
    if (bytes > S.stage_bytes) {
        if (S.h_stage) cudaFreeHost(S.h_stage);
        if (S.d_stage) cudaFree(S.d_stage);
        S.h_stage = nullptr; S.d_stage = nullptr; S.stage_bytes = 0;
        CUDA_TRY(cudaMallocHost((void **)&S.h_stage, bytes));
        CUDA_TRY(cudaMalloc((void **)&S.d_stage, bytes));
        S.stage_bytes = bytes;
    }
    // A handful of problems (the reference's own use: ONE updateMPC per control tick, mpc_to_wbid_node.cpp:494): no copy engine at
    // all.  The kernel reads its 2.7 KB of inputs straight from the slot's page-locked block over the bus and writes its results
    // there (pinned memory is mapped into the device's address space under unified addressing) -- two cudaMemcpyAsync calls and
    // their DMA start-up latency (~10-15 us) less on the one path where latency is the metric.
    const bool zero_copy = B <= g1mpc_handle::kZeroCopyMax && !h->use_global;
    S.stride = stride; S.cap = cap; S.staged_out = !direct_out || zero_copy;
    S.u_opt = u_opt; S.x_opt = x_opt; S.resid = resid; S.status = status; S.iters = iters;
    const size_t D = sizeof(double);
    if (zero_copy) {
        const size_t nb = (size_t)B;
        S.nbs[0] = nb; S.nchunk = 1;
        double *hs = (double *)S.h_stage;
        g1mpc_handle::Off &f = S.off[0];
        size_t q = 0;
        f.x0 = q; q += nb * NS;  f.xr = q; q += nb * N * NS;  f.ct = q; q += nb * N * NC;  f.ch = q; q += nb * N * NU;
        f.pc = q; q += nb * N * 3;  f.xc = q; q += nb * NS;  f.xo = q; q += nb * N * NS;
        f.u = q; q += nb * N * NU;  f.res = q; q += nb * 4;  f.st = q;
        memcpy(hs + f.x0, x0, D * nb * NS);
        memcpy(hs + f.xr, x_ref, D * nb * N * NS);
        memcpy(hs + f.ct, contact, D * nb * N * NC);
        memcpy(hs + f.ch, c_horizon, D * nb * N * NU);
        memcpy(hs + f.pc, p_com, D * nb * N * 3);
        if (x_current) memcpy(hs + f.xc, x_current, D * nb * NS);
        if (xopt_in) memcpy(hs + f.xo, x_opt, D * nb * N * NS);
        int32_t *hs_i = (int32_t *)(hs + f.st);
        // (a warm start reads the previous call's u_opt where that call left it: the same place of the same block)
        int rc = solve_batch_impl(h, (int64_t)nb, B, hs + f.x0, hs + f.xr, hs + f.ct, hs + f.ch, hs + f.pc, x_current ? hs + f.xc : nullptr,
                                  one_rollout, hs + f.u, hs + f.xo, hs_i, hs_i + nb, hs + f.res, nullptr, S.pipe[0],
                                  (warm && S.prev_zero_copy) ? hs + f.u : nullptr, h->host_warm_shift);
        if (rc) return rc;
        S.prev_zero_copy = true;
        S.pending = true;
        return G1MPC_OK;
    }
    S.prev_zero_copy = false;
    for (int c = 0; c < nchunk; c++) {
        const size_t b0 = cap * c, nb = b0 < (size_t)B ? (((size_t)B - b0) < cap ? ((size_t)B - b0) : cap) : 0;
        S.nbs[c] = nb;
        S.nchunk = c + 1;
        if (nb == 0) continue;
        double *hs = (double *)(S.h_stage + stride * c), *ds = (double *)(S.d_stage + stride * c);
        g1mpc_handle::Off &f = S.off[c];
        size_t q = 0;
        f.x0 = q; q += nb * NS;  f.xr = q; q += nb * N * NS;  f.ct = q; q += nb * N * NC;  f.ch = q; q += nb * N * NU;
        f.pc = q; q += nb * N * 3;  f.xc = q; q += nb * NS;  f.xo = q; q += nb * N * NS;
        f.u = q; q += nb * N * NU;  f.res = q; q += nb * 4;  f.st = q;      // int32 status[nb] | iters[nb]
        cudaStream_t sp = S.pipe[c];
        if (direct_in) {
            CUDA_TRY(cudaMemcpyAsync(ds + f.x0, x0 + b0 * NS, D * nb * NS, cudaMemcpyHostToDevice, sp));
            CUDA_TRY(cudaMemcpyAsync(ds + f.xr, x_ref + b0 * N * NS, D * nb * N * NS, cudaMemcpyHostToDevice, sp));
            CUDA_TRY(cudaMemcpyAsync(ds + f.ct, contact + b0 * N * NC, D * nb * N * NC, cudaMemcpyHostToDevice, sp));
            CUDA_TRY(cudaMemcpyAsync(ds + f.ch, c_horizon + b0 * N * NU, D * nb * N * NU, cudaMemcpyHostToDevice, sp));
            CUDA_TRY(cudaMemcpyAsync(ds + f.pc, p_com + b0 * N * 3, D * nb * N * 3, cudaMemcpyHostToDevice, sp));
            if (x_current) CUDA_TRY(cudaMemcpyAsync(ds + f.xc, x_current + b0 * NS, D * nb * NS, cudaMemcpyHostToDevice, sp));
            if (xopt_in) CUDA_TRY(cudaMemcpyAsync(ds + f.xo, x_opt + b0 * N * NS, D * nb * N * NS, cudaMemcpyHostToDevice, sp));
        } else {
            memcpy(hs + f.x0, x0 + b0 * NS, D * nb * NS);
            memcpy(hs + f.xr, x_ref + b0 * N * NS, D * nb * N * NS);
            memcpy(hs + f.ct, contact + b0 * N * NC, D * nb * N * NC);
            memcpy(hs + f.ch, c_horizon + b0 * N * NU, D * nb * N * NU);
            memcpy(hs + f.pc, p_com + b0 * N * 3, D * nb * N * 3);
            if (x_current) memcpy(hs + f.xc, x_current + b0 * NS, D * nb * NS);
            if (xopt_in) memcpy(hs + f.xo, x_opt + b0 * N * NS, D * nb * N * NS);
            CUDA_TRY(cudaMemcpyAsync(ds, hs, D * (xopt_in ? f.u : f.xo), cudaMemcpyHostToDevice, sp));
        }
        int32_t *ds_i = (int32_t *)(ds + f.st);
        int rc = solve_batch_impl(h, (int64_t)nb, B, ds + f.x0, ds + f.xr, ds + f.ct, ds + f.ch, ds + f.pc,
                                  x_current ? ds + f.xc : nullptr, one_rollout, ds + f.u, ds + f.xo, ds_i, ds_i + nb,
                                  ds + f.res, nullptr, sp, warm ? ds + f.u : nullptr, h->host_warm_shift);
        if (rc) return rc;
        if (direct_out) {
            CUDA_TRY(cudaMemcpyAsync(x_opt + b0 * N * NS, ds + f.xo, D * nb * N * NS, cudaMemcpyDeviceToHost, sp));
            CUDA_TRY(cudaMemcpyAsync(u_opt + b0 * N * NU, ds + f.u, D * nb * N * NU, cudaMemcpyDeviceToHost, sp));
            if (resid) CUDA_TRY(cudaMemcpyAsync(resid + b0 * 4, ds + f.res, D * nb * 4, cudaMemcpyDeviceToHost, sp));
            if (status) CUDA_TRY(cudaMemcpyAsync(status + b0, ds_i, sizeof(int32_t) * nb, cudaMemcpyDeviceToHost, sp));
            if (iters) CUDA_TRY(cudaMemcpyAsync(iters + b0, ds_i + nb, sizeof(int32_t) * nb, cudaMemcpyDeviceToHost, sp));
        } else {
            CUDA_TRY(cudaMemcpyAsync(hs + f.xo, ds + f.xo, D * (f.st - f.xo) + sizeof(int32_t) * 2 * nb, cudaMemcpyDeviceToHost, sp));
        }
    }
    S.pending = true;
    return G1MPC_OK;
}

int g1mpc_update_host_end(g1mpc_handle *h, int slot)
{
    if (!h || slot < 0 || slot >= g1mpc_handle::kSlots) { g_last_error = "g1mpc_update_host_end: bad handle or slot"; return G1MPC_ERR_INVALID_ARGUMENT; }
    g1mpc_handle::Slot &S = h->slot[slot];
    if (!S.pending) { g_last_error = "g1mpc_update_host_end: nothing in flight on this slot"; return G1MPC_ERR_INVALID_ARGUMENT; }
    S.pending = false;
    CUDA_TRY(cudaSetDevice(h->device));
    const size_t N = h->params.horizon, NC = h->params.num_contacts, NU = 3 * NC;
    for (int c = 0; c < S.nchunk; c++) {
        const size_t b0 = S.cap * c, nb = S.nbs[c];
        if (nb == 0) continue;
        CUDA_TRY(cudaStreamSynchronize(S.pipe[c]));
        if (!S.staged_out) continue;
        const double *hs = (const double *)(S.h_stage + S.stride * c);
        const g1mpc_handle::Off &f = S.off[c];
        memcpy(S.u_opt + b0 * N * NU, hs + f.u, sizeof(double) * nb * N * NU);
        memcpy(S.x_opt + b0 * N * NS, hs + f.xo, sizeof(double) * nb * N * NS);
        if (S.resid) memcpy(S.resid + b0 * 4, hs + f.res, sizeof(double) * nb * 4);
        const int32_t *hs_i = (const int32_t *)(hs + f.st);
        if (S.status) memcpy(S.status + b0, hs_i, sizeof(int32_t) * nb);
        if (S.iters) memcpy(S.iters + b0, hs_i + nb, sizeof(int32_t) * nb);
    }
    return G1MPC_OK;
}

int g1mpc_update_host(g1mpc_handle *h, int64_t B, const double *x0, const double *x_ref, const double *contact,
                      const double *c_horizon, const double *p_com, const double *x_current, int one_rollout,
                      double *u_opt, double *x_opt, int32_t *status, int32_t *iters, double *resid)
{
    if (h && h->slot[0].pending) { g_last_error = "g1mpc_update_host: slot 0 is in flight (g1mpc_update_host_begin without _end)"; return G1MPC_ERR_INVALID_ARGUMENT; }
    int rc = g1mpc_update_host_begin(h, 0, B, x0, x_ref, contact, c_horizon, p_com, x_current, one_rollout, u_opt, x_opt, status, iters, resid);
    if (rc) { if (h) h->slot[0].pending = false; return rc; }
    return g1mpc_update_host_end(h, 0);
}

int g1mpc_walk_inputs_batch(g1mpc_handle *h, int64_t B, const double *x0, const double *contacts_now, const double *sim_time,
                            const double *x_opt_prev, double *x_ref, double *contact, double *c_horizon, double *p_com,
                            double *landing, void *stream)
{
    if (!h || B < 0 || (B > 0 && (!x0 || !contacts_now || !sim_time || !x_opt_prev || !x_ref || !contact || !c_horizon || !p_com))) {
        g_last_error = "g1mpc_walk_inputs_batch: null argument"; return G1MPC_ERR_INVALID_ARGUMENT;
    }
    if (h->params.num_contacts != 4 || h->params.horizon > 10) {
        g_last_error = "g1mpc_walk_inputs_batch: the node's gait table needs num_contacts == 4 and horizon <= 10";
        return G1MPC_ERR_UNSUPPORTED_SHAPE;
    }
    if (B == 0) return G1MPC_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    adaptors::WalkPtrs io{x0, contacts_now, sim_time, x_opt_prev, x_ref, contact, c_horizon, p_com, landing};
    adaptors::walk_inputs_kernel<<<(unsigned)((B + 3) / 4), 128, 0, (cudaStream_t)stream>>>(io, (long long)B, h->params.horizon,
                                                                                            h->params.dt, h->params.gravity);
    CUDA_TRY(cudaGetLastError());
    return G1MPC_OK;
}

int g1mpc_wbid_refs_batch(g1mpc_handle *h, int64_t B, const double *u_opt, const double *x_opt, const double *foot_positions,
                          double wbid_mass, double *refs, void *stream)
{
    if (!h || B < 0 || (B > 0 && (!u_opt || !x_opt || !foot_positions || !refs)) || !(wbid_mass > 0)) {
        g_last_error = "g1mpc_wbid_refs_batch: bad argument"; return G1MPC_ERR_INVALID_ARGUMENT;
    }
    if (h->params.num_contacts != 4 || h->params.horizon < 2) {
        g_last_error = "g1mpc_wbid_refs_batch: the WBID layer has four contact frames (wbid.py:39-44)";
        return G1MPC_ERR_UNSUPPORTED_SHAPE;
    }
    if (B == 0) return G1MPC_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    adaptors::wbid_refs_kernel<<<(unsigned)((B + 127) / 128), 128, 0, (cudaStream_t)stream>>>(u_opt, x_opt, foot_positions, refs, (long long)B,
                                                                                              h->params.horizon, wbid_mass);
    CUDA_TRY(cudaGetLastError());
    return G1MPC_OK;
}

int g1mpc_closed_loop_batch_ex(g1mpc_handle *h, int64_t B, int32_t ticks, int warm_start, double *x0, double *contacts_now,
                               double *sim_time, double *x_ref, double *x_opt, double *u_opt, double *traj_x, double *traj_u,
                               int32_t *status_any, int32_t *iters_sum, void *stream)
{
    if (!h || B < 0 || ticks < 0 || (B > 0 && (!x0 || !contacts_now || !sim_time || !x_ref || !x_opt || !u_opt))) {
        g_last_error = "g1mpc_closed_loop_batch: bad argument"; return G1MPC_ERR_INVALID_ARGUMENT;
    }
    if (h->params.num_contacts != 4 || h->params.horizon > 10) {
        g_last_error = "g1mpc_closed_loop_batch: the node's gait table needs num_contacts == 4 and horizon <= 10";
        return G1MPC_ERR_UNSUPPORTED_SHAPE;
    }
    if (B == 0 || ticks == 0) return G1MPC_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    const size_t N = h->params.horizon;
    // workspace: contact [B,N,4] | c_horizon [B,N,12] | p_com [B,N,3] | x0 copy, carry [B,13] each | status [B] int32 | iters [B] int32
    const size_t need = sizeof(double) * ((size_t)B * N * 19 + 2 * (size_t)B * NS) + sizeof(int32_t) * 2 * (size_t)B;
    if (need > h->loop_bytes) {
        CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
        if (h->d_loop) cudaFree(h->d_loop);
        h->d_loop = nullptr; h->loop_bytes = 0;
        CUDA_TRY(cudaMalloc((void **)&h->d_loop, need));
        h->loop_bytes = need;
    }
    double *contact = (double *)h->d_loop, *c_hor = contact + (size_t)B * N * 4, *p_com = c_hor + (size_t)B * N * 12;
    int32_t *status = (int32_t *)(p_com + (size_t)B * N * 3 + 2 * (size_t)B * NS), *iters = status + B;
    for (int t = 0; t < ticks; t++) {
        int rc = g1mpc_walk_inputs_batch(h, B, x0, contacts_now, sim_time, x_opt, x_ref, contact, c_hor, p_com, nullptr, stream);
        if (rc) return rc;
        rc = solve_batch_impl(h, B, B, x0, x_ref, contact, c_hor, p_com, x0, 0, u_opt, x_opt, status, iters, nullptr, nullptr, stream,
                              (warm_start && t > 0) ? u_opt : nullptr, 1);
        if (rc) return rc;
        adaptors::advance_kernel<<<(unsigned)((B + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
            x0, contacts_now, sim_time, x_opt, u_opt, c_hor, status, iters, traj_x, traj_u, status_any, iters_sum, (long long)B, (int)N, t, ticks,
            h->params.dt);
        CUDA_TRY(cudaGetLastError());
    }
    return G1MPC_OK;
}

int g1mpc_closed_loop_batch(g1mpc_handle *h, int64_t B, int32_t ticks, double *x0, double *contacts_now, double *sim_time,
                            double *x_ref, double *x_opt, double *u_opt, double *traj_x, double *traj_u, int32_t *status_any,
                            void *stream)
{
    return g1mpc_closed_loop_batch_ex(h, B, ticks, 0, x0, contacts_now, sim_time, x_ref, x_opt, u_opt, traj_x, traj_u, status_any, nullptr, stream);
}

int g1mpc_plan_gait_batch(g1mpc_handle *h, int64_t B, const double *planner, const double *current_time, double foot_offset,
                          double *contact, double *c_horizon, void *stream)
{
    if (!h || B < 0 || (B > 0 && (!planner || !current_time || !contact || !c_horizon))) {
        g_last_error = "g1mpc_plan_gait_batch: null argument"; return G1MPC_ERR_INVALID_ARGUMENT;
    }
    if (h->params.num_contacts != 4) {
        g_last_error = "g1mpc_plan_gait_batch: the planner's mapping has heel and toe of two feet (main.py:273-283): num_contacts == 4";
        return G1MPC_ERR_UNSUPPORTED_SHAPE;
    }
    if (B == 0) return G1MPC_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    const long long n = (long long)B * h->params.horizon;
    adaptors::plan_gait_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(planner, current_time, foot_offset, h->params.dt,
                                                                                               contact, c_horizon, (long long)B, h->params.horizon);
    CUDA_TRY(cudaGetLastError());
    return G1MPC_OK;
}

int g1mpc_main_loop_batch(g1mpc_handle *h, int64_t B, int32_t ticks, const double *planner, double total_duration, double foot_offset,
                          int warm_start, double *current_time, double *x_ref, double *x_opt, double *u_opt, double *traj_x,
                          double *traj_u, int32_t *status_any, int32_t *iters_sum, void *stream)
{
    if (!h || B < 0 || ticks < 0 || (B > 0 && (!planner || !current_time || !x_ref || !x_opt || !u_opt))) {
        g_last_error = "g1mpc_main_loop_batch: bad argument"; return G1MPC_ERR_INVALID_ARGUMENT;
    }
    if (h->params.num_contacts != 4 || h->params.horizon < 3) {
        g_last_error = "g1mpc_main_loop_batch: needs num_contacts == 4 and horizon >= 3 (x_opt[2] is fed back, main.py:228)";
        return G1MPC_ERR_UNSUPPORTED_SHAPE;
    }
    if (B == 0 || ticks == 0) return G1MPC_OK;
    CUDA_TRY(cudaSetDevice(h->device));
    const size_t N = h->params.horizon;
    const size_t need = sizeof(double) * ((size_t)B * N * 19 + 2 * (size_t)B * NS) + sizeof(int32_t) * 2 * (size_t)B;
    if (need > h->loop_bytes) {
        CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
        if (h->d_loop) cudaFree(h->d_loop);
        h->d_loop = nullptr; h->loop_bytes = 0;
        CUDA_TRY(cudaMalloc((void **)&h->d_loop, need));
        h->loop_bytes = need;
    }
    double *contact = (double *)h->d_loop, *c_hor = contact + (size_t)B * N * 4, *p_com = c_hor + (size_t)B * N * 12;
    double *x0 = p_com + (size_t)B * N * 3, *carry = x0 + (size_t)B * NS;
    int32_t *status = (int32_t *)(carry + (size_t)B * NS), *iters = status + B;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned gb = (unsigned)((B + 127) / 128), gn = (unsigned)((B * (long long)N + 127) / 128);
    for (int t = 0; t < ticks; t++) {
        adaptors::main_pre_kernel<<<gb, 128, 0, st>>>(x0, carry, x_ref, x_opt, (long long)B, (int)N, t);
        if (t > 0) adaptors::double_support_ref_kernel<<<gn, 128, 0, st>>>(current_time, total_duration, h->params.dt, x_ref, (long long)B, (int)N);
        int rc = g1mpc_plan_gait_batch(h, B, planner, current_time, foot_offset, contact, c_hor, stream);
        if (rc) return rc;
        // p_com_horizon = x_ref_hor[:, 3:6]  (main.py:286): strided copy
        CUDA_TRY(cudaMemcpy2DAsync(p_com, 3 * sizeof(double), x_ref + 3, NS * sizeof(double), 3 * sizeof(double), (size_t)B * N,
                                   cudaMemcpyDeviceToDevice, st));
        rc = solve_batch_impl(h, B, B, x0, x_ref, contact, c_hor, p_com, carry, 0, u_opt, x_opt, status, iters, nullptr, nullptr, stream,
                              (warm_start && t > 0) ? u_opt : nullptr, 1);
        if (rc) return rc;
        adaptors::main_post_kernel<<<gb, 128, 0, st>>>(current_time, x_opt, u_opt, status, iters, traj_x, traj_u, status_any, iters_sum,
                                                        (long long)B, (int)N, t, ticks, h->params.dt);
        CUDA_TRY(cudaGetLastError());
    }
    return G1MPC_OK;
}

int g1mpc_fp64_peak_tflops(int device, double *tflops)
{
    if (!tflops) return G1MPC_ERR_INVALID_ARGUMENT;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { g_last_error = "no CUDA device"; return G1MPC_ERR_NO_DEVICE; }
    CUDA_TRY(cudaSetDevice(device));
    int sms = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    const int blocks = sms * 8, threads = 256, iters = 1 << 15;
    double *buf = nullptr;
    CUDA_TRY(cudaMalloc((void **)&buf, sizeof(double) * blocks * threads));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    fp64_peak_kernel<<<blocks, threads>>>(buf, 1 << 10);
    double best = 0.0;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        fp64_peak_kernel<<<blocks, threads>>>(buf, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        const double tf = 2.0 * 8.0 * (double)iters * blocks * threads / (ms * 1e-3) / 1e12;
        if (tf > best) best = tf;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(buf);
    CUDA_TRY(cudaGetLastError());
    *tflops = best;
    return G1MPC_OK;
}

} // extern "C"
