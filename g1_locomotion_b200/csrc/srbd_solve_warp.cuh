// srbd_solve_warp.cuh -- v2 solve kernel: ONE WARP PER PROBLEM, stage-wise (Riccati) interior point, FP64.
// Included by g1mpc_kernels.cu inside its anonymous namespace (uses DevParams, BatchPtrs, Lin, linearise, NS).
//
// Reference path replaced: g1_mpc::MPC::updateMPC (g1_mpc/src/mpc.cpp:448-473) = linearise/discretise
// (mpc.cpp:153-254), condensed build (mpc.cpp:256-291,359-376), constraints (mpc.cpp:293-357), QP solve
// (mpc.cpp:378-425, OSQP in the reference) and rollout (mpc.cpp:427-446).
//
// Mathematics (checked on the CPU against the oracle by tests/ipm_model.py + tests/test_ipm_model_cpu.py):
//   * states y = (theta, p | omega, v) (12 weighted states; gravity is constant), per-step wrench increment
//     w_k = Bm (sum_c r_c x f_c, sum_c f_c),  Bm = blockdiag(dt Iw^-1, dt/m I)         (mpc.cpp:213-254)
//     y_{k+1} = [I Phi; 0 I] y_k + [0; I] w_k,  Phi = dt blockdiag(Rz(psi), I)           (mpc.cpp:196-211)
//     so the reference Hessian/gradient (mpc.cpp:361,375) are H = s r I + W'MW, g = W'h with M = G' sQ G.
//   * QP: Mehrotra predictor-corrector on the 6 one-sided rows per stance contact (mpc.cpp:293-357); swing
//     contacts (flag 0 => f = 0) drop out.  Newton systems (Dblk + W'MW) du = rhs with Dblk block-diagonal 3x3
//     reduce to (M^-1 + S) y = t in wrench space, S_k = W_k Dblk_k^-1 W_k' (6x6 per stage), solved by a
//     backward Riccati sweep over the horizon: per stage two 6x6 LDL' factorisations, explicit closed-loop
//     blocks Gamma' and Psi, then one small mat-vec per stage in each of the backward/forward vector sweeps.
//     Cost O(N) per iteration, shared memory O(N): the condensed 3ncN x 3ncN Hessian is never formed.
//   * one warp owns one problem: no block-level barrier anywhere, only __syncwarp(); the sequential 6x6
//     factorisations run redundantly in every lane's registers (no shuffles on the critical path), the wide
//     phases (per contact, per stage, per matrix column) are spread over the lanes.
#pragma once

namespace v2 {

// optional per-phase cycle counters (development builds: -DG1MPC_PHASE_CLOCKS; written over the first 8 entries of y_opt)
#ifdef G1MPC_PHASE_CLOCKS
#define PH_DECL long long ph_t[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}; long long ph_last = clock64();
#define PH(i) { const long long ph_now = clock64(); ph_t[i] += ph_now - ph_last; ph_last = ph_now; }
#define PH_PARAMS , long long (&ph_t)[16], long long &ph_last
#define PH_ARGS , ph_t, ph_last
#ifdef G1MPC_FACTOR_CLOCKS      // counters 8..12 split the factor stage instead of the solve
#define PHF(i) PH(i)
#define PHS(i)
#else
#define PHF(i)
#define PHS(i) PH(i)
#endif
#else
#define PHF(i)
#define PHS(i)
#define PH_DECL
#define PH(i)
#define PH_PARAMS
#define PH_ARGS
#endif

constexpr unsigned kFullMask = 0xffffffffu;
constexpr int kWarmCheckIt = 5;         // iteration at which a warm-started problem must have halved its gap (else: cold restart)

// Reductions over the HL (32 or 16) lanes that own one problem; lanes of the other half-warp never mix in (xor offsets < HL).
template <int HL>
__device__ __forceinline__ double wsum(double v)
{
#pragma unroll
    for (int o = HL / 2; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
    return v;
}
template <int HL>
__device__ __forceinline__ double wmax(double v)
{
#pragma unroll
    for (int o = HL / 2; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(kFullMask, v, o));
    return v;
}
template <int HL>
__device__ __forceinline__ double wmin(double v)
{
#pragma unroll
    for (int o = HL / 2; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(kFullMask, v, o));
    return v;
}
// Sum over the contacts of a problem in an order that does not depend on HL: contact i belongs to "virtual lane" i % 32;
// with 16 lanes per problem the two virtual lanes of a lane are kept apart (v[0]: rounds 0, 2, ..; v[1]: rounds 1, 3, ..)
// and added as the 32-lane butterfly's first step would add them, so both layouts give the same bits.
template <int HL>
__device__ __forceinline__ double wsum2(const double (&v)[2])
{
    return wsum<HL>(HL == 32 ? v[0] : v[0] + v[1]);
}

// packed lower-triangular index of a symmetric 6x6, i >= j
__host__ __device__ constexpr int tri(int i, int j) { return i * (i + 1) / 2 + j; }
__device__ __forceinline__ int trs(int i, int j) { return i >= j ? i * (i + 1) / 2 + j : j * (j + 1) / 2 + i; }

// Shared-memory layout of one problem (offsets in doubles).  Regions that are never live at the same time
// share storage; the comments say who owns what when.
__host__ __device__ constexpr int ev(int x) { return (x + 1) & ~1; }

template <int N_, int NC_>
struct WLay {
    static constexpr int N = N_, NC = NC_;
    static constexpr int NCN = NC * N;
    // ---- per contact, persistent over the solve ----
    static constexpr int oR = 0;                    // [NCN*3] lever arm r = c - p_com            (mpc.cpp:229-233)
    static constexpr int oCf = ev(oR + NCN * 3);        // [NCN]   contact flag (multiplies fz bounds, mpc.cpp:344,350)
    static constexpr int oU = ev(oCf + NCN);            // [NCN*3] forces
    static constexpr int oSl = ev(oU + NCN * 3);        // [NCN*6] slacks
    static constexpr int oLam = oSl + NCN * 6;      // [NCN*6] multipliers
    static constexpr int oGf = oLam + NCN * 6;      // [NCN*3] grad f(u)
    static constexpr int oDst = ev(oGf + NCN * 3);      // [NCN*3] 1/a, 1/b, 1/d2 of the 3x3 block (1/s is recomputed where needed)
    static constexpr int oProd = ev(oDst + NCN * 3);    // [NCN*6] predictor ds*dl (Mehrotra second-order term), then q, then W_c^T y
    static constexpr bool kTvInScratch = 3 * NCN <= 12 * N;     // Dblk^-1 rhs / du fit into vB|vC, dead whenever they are live
    static constexpr int oTvOwn = oProd + NCN * 6;      // [NCN*3] Dblk^-1 rhs, then du (only when they do not fit into the scratch)
    // ---- per stage ----  (every region below starts on a 16-byte boundary: rows are read as double2)
    static constexpr int oSt = ev(oTvOwn + (kTvInScratch ? 0 : NCN * 3));   // [(NCN+1)/2] int32 list of stance contacts (compacted)
    static constexpr int oE = ev(oSt + (NCN + 1) / 2);   // [N*12]  free-response error A^(k+1) x0 - xref_k (mpc.cpp:369-375)
    static constexpr int SST = 22;                  // S_k packed (21) + pad
    static constexpr int oS = oE + N * 12;          // [N*SST]
    // per stage: G2 = Gamma - Phi^T Psi | Psi | P12 of P_{k+1} | Gamma | P22 of P_{k+1} (packed 21 + pad);
    // G2, Psi and Gamma are stored by columns (the factor lanes own columns), P12 by rows
    // (Gamma is kept beside G2: rebuilding it as G2 + Phi^T Psi cancels badly on long horizons -- measured at N = 40)
    static constexpr int sG2 = 0, sPsi = 36, sP12 = 72, sGam = 108, sP22 = 144, STG = 166;
    static constexpr int oRic = oS + N * SST;       // [N*STG] stage factors.  The region is dead from the end of the corrector
                                                    // solve to the next factorisation and then holds, one after the other:
                                                    //   ds | dl  (needs 12 NCN <= STG N)   and the gradient sweeps' vectors
    static constexpr int oW = oRic;                 //   gradient: w  [N*6]      (these four also survive the last iteration:
    static constexpr int oXs = oW + N * 6;          //             X  [N*12]      the outputs read w, X and V)
    static constexpr int oL1 = oXs + N * 12;        //             L1 [N*6]
    static constexpr int oV = oL1 + N * 6;          //             V  [N*6]  (then Bm' V in place)
    // ---- scratch of the solves and of the factor exchange ----
    static constexpr int oX = oRic + N * STG;
    static constexpr int oVa = oX;                  // solve: vA [N*6]  t -> Gamma't -> b -> y2
    static constexpr int oVb = oVa + N * 6;         //        vB [N*6]  c1 -> y1
    static constexpr int oVc = oVb + N * 6;         //        vC [N*6]  c2' -> y (multiplier wrench, output of the solve)
    static constexpr int oVd = oVc + N * 6;         //        vD [N*6]  p2 of P_{k+1} (incoming co-state of stage k)
    static constexpr int oPb = oVd + N * 6;         //        p1 exchange, double buffered [12]
    static constexpr int oEx = oVd;                 // factor exchange: A21 [36] (dead before the solve starts)
    static constexpr int oTv = kTvInScratch ? oVb : oTvOwn;     // written after a solve has delivered vA, read before the next one starts
    static constexpr int scratch = (N * 24 + 12 > 18 * N + 36) ? N * 24 + 12 : 18 * N + 36;
    static constexpr int oEnd = ev(oX + scratch);
    static_assert(12 * NCN <= STG * N, "ds|dl alias the stage storage: needs num_contacts <= 10");
    static_assert(30 * N <= STG * N, "gradient vectors alias the stage storage");
    static constexpr size_t bytes = sizeof(double) * (size_t)oEnd;
};

// The same layout with the horizon and the number of contacts known only at run time (the reference takes both at run
// time: HORIZON_LENGTH is a constructor argument, mpc.py:12 / mpc.hpp:22, NUM_CONTACTS an attribute, mpc.py:17): every
// member of WLay as a plain int, computed once per launch on the host and passed to the kernel by value.  The device code
// is written against `L.member`, so one source serves the compile-time shapes (members are constants, everything folds
// exactly as before) and this one (members live in registers / constant bank).
struct DLay {
    int N, NC, NCN, oR, oCf, oU, oSl, oLam, oGf, oDst, oProd, oTvOwn, oSt, oE, oS, oRic, oW, oXs, oL1, oV, oX, oVa, oVb, oVc,
        oVd, oPb, oEx, oTv, scratch, oEnd;
    bool kTvInScratch;
    static constexpr int SST = 22, sG2 = 0, sPsi = 36, sP12 = 72, sGam = 108, sP22 = 144, STG = 166;
    size_t bytes;
    __host__ __device__ DLay() {}
    __host__ DLay(int n, int nc)
    {
        N = n; NC = nc; NCN = nc * n;
        oR = 0; oCf = ev(oR + NCN * 3); oU = ev(oCf + NCN); oSl = ev(oU + NCN * 3); oLam = oSl + NCN * 6; oGf = oLam + NCN * 6;
        oDst = ev(oGf + NCN * 3); oProd = ev(oDst + NCN * 3);
        kTvInScratch = 3 * NCN <= 12 * N;
        oTvOwn = oProd + NCN * 6;
        oSt = ev(oTvOwn + (kTvInScratch ? 0 : NCN * 3));
        oE = ev(oSt + (NCN + 1) / 2);
        oS = oE + N * 12; oRic = oS + N * SST;
        oW = oRic; oXs = oW + N * 6; oL1 = oXs + N * 12; oV = oL1 + N * 6;
        oX = oRic + N * STG; oVa = oX; oVb = oVa + N * 6; oVc = oVb + N * 6; oVd = oVc + N * 6; oPb = oVd + N * 6; oEx = oVd;
        oTv = kTvInScratch ? oVb : oTvOwn;
        scratch = (N * 24 + 12 > 18 * N + 36) ? N * 24 + 12 : 18 * N + 36;
        oEnd = ev(oX + scratch);
        bytes = sizeof(double) * (size_t)oEnd;
    }
    // what the static_asserts of WLay demand
    __host__ bool valid() const { return N >= 2 && NC >= 1 && 12 * NCN <= STG * N && 12 * N >= 22; }
};

// 1/d: MUFU.RCP64H seed (relative error 9.6e-7 measured on B200, profiles/micro/lat.cu) + one cubically convergent
// step x (1 + e + e^2), e = 1 - d x: 2.2e-16 maximum relative error measured, 39 cycles dependent latency against 74 of
// `1.0 / d`.  No special cases: every caller passes a finite, normal, non-zero value or the problem is discarded
// afterwards by the NaN checks.
__device__ __forceinline__ double frcp(double d)
{
    double x;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(d));
    const double e = fma(-d, x, 1.0);
    return fma(x, fma(e, e, e), x);
}

// LDL' of a symmetric positive definite 6x6 held packed (21) in registers; afterwards a[tri(i,j)], i>j, holds
// the unit-lower factor and dinv[j] = 1/d_j.  `neg` collects the sign bits of the pivots (negative => not positive definite).
__device__ __forceinline__ void ldl6(double (&a)[21], double (&dinv)[6], int &neg)
{
#pragma unroll
    for (int j = 0; j < 6; j++) {
        const double dj = a[tri(j, j)];
        neg |= __double2hiint(dj);                   // sign bit of any pivot; a zero pivot shows up as Inf/NaN later
        const double r = frcp(dj);
        dinv[j] = r;
        double col[6];
#pragma unroll
        for (int i = j + 1; i < 6; i++) col[i] = a[tri(i, j)];
        // the next pivot is the dependent chain: its update waits for r with one operation (the square is formed
        // while the reciprocal runs), everything else has a whole reciprocal of slack
        if (j < 5) a[tri(j + 1, j + 1)] = fma(-(col[j + 1] * col[j + 1]), r, a[tri(j + 1, j + 1)]);
#pragma unroll
        for (int i = j + 1; i < 6; i++) {
            const double lij = col[i] * r;
#pragma unroll
            for (int c = j + 1; c <= i; c++) if (!(i == j + 1 && c == j + 1)) a[tri(i, c)] = fma(-lij, col[c], a[tri(i, c)]);
            a[tri(i, j)] = lij;
        }
    }
}
// x <- L^-1 x  (unit lower, packed)
__device__ __forceinline__ void fsub6(const double (&l)[21], double (&x)[6])
{
#pragma unroll
    for (int i = 1; i < 6; i++)
#pragma unroll
        for (int j = 0; j < i; j++) x[i] = fma(-l[tri(i, j)], x[j], x[i]);
}
// x <- L^-T x
__device__ __forceinline__ void bsub6(const double (&l)[21], double (&x)[6])
{
#pragma unroll
    for (int j = 5; j >= 1; j--)
#pragma unroll
        for (int i = 0; i < j; i++) x[i] = fma(-l[tri(j, i)], x[j], x[i]);
}

// adjugate (packed 00,10,11,20,21,22) and determinant of the symmetric 3x3 [[a,b,d],[b,c,e],[d,e,f]]
__device__ __forceinline__ void adj3(double a, double b, double c, double d, double e, double f, double (&A)[6], double &det)
{
    A[0] = fma(c, f, -(e * e)); A[1] = fma(d, e, -(b * f)); A[2] = fma(a, f, -(d * d));
    A[3] = fma(b, e, -(c * d)); A[4] = fma(b, d, -(a * e)); A[5] = fma(a, c, -(b * b));
    det = fma(a, A[0], fma(b, A[1], d * A[3]));
}

// Inverse of a symmetric positive definite 6x6 (packed lower, 21) through its 3x3 blocks M = [[A,B'],[B,C]] in
// closed form: two reciprocals on the dependent chain instead of the six of an LDL' (measured: the reciprocal
// chain was the largest single item of the v3 profile).  `neg` collects the sign bits of the leading minors.
__device__ __forceinline__ void inv6(const double (&m)[21], double (&o)[21], int &neg)
{
    double aA[6], detA;
    adj3(m[0], m[1], m[2], m[3], m[4], m[5], aA, detA);
    const double r = frcp(detA);
    const double B[3][3] = {{m[6], m[7], m[8]}, {m[10], m[11], m[12]}, {m[15], m[16], m[17]}};
    double X0[3][3];                                   // B adj(A)
#pragma unroll
    for (int i = 0; i < 3; i++) {
        X0[i][0] = fma(B[i][0], aA[0], fma(B[i][1], aA[1], B[i][2] * aA[3]));
        X0[i][1] = fma(B[i][0], aA[1], fma(B[i][1], aA[2], B[i][2] * aA[4]));
        X0[i][2] = fma(B[i][0], aA[3], fma(B[i][1], aA[4], B[i][2] * aA[5]));
    }
    // Schur complement C - r X0 B'
    double Sc[6];
    {
        const double Cm[6] = {m[9], m[13], m[14], m[18], m[19], m[20]};
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
            for (int j = 0; j <= i; j++) {
                const double g = fma(X0[i][0], B[j][0], fma(X0[i][1], B[j][1], X0[i][2] * B[j][2]));
                Sc[tri(i, j)] = fma(-r, g, Cm[tri(i, j)]);
            }
    }
    double aS[6], detS;
    adj3(Sc[0], Sc[1], Sc[2], Sc[3], Sc[4], Sc[5], aS, detS);
    const double r2 = frcp(detS);
    const double aSf[3][3] = {{aS[0], aS[1], aS[3]}, {aS[1], aS[2], aS[4]}, {aS[3], aS[4], aS[5]}};
    double Y0[3][3];                                   // adj(Sc) X0
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) Y0[i][j] = fma(aSf[i][0], X0[0][j], fma(aSf[i][1], X0[1][j], aSf[i][2] * X0[2][j]));
    const double rr2 = r * r2, rrr2 = r * rr2;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j <= i; j++) {
            const double w = fma(X0[0][i], Y0[0][j], fma(X0[1][i], Y0[1][j], X0[2][i] * Y0[2][j]));
            o[tri(i, j)] = fma(rrr2, w, r * aA[tri(i, j)]);
            o[tri(3 + i, 3 + j)] = r2 * aS[tri(i, j)];
        }
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) o[tri(3 + i, j)] = -rr2 * Y0[i][j];
    neg |= __double2hiint(m[0]) | __double2hiint(aA[5]) | __double2hiint(detA) | __double2hiint(Sc[0]) | __double2hiint(aS[5]) | __double2hiint(detS);
}

// 16-byte shared-memory accesses: p must be 16-byte aligned (every row / packed block of WLay is)
__device__ __forceinline__ void ld6(const double *p, double (&v)[6])
{
    const double2 *q = reinterpret_cast<const double2 *>(p);
    const double2 a = q[0], b = q[1], c = q[2];
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y;
}
__device__ __forceinline__ void st6(double *p, const double (&v)[6])
{
    double2 *q = reinterpret_cast<double2 *>(p);
    q[0] = make_double2(v[0], v[1]); q[1] = make_double2(v[2], v[3]); q[2] = make_double2(v[4], v[5]);
}
__device__ __forceinline__ void ld22(const double *p, double (&v)[22])
{
    const double2 *q = reinterpret_cast<const double2 *>(p);
#pragma unroll
    for (int i = 0; i < 11; i++) { const double2 a = q[i]; v[2 * i] = a.x; v[2 * i + 1] = a.y; }
}
__device__ __forceinline__ void st21(double *p, const double (&v)[21])
{
    double2 *q = reinterpret_cast<double2 *>(p);
#pragma unroll
    for (int i = 0; i < 10; i++) q[i] = make_double2(v[2 * i], v[2 * i + 1]);
    p[20] = v[20];
}
__device__ __forceinline__ double dot6(const double (&a)[6], const double (&b)[6])
{
    const double e = fma(a[4], b[4], fma(a[2], b[2], a[0] * b[0])), o = fma(a[5], b[5], fma(a[3], b[3], a[1] * b[1]));
    return e + o;
}

// y = M x for a packed symmetric 6x6 in registers (S = 21, or 22 with the pad)
template <int S>
__device__ __forceinline__ void symmv6(const double (&M)[S], const double (&x)[6], double (&y)[6])
{
#pragma unroll
    for (int i = 0; i < 6; i++) {
        double acc = M[tri(i, 0)] * x[0];
#pragma unroll
        for (int j = 1; j < 6; j++) acc = fma(M[i >= j ? tri(i, j) : tri(j, i)], x[j], acc);
        y[i] = acc;
    }
}

// (Phi x)[c] and (Phi' x)[c] for a 6-vector in registers, Phi = dt blockdiag(Rz, I)   (mpc.cpp:199,210)
struct PhiC { double dc, dsn, dt; };   // dt cos(psi), dt sin(psi), dt
__device__ __forceinline__ void phiT6(const PhiC &F, const double (&x)[6], double (&y)[6])
{
    y[0] = F.dc * x[0] + F.dsn * x[1];
    y[1] = -F.dsn * x[0] + F.dc * x[1];
    y[2] = F.dt * x[2]; y[3] = F.dt * x[3]; y[4] = F.dt * x[4]; y[5] = F.dt * x[5];
}


// 3x3 block of one stance contact, Dblk = sr I + Cb' D Cb = [[a,0,e],[0,b,f],[e,f,c]], from D0..D3 = lam/s of the
// four friction rows and the stored reciprocals ia = 1/a, ib = 1/b, id2 = 1/(Schur complement):
// Dinv (symmetric, 6: 00,01,02,11,12,22)
__device__ __forceinline__ void contact_dinv(double D0, double D1, double D2, double D3, double ia, double ib, double id2,
                                             double mu, double (&Di)[6])
{
    const double e = -mu * (D0 - D1), f = -mu * (D2 - D3);
    const double l20 = e * ia, l21 = f * ib;
    Di[0] = ia + l20 * l20 * id2; Di[1] = l20 * l21 * id2; Di[2] = -l20 * id2;
    Di[3] = ib + l21 * l21 * id2; Di[4] = -l21 * id2; Di[5] = id2;
}
// X = Dblk^-1 Cb' (6 rows x 3) in the cancellation-free closed form: c_i' du must stay accurate to << s_i
// near the solution, which a 3x3 mat-vec in (fx,fy,fz) coordinates cannot deliver.
__device__ __forceinline__ void contact_X(double D0, double D1, double D2, double D3, double ia, double ib, double id2,
                                          double sr, double mu, double (&X)[18])
{
    const double e = -mu * (D0 - D1), f = -mu * (D2 - D3);
    const double iad = ia * id2, ibd = ib * id2;
    double x2 = -mu * (sr + 2.0 * D1) * iad;
    X[0] = (1.0 - e * x2) * ia; X[1] = -f * x2 * ib; X[2] = x2;
    x2 = -mu * (sr + 2.0 * D0) * iad;
    X[3] = (-1.0 - e * x2) * ia; X[4] = -f * x2 * ib; X[5] = x2;
    x2 = -mu * (sr + 2.0 * D3) * ibd;
    X[6] = -e * x2 * ia; X[7] = (1.0 - f * x2) * ib; X[8] = x2;
    x2 = -mu * (sr + 2.0 * D2) * ibd;
    X[9] = -e * x2 * ia; X[10] = (-1.0 - f * x2) * ib; X[11] = x2;
    X[12] = -e * id2 * ia; X[13] = -f * id2 * ib; X[14] = id2;
    X[15] = -X[12]; X[16] = -X[13]; X[17] = -id2;
}

// Per-stage constants every lane keeps in registers.
struct StageC {
    double J[9];      // dt Iw^-1                       (mpc.cpp:220-221,237,251)
    double bv;        // dt / m                         (mpc.cpp:238,251)
    PhiC F;
};

// t_k = Bm sum_c [r x tv_c ; tv_c]  (lane k), optionally also S_k = Bm (sum_c W~ Dinv W~') Bm'
template <class LAY, int HL, bool WITH_S>
__device__ __forceinline__ void assemble_stage(const LAY &L, double *sm, double *smc, const DevParams &P, const StageC &C, int lane)
{
    const int N = L.N, NC = L.NC;
    for (int k = lane; k < N; k += HL) {
        double tt[3] = {0, 0, 0}, tf[3] = {0, 0, 0};
        double Stt[6] = {0, 0, 0, 0, 0, 0}, Sff[6] = {0, 0, 0, 0, 0, 0}, Stf[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};   // tau-tau, F-F (sym), tau-F (full)
        for (int j = 0; j < NC; j++) {
            const int c = k * NC + j;
            const double cf = smc[L.oCf + c];
            if (!(P.fz_max * cf > P.fz_min * cf)) continue;          // swing / eliminated
            const double *r = smc + L.oR + 3 * c, *tv = (L.kTvInScratch ? sm : smc) + L.oTv + 3 * c;
            const double r0 = r[0], r1 = r[1], r2 = r[2];
            tt[0] += r1 * tv[2] - r2 * tv[1]; tt[1] += r2 * tv[0] - r0 * tv[2]; tt[2] += r0 * tv[1] - r1 * tv[0];
            tf[0] += tv[0]; tf[1] += tv[1]; tf[2] += tv[2];
            if (WITH_S) {
                double Di[6];
                ld6(smc + L.oProd + 6 * c, Di);        // Dblk^-1 of this iteration, left there by the per-contact pass (2)
                const double D00 = Di[0], D01 = Di[1], D02 = Di[2], D11 = Di[3], D12 = Di[4], D22 = Di[5];
                // RD = skew(r) Dinv
                const double a00 = -r2 * D01 + r1 * D02, a01 = -r2 * D11 + r1 * D12, a02 = -r2 * D12 + r1 * D22;
                const double a10 = r2 * D00 - r0 * D02, a11 = r2 * D01 - r0 * D12, a12 = r2 * D02 - r0 * D22;
                const double a20 = -r1 * D00 + r0 * D01, a21 = -r1 * D01 + r0 * D11, a22 = -r1 * D02 + r0 * D12;
                // tau-tau += RD skew(r)'   ( (RD R')[i][j] = sum_m RD[i][m] R[j][m] )
                Stt[0] += -r2 * a01 + r1 * a02;                       // (0,0)
                Stt[1] += -r2 * a11 + r1 * a12;                       // (1,0)
                Stt[2] += r2 * a10 - r0 * a12;                        // (1,1)
                Stt[3] += -r2 * a21 + r1 * a22;                       // (2,0)
                Stt[4] += r2 * a20 - r0 * a22;                        // (2,1)
                Stt[5] += -r1 * a20 + r0 * a21;                       // (2,2)
                Stf[0] += a00; Stf[1] += a01; Stf[2] += a02; Stf[3] += a10; Stf[4] += a11; Stf[5] += a12;
                Stf[6] += a20; Stf[7] += a21; Stf[8] += a22;
                Sff[0] += D00; Sff[1] += D01; Sff[2] += D11; Sff[3] += D02; Sff[4] += D12; Sff[5] += D22;   // packed (0,0)(1,0)(1,1)(2,0)(2,1)(2,2)
            }
        }
        const double *J = C.J;
        double *t = sm + L.oVa + 6 * k;
        t[0] = J[0] * tt[0] + J[1] * tt[1] + J[2] * tt[2];
        t[1] = J[3] * tt[0] + J[4] * tt[1] + J[5] * tt[2];
        t[2] = J[6] * tt[0] + J[7] * tt[1] + J[8] * tt[2];
        t[3] = C.bv * tf[0]; t[4] = C.bv * tf[1]; t[5] = C.bv * tf[2];
        if (WITH_S) {
            double *S = sm + L.oS + L.SST * k;
            // omega-omega = J Stt J'
            double JS[9];
#pragma unroll
            for (int i = 0; i < 3; i++)
#pragma unroll
                for (int m = 0; m < 3; m++) {
                    double acc = 0.0;
#pragma unroll
                    for (int q = 0; q < 3; q++) acc += J[3 * i + q] * Stt[q >= m ? tri(q, m) : tri(m, q)];
                    JS[3 * i + m] = acc;
                }
#pragma unroll
            for (int i = 0; i < 3; i++)
#pragma unroll
                for (int m = 0; m <= i; m++) S[tri(i, m)] = JS[3 * i] * J[3 * m] + JS[3 * i + 1] * J[3 * m + 1] + JS[3 * i + 2] * J[3 * m + 2];
            // v-omega = bv (Stf)' J'  :  S[3+i][m] = bv sum_q Stf[q][i] J[m][q]
#pragma unroll
            for (int i = 0; i < 3; i++)
#pragma unroll
                for (int m = 0; m < 3; m++)
                    S[tri(3 + i, m)] = C.bv * (Stf[i] * J[3 * m] + Stf[3 + i] * J[3 * m + 1] + Stf[6 + i] * J[3 * m + 2]);
#pragma unroll
            for (int i = 0; i < 3; i++)
#pragma unroll
                for (int m = 0; m <= i; m++) S[tri(3 + i, 3 + m)] = C.bv * C.bv * Sff[tri(i, m)];
        }
    }
}


// Backward Riccati sweep: for k = N-1..0 factor stage k against P_{k+1} and store G2 | Psi | Gamma; P12 | P22 of
// P_{k+1} are written straight into the stage slot by the previous step.  Per stage, with Pi = P22^-1 (closed form)
// and T = Pi + S_k = L D L' (every lane, registers):  Gamma = T^-1 Pi, A21 = Gamma P21, A22 = Gamma P22 -- lane =
// column of [ I | P21 | P22 ] (18 lanes), one 6x6 mat-vec and two substitutions each; then Psi = P12 S Gamma,
// A11 = P11 - P12 S A21.
// P11 stays in the registers of the six A21 lanes (one column each).  Returns false if a minor was not positive.
// Latency mode (LAT, one problem per two-warp CTA, see srbd_mpc_solve_lat_kernel): a partner warp runs the predictor's pre-pass
// and backward sweep one stage behind this loop.  `lat` = 40 doubles of extra shared memory (the A21 exchange buffer, which
// otherwise lives in vD where the partner writes, and dt cos, dt sin, dt); `fcount` = how
// many factorisations this problem has had.  `lat` costs no shared memory of its own: it is the `prod` region of the layout
// (6 NCN >= 39 doubles), which is dead from the assembly of S_k to the predictor's per-contact pass and which the partner never touches.
// The two warps meet on NAMED BARRIERS (bar.arrive by the producer, bar.sync by the consumer, 64 threads each): barrier 1 + k
// = "stage k of the running factorisation is stored", barrier 0 (free again once the kernel's opening __syncthreads has
// passed) = "the partner has finished its sweep" -- sixteen hardware barriers, hence N <= 15.  A
// barrier id is only re-armed after the main warp has waited on barrier 0, so a lagging partner can never see two arrivals on
// one id.  (Polling a shared-memory word instead -- volatile, or ld.acquire with a back-off -- slowed the MAIN warp's own
// shared-memory chains 1.7-2.5x: measured 524 / 356 us against 213 us for the single warp.  The barrier costs nothing while waiting.)
__device__ __forceinline__ void bar_arrive64(int id) { asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void bar_sync64(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }
constexpr int kLatDoneBar = 0;

template <class LAY, int HL, bool LAT = false>
__device__ __forceinline__ bool factor_stages(const LAY &L, double *sm, const DevParams &P, const StageC &C, int lane PH_PARAMS,
                                              double *lat = nullptr, int fcount = 0)
{
    const int N = L.N;
    (void)fcount;
    double *Ex = LAT ? lat : sm + L.oEx;
    // HL = 32: lanes >= 18 shadow group 2 and never store.  HL = 16 (two problems per warp): groups 0 and 1 as before,
    // lanes 12..15 are columns 0..3 of group 2, and lanes 4, 5 (group 0) take over the P22 update of columns 4, 5.
    const int grp = lane < 18 ? lane / 6 : 2, cb = lane % 6;
    const bool act = lane < 18;
    // Phi column cb:  (M Phi)[:,cb] = ga M[:,ja] + gb M[:,jb]
    const double ga = cb == 0 ? C.F.dc : (cb == 1 ? -C.F.dsn : C.F.dt), gb = cb == 0 ? C.F.dsn : (cb == 1 ? C.F.dc : 0.0);
    const int ja = cb < 2 ? 0 : cb, jb = cb < 2 ? 1 : cb;
    double sq1 = 0.0, sq2 = 0.0;
#pragma unroll
    for (int i = 0; i < 6; i++) if (cb == i) { sq1 = P.scale * P.q[i]; sq2 = P.scale * P.q[6 + i]; }
    asm volatile("" : "+d"(sq1), "+d"(sq2));                    // keep them in registers (no re-derivation per stage)
    // P_N = s Q
    {
        double *St = sm + L.oRic + (N - 1) * L.STG;
        for (int e = lane; e < 58; e += HL) {
            double v = 0.0;
            if (e >= 36) {
#pragma unroll
                for (int d = 0; d < 6; d++) if (e - 36 == tri(d, d)) v = P.scale * P.q[6 + d];
            }
            St[(e < 36 ? L.sP12 : L.sP22 - 36) + e] = v;
        }
    }
    const double m1 = grp == 1 ? 1.0 : 0.0, e1 = 1.0 - m1;
    const bool cf2 = HL == 16 && grp == 0 && cb >= 4;           // closed-form A22 column (see the P22 update)
    const bool st2 = (act && grp == 2) || cf2;
    // stage-slot offsets of the two column stores shared by groups 0 and 1 (group 1 writes into the slot of stage k-1 and
    // into the exchange buffer)
    const int offr = grp == 0 ? L.sG2 + 6 * cb : L.sP12 - L.STG + 6 * cb;
    double *pcol = grp == 0 ? sm + L.oRic + (N - 1) * L.STG + L.sGam + 6 * cb : Ex + 6 * cb;
    const int dcol = grp == 0 ? L.STG : 0;
    double p11[6];                                              // column cb of P11 (meaningful in group 1)
#pragma unroll
    for (int i = 0; i < 6; i++) p11[i] = i == cb ? sq1 : 0.0;
    __syncwarp();
    int neg = 0;
#pragma unroll 1
    for (int k = N - 1; k >= 0; k--) {
        double *St = sm + L.oRic + k * L.STG;
        double *Sk = sm + L.oS + L.SST * k;
        // Pi = P22^-1 in closed form (P22 is well conditioned: adjugate blocks are accurate enough), T = Pi + S_k by
        // LDL' (T spans the full range of the barrier weights: the closed form stalls the interior point there --
        // measured at N = 40, nc = 8; tests/test_ipm_model_cpu.py keeps the experiment)
        double c[6];                                            // issued first: nothing below depends on it until the mat-vec
        ld6(St + L.sP12 + 6 * cb, c);                         // column cb of P21 = row cb of P12 (group 1)
#pragma unroll
        for (int i = 0; i < 6; i++) c[i] = fma(m1, c[i], i == cb ? e1 : 0.0);      // groups 0, 2: unit vector (no branch)
        double Pi[21], T[21], dT[6];
        {
            double ld[22];
            ld22(St + L.sP22, ld);
#pragma unroll
            for (int e = 0; e < 21; e++) T[e] = ld[e];
            inv6(T, Pi, neg);
            ld22(Sk, ld);
#pragma unroll
            for (int e = 0; e < 21; e++) T[e] = Pi[e] + ld[e];
        }
        PHF(8)
        ldl6(T, dT, neg);
        PHF(9)
        // column pipeline: Gamma [ I | P21 | P22 ] -> Gamma | A21 | A22,  Gamma = T^-1 Pi  (A22 = T^-1: Pi P22 = I exactly)
        double col[6];
        {
            symmv6(Pi, c, col);
            if (grp == 2) {
#pragma unroll
                for (int i = 0; i < 6; i++) col[i] = c[i];
            }
            fsub6(T, col);
#pragma unroll
            for (int i = 0; i < 6; i++) col[i] *= dT[i];
            bsub6(T, col);
        }
        PHF(10)
        // S col without the product S * (small): S Gamma = Pi (I - Gamma) =: Theta (symmetric), S A21 = Pi (P21 - A21), i.e.
        // v = Pi (c - col) in groups 0 and 1.  (S (Gamma x) loses eps |S| |P22| relative where the barrier weights are
        // large against P22^-1 -- cond(H) 4e11 stalls on it, tests/golden/ipm_illconditioned_problems.npz -- this form
        // is exact to eps |Pi| there and costs the same.)
        double wv[6], v[6];                                     // v = S col,  wv = P12 v
        {
            double row[6];
#pragma unroll
            for (int i = 0; i < 6; i++) c[i] -= col[i];
            symmv6(Pi, c, v);
#pragma unroll
            for (int i = 0; i < 6; i++) { ld6(St + L.sP12 + 6 * i, row); wv[i] = dot6(row, v); }
        }
        PHF(11)
        // One code path for groups 0 and 1 (no divergence), x = [group 1] P11[:,cb] - wv:
        //   group 0: Psi[:,cb] = wv, G2[:,cb] = Gamma[:,cb] - Phi' wv = col + Phi' x, Gamma[:,cb] = col   (stored as columns)
        //   group 1: A11[:,cb] = x;  row cb of the next P12 = A12[cb,:] + (A11 Phi)[cb,:] = col + Phi' x (A11 symmetric);
        //            A21[:,cb] = col goes to the exchange buffer
        {
            double x[6], px[6];
#pragma unroll
            for (int i = 0; i < 6; i++) x[i] = fma(m1, p11[i], -wv[i]);
            phiT6(C.F, x, px);
#pragma unroll
            for (int i = 0; i < 6; i++) {
                px[i] += col[i];
                p11[i] = i == cb ? x[i] + sq1 : x[i];           // (finite junk outside group 1, multiplied by m1 = 0)
            }
            const bool st01 = lane < 12 && (grp == 0 || k > 0);
            if (st01) { st6(St + offr, px); st6(pcol, col); }
            pcol -= dcol;
            if (grp == 0) st6(St + L.sPsi + 6 * cb, wv);
        }
        double *Sn = St - L.STG;                              // slot of stage k-1: receives P12 | P22 of P_k
        __syncwarp();
        if (LAT) {      // G2 | Psi | Gamma of stage k and P12 | P22 of P_{k+1} are in place: the partner warp may run stage k
            bar_arrive64(1 + k);
        }
        // Theta_k (lower triangle, column cb from group 0) replaces S_k, which nobody reads after this point: the solves
        // form b_k = Gamma' t - Theta p2 with it
        if (lane < 6) {
#pragma unroll
            for (int i = 0; i < 6; i++) if (i >= cb) Sk[tri(i, 0) + cb] = v[i];
        }
        PHF(12)
        if (k > 0) {
            // P22 of P_k, column cb = s Q2 + A22 + Phi' M12 + A21 Phi      (group 2)
            // (every lane runs the arithmetic -- no divergence; only group 2 holds A22 in `col` and stores)
            double m12[6], pm[6], ca[6], cbv[6];
            if (HL == 16) {
                // columns 4 and 5 of A22 = T^-1 = L^-T D^-1 L^-1: their lower parts, operation for operation what the
                // column solve of e_4 / e_5 produces (so that both lane layouts give the same bits)
                const double l54 = T[tri(5, 4)];
                const double x5 = __dmul_rn(-l54, dT[5]), x44 = fma(-l54, x5, dT[4]);
                col[4] = cf2 ? x44 : col[4];
                col[5] = cf2 ? (cb == 4 ? x5 : dT[5]) : col[5];
            }
#pragma unroll
            for (int i = 0; i < 6; i++) m12[i] = Sn[L.sP12 + 6 * i + cb];
            ld6(Ex + 6 * ja, ca); ld6(Ex + 6 * jb, cbv);
            phiT6(C.F, m12, pm);
            double *dst = Sn + L.sP22 + cb;                   // lower triangle: entry (i, cb), i >= cb, at tri(i,0) + cb
#pragma unroll
            for (int i = 0; i < 6; i++) {
                const double val = fma(ga, ca[i], fma(gb, cbv[i], col[i] + pm[i])) + (i == cb ? sq2 : 0.0);
                if (st2 && i >= cb) dst[tri(i, 0)] = val;
            }
            __syncwarp();
        }
        PHF(3)
    }
    return neg >= 0;
}


// Solve (M^-1 + S) y = t with the stored stage factors.  In: t in vA.  Out: zeta~ = Bm^T y in vA (what the
// per-contact step needs: W_c^T y = zeta~_tau x r + zeta~_F).  The two sequential sweeps run on 12 lanes (one state
// component each) and exchange the 12-vector through shared memory once per stage (42 cycles measured for
// STS + sync + LDS against 27 for one dependent shuffle, but 8 instructions instead of 16-28 SHFL).
// pre-pass of one stage (one lane): gt = Gamma^T t, c1 = P12 gt, c2 = Phi^T c1 + P22 gt
template <class LAY>
__device__ __forceinline__ void solve_pre_stage(const LAY &L, double *sm, const PhiC &F, int k)
{
    double *vA = sm + L.oVa, *vB = sm + L.oVb, *vC = sm + L.oVc;
    const double *St = sm + L.oRic + k * L.STG;
    double t[6], gt[6], c1[6], c2[6], pc1[6], row[6], p22[22];
    ld6(vA + 6 * k, t);
#pragma unroll
    for (int j = 0; j < 6; j++) { ld6(St + L.sGam + 6 * j, row); gt[j] = dot6(row, t); }      // (columns of Gamma are stored)
#pragma unroll
    for (int i = 0; i < 6; i++) { ld6(St + L.sP12 + 6 * i, row); c1[i] = dot6(row, gt); }
    ld22(St + L.sP22, p22);
    symmv6(p22, gt, c2);
    phiT6(F, c1, pc1);
#pragma unroll
    for (int i = 0; i < 6; i++) c2[i] += pc1[i];
    st6(vA + 6 * k, gt); st6(vB + 6 * k, c1); st6(vC + 6 * k, c2);
}

// LATP: the pre-pass and the backward sweep of THIS solve have been done by the partner warp (latency mode, predictor only)
template <class LAY, int HL, bool LATP = false>
__device__ __forceinline__ void solve_stages(const LAY &L, double *sm, const StageC &C, int lane PH_PARAMS)
{
    const int N = L.N;
    double *vA = sm + L.oVa, *vB = sm + L.oVb, *vC = sm + L.oVc, *vD = sm + L.oVd, *pbuf = sm + L.oPb;
    const PhiC F = C.F;
    // ---- pre (lane = stage)
    if (!LATP) {
        for (int k = lane; k < N; k += HL) solve_pre_stage<LAY>(L, sm, F, k);
    }
    PHS(8)
    const int r6 = lane % 6;
    if (!LATP) {
        if (lane < 6) { vD[6 * (N - 1) + r6] = 0.0; pbuf[6 * (N & 1) + r6] = 0.0; }     // zero terminal co-state
    }
    __syncwarp();
    // Phi column r6 (for Phi^T x) and Phi row r6 (for Phi x)
    const double ga = r6 == 0 ? F.dc : (r6 == 1 ? -F.dsn : F.dt), gb = r6 == 0 ? F.dsn : (r6 == 1 ? F.dc : 0.0);
    const double fa = r6 == 0 ? F.dc : (r6 == 1 ? F.dsn : F.dt), fb = r6 == 0 ? -F.dsn : (r6 == 1 ? F.dc : 0.0);
    const int ja = r6 < 2 ? 0 : r6, jb = r6 < 2 ? 1 : r6;
    // ---- backward sweep (lane r < 6: p1_r, lane 6+r: p2_r; lanes >= 12 shadow lanes 6..11 and never store):
    //      p1 <- p1 - Psi p2 + c1 ;  p2 <- Phi^T p1 + G2 p2 + c2 ;  vD[k] keeps the p2 that stage k receives.
    //      Branch-free: one predicated store per step; the loads that do not depend on the recurrence (matrix row,
    //      constant term) are issued one step ahead, before the barrier.
    const bool lo6 = lane < 6;
    if (!LATP) {
        const double *rowp = sm + L.oRic + (lo6 ? L.sPsi : L.sG2) + r6;          // row r6 of column storage: + k STG + 6 j
        const double *cp = (lo6 ? vB : vC) + r6;                                        // + 6 k
        double *dstp = lo6 ? pbuf + r6 : vD + r6 - 6;                                   // p1 -> pbuf[6 (k&1)], p2 -> vD[6 (k-1)]
        const bool wr = lane < 12;
        double pown = 0.0, row[6], cadd;
#pragma unroll
        for (int j = 0; j < 6; j++) row[j] = rowp[(N - 1) * L.STG + 6 * j];
        cadd = cp[6 * (N - 1)];
#pragma unroll 2
        for (int k = N - 1; k >= 0; k--) {
            double p2[6];
            ld6(vD + 6 * k, p2);
            const double *pp = pbuf + 6 * ((k + 1) & 1);
            const double pa = pp[ja], pb = pp[jb];
            double acc = dot6(row, p2);
            double b_lo = pown + cadd, b_hi = fma(ga, pa, fma(gb, pb, cadd));           // off the dot-product chain
            // every lane computes both forms: without the barrier ptxas splits the warp into three divergent groups and
            // runs the dot product twice (measured: 274 -> 444 cycles per step when it did)
            asm volatile("" : "+d"(acc), "+d"(b_lo), "+d"(b_hi));
            pown = lo6 ? b_lo - acc : b_hi + acc;
            if (wr && (lo6 || k > 0)) dstp[lo6 ? 6 * (k & 1) : 6 * k] = pown;
            if (k > 0) {
#pragma unroll
                for (int j = 0; j < 6; j++) row[j] = rowp[(k - 1) * L.STG + 6 * j];
                cadd = cp[6 * (k - 1)];
            }
            __syncwarp();
        }
    }
    PHS(9)
    // ---- b_k = gt_k - S_k Gamma p2 = gt_k - Theta_k p2   (lane = stage; Theta_k sits where S_k was, see factor_stages)
    for (int k = lane; k < N; k += HL) {
        double p2[6], sk[22], sg[6], b[6];
        ld6(vD + 6 * k, p2);
        ld22(sm + L.oS + L.SST * k, sk);
        symmv6(sk, p2, sg);
        ld6(vA + 6 * k, b);
#pragma unroll
        for (int i = 0; i < 6; i++) b[i] -= sg[i];
        st6(vA + 6 * k, b);
    }
    if (lane < 6) vB[r6] = 0.0;                     // y1 after stage 0 (y2 after stage 0 is b_0, already in vA)
    __syncwarp();
    PHS(10)
    // ---- forward sweep:  y1 <- y1 + Phi y2 ;  y2 <- b + G2^T y2 - Psi^T y1   (old values on the right)
    {
        const double *g2p = sm + L.oRic + L.sG2 + 6 * r6, *psp = sm + L.oRic + L.sPsi + 6 * r6;   // column r6: + k STG
        double *dstp = (lo6 ? vB : vA) + r6;                                                      // + 6 k
        const bool wr = lane < 12;
        double yown = lo6 ? 0.0 : vA[r6];
        double g2c[6], psc[6], bk;
        ld6(g2p + L.STG, g2c); ld6(psp + L.STG, psc);
        bk = vA[6 + r6];
#pragma unroll 2
        for (int k = 1; k < N; k++) {
            double y1[6], y2[6];
            ld6(vB + 6 * (k - 1), y1); ld6(vA + 6 * (k - 1), y2);
            const double ya = vA[6 * (k - 1) + ja], yb = vA[6 * (k - 1) + jb];
            // four chains of three
            const double a0 = fma(g2c[4], y2[4], fma(g2c[2], y2[2], fma(g2c[0], y2[0], bk)));
            const double a1 = fma(g2c[5], y2[5], fma(g2c[3], y2[3], g2c[1] * y2[1]));
            const double b0 = fma(psc[4], y1[4], fma(psc[2], y1[2], psc[0] * y1[0]));
            const double b1 = fma(psc[5], y1[5], fma(psc[3], y1[3], psc[1] * y1[1]));
            double v_lo = fma(fa, ya, fma(fb, yb, yown)), v_hi = (a0 + a1) - (b0 + b1);
            asm volatile("" : "+d"(v_lo), "+d"(v_hi));          // both forms in every lane: no divergent split (see above)
            yown = lo6 ? v_lo : v_hi;
            if (wr) dstp[6 * k] = yown;
            if (k + 1 < N) { ld6(g2p + (k + 1) * L.STG, g2c); ld6(psp + (k + 1) * L.STG, psc); bk = vA[6 * (k + 1) + r6]; }
            __syncwarp();
        }
    }
    PHS(11)
    // ---- output (lane = stage): y_k = P12^T y1 + P22 y2 + p2 ; zeta~ = Bm^T y_k
    for (int k = lane; k < N; k += HL) {
        const double *St = sm + L.oRic + k * L.STG;
        double y1[6], y2[6], z[6], row[6], p22[22], q[6];
        ld6(vB + 6 * k, y1); ld6(vA + 6 * k, y2); ld6(vD + 6 * k, z);
        ld22(St + L.sP22, p22);
        symmv6(p22, y2, q);
#pragma unroll
        for (int i = 0; i < 6; i++) z[i] += q[i];
#pragma unroll
        for (int j = 0; j < 6; j++) {
            ld6(St + L.sP12 + 6 * j, row);
#pragma unroll
            for (int i = 0; i < 6; i++) z[i] = fma(row[i], y1[j], z[i]);
        }
        const double *J = C.J;
        double o[6];
        o[0] = J[0] * z[0] + J[3] * z[1] + J[6] * z[2];
        o[1] = J[1] * z[0] + J[4] * z[1] + J[7] * z[2];
        o[2] = J[2] * z[0] + J[5] * z[1] + J[8] * z[2];
        o[3] = C.bv * z[3]; o[4] = C.bv * z[4]; o[5] = C.bv * z[5];
        st6(vA + 6 * k, o);
    }
    __syncwarp();
    PHS(12)
}


// Latency mode, partner warp: for every factorisation of the problem, pre-pass + backward-sweep step of stage k as soon as the
// main warp has signalled that stage (one stage behind the factor loop: ~600 cycles of work against ~1,900 per factor stage).
// Same functions, same operations as solve_stages<.., false>: the results are bit-identical.  Returns when the main warp
// signals the end of the problem (a negative value).
template <class LAY>
__device__ __forceinline__ void lat_helper(const LAY &L, double *sm, double *lat, const int *sig, int lane)
{
    const int N = L.N;
    double *vB = sm + L.oVb, *vC = sm + L.oVc, *vD = sm + L.oVd, *pbuf = sm + L.oPb;
    const int r6 = lane % 6;
    const bool lo6 = lane < 6, wr = lane < 12;
    const int ja = r6 < 2 ? 0 : r6, jb = r6 < 2 ? 1 : r6;
    const double *rowp = sm + L.oRic + (lo6 ? L.sPsi : L.sG2) + r6;
    const double *cp = (lo6 ? vB : vC) + r6;
    double *dstp = lo6 ? pbuf + r6 : vD + r6 - 6;
    for (;;) {
        PhiC F = {0.0, 0.0, 0.0};
        double ga = 0.0, gb = 0.0, pown = 0.0;
        for (int k = N - 1; k >= 0; k--) {
            bar_sync64(1 + k);                          // stage k is stored (or, at k = N-1, the problem is over)
            if (k == N - 1) {
                if (sig[0] != 0) return;
                F.dc = lat[36]; F.dsn = lat[37]; F.dt = lat[38];
                ga = r6 == 0 ? F.dc : (r6 == 1 ? -F.dsn : F.dt); gb = r6 == 0 ? F.dsn : (r6 == 1 ? F.dc : 0.0);
                if (lane < 6) { vD[6 * (N - 1) + r6] = 0.0; pbuf[6 * (N & 1) + r6] = 0.0; }     // zero terminal co-state
            }
            if (lane == 0) solve_pre_stage<LAY>(L, sm, F, k);
            __syncwarp();
            double row[6], p2[6];
#pragma unroll
            for (int j = 0; j < 6; j++) row[j] = rowp[k * L.STG + 6 * j];
            const double cadd = cp[6 * k];
            ld6(vD + 6 * k, p2);
            const double *pp = pbuf + 6 * ((k + 1) & 1);
            const double pa = pp[ja], pb = pp[jb];
            double acc = dot6(row, p2);
            double b_lo = pown + cadd, b_hi = fma(ga, pa, fma(gb, pb, cadd));
            asm volatile("" : "+d"(acc), "+d"(b_lo), "+d"(b_hi));
            pown = lo6 ? b_lo - acc : b_hi + acc;
            if (wr && (lo6 || k > 0)) dstp[lo6 ? 6 * (k & 1) : 6 * k] = pown;
            __syncwarp();
        }
        bar_arrive64(kLatDoneBar);
    }
}

// grad f(u) = s r u + W^T (M W u + h): per-stage wrench, state rollout, adjoint sweep (all O(N)), then the
// per-stage co-state wrench in physical units (Bm^T v) is left in V for the per-contact pass.
template <class LAY, int HL>
__device__ __forceinline__ void gradient_sweeps(const LAY &L, double *sm, double *smc, const DevParams &P, const StageC &C, int lane)
{
    const int N = L.N, NC = L.NC;
    double *w = sm + L.oW, *X = sm + L.oXs, *L1 = sm + L.oL1, *V = sm + L.oV;
    const double *E = sm + L.oE;
    const PhiC F = C.F;
    // w_k = Bm sum_c [r x f ; f]    (lane = stage)                                   (mpc.cpp:213-254)
    for (int k = lane; k < N; k += HL) {
        double tt[3] = {0, 0, 0}, tf[3] = {0, 0, 0};
        for (int j = 0; j < NC; j++) {
            const int c = k * NC + j;
            const double *r = smc + L.oR + 3 * c, *f = smc + L.oU + 3 * c;
            tt[0] += r[1] * f[2] - r[2] * f[1]; tt[1] += r[2] * f[0] - r[0] * f[2]; tt[2] += r[0] * f[1] - r[1] * f[0];
            tf[0] += f[0]; tf[1] += f[1]; tf[2] += f[2];
        }
        const double *J = C.J;
        w[6 * k + 0] = J[0] * tt[0] + J[1] * tt[1] + J[2] * tt[2];
        w[6 * k + 1] = J[3] * tt[0] + J[4] * tt[1] + J[5] * tt[2];
        w[6 * k + 2] = J[6] * tt[0] + J[7] * tt[1] + J[8] * tt[2];
        w[6 * k + 3] = C.bv * tf[0]; w[6 * k + 4] = C.bv * tf[1]; w[6 * k + 5] = C.bv * tf[2];
    }
    __syncwarp();
    const int r6 = lane < 6 ? lane : (lane < 12 ? lane - 6 : 0);
    const double ga = r6 == 0 ? F.dc : (r6 == 1 ? -F.dsn : F.dt), gb = r6 == 0 ? F.dsn : (r6 == 1 ? F.dc : 0.0);   // Phi column
    const double fa = r6 == 0 ? F.dc : (r6 == 1 ? F.dsn : F.dt), fb = r6 == 0 ? -F.dsn : (r6 == 1 ? F.dc : 0.0);   // Phi row
    const int ja = r6 < 2 ? 0 : r6, jb = r6 < 2 ? 1 : r6;
    const double sq = lane < 12 ? P.scale * P.q[lane] : 0.0;
    // X[k] = response of state k+1 to the wrenches (zero initial state): (omega, v) part, then (theta, p) part
    if (lane >= 6 && lane < 12) {
        double acc = 0.0;
#pragma unroll
        for (int k = 0; k < N; k++) { acc += w[6 * k + r6]; X[12 * k + lane] = acc; }
    }
    __syncwarp();
    if (lane < 6) {
        double acc = 0.0;
        X[lane] = 0.0;
#pragma unroll
        for (int k = 1; k < N; k++) { acc += fa * X[12 * (k - 1) + 6 + ja] + fb * X[12 * (k - 1) + 6 + jb]; X[12 * k + lane] = acc; }
    }
    __syncwarp();
    // adjoint: lam1 (theta, p) suffix sums, then lam2 (omega, v);  V[k] = lam2 at state k+1
    if (lane < 6) {
        double acc = 0.0;
#pragma unroll
        for (int k = N - 1; k >= 0; k--) { acc += sq * (E[12 * k + lane] + X[12 * k + lane]); L1[6 * k + lane] = acc; }
    }
    __syncwarp();
    if (lane >= 6 && lane < 12) {
        double acc = 0.0;
#pragma unroll
        for (int k = N - 1; k >= 0; k--) {
            acc += sq * (E[12 * k + lane] + X[12 * k + lane]);
            if (k < N - 1) acc += ga * L1[6 * (k + 1) + ja] + gb * L1[6 * (k + 1) + jb];
            V[6 * k + r6] = acc;
        }
    }
    __syncwarp();
    for (int k = lane; k < N; k += HL) {                          // V <- Bm^T V
        double *v = V + 6 * k;
        const double *J = C.J;
        const double v0 = v[0], v1 = v[1], v2 = v[2];
        v[0] = J[0] * v0 + J[3] * v1 + J[6] * v2;
        v[1] = J[1] * v0 + J[4] * v1 + J[7] * v2;
        v[2] = J[2] * v0 + J[5] * v1 + J[8] * v2;
        v[3] *= C.bv; v[4] *= C.bv; v[5] *= C.bv;
    }
    __syncwarp();
}


// Rollout (mpc.cpp:427-446) in closed form; x_opt (gx) is in/out.  Row i >= 1 is the free response of x_current over i
// steps (A_c is nilpotent: base + i rate + i (i-1) quad per component) plus the response X to the wrenches, which the
// last gradient evaluation left in shared memory (zero initial state, same recurrence as the reference's).
// Not inlined: keeps the register allocation of the iteration loop independent of the epilogue.
__device__ __noinline__ void rollout_rows(double *gx, const double *x_current, int last, double *fr, const double *Xs,
                                          double dt, double cz, double sz, int lane, int nl, bool write)
{
    const double *xc = x_current ? x_current : gx;
    if (lane < NS) {
        const int r = lane;
        const double base = xc[r];
        double rate = 0.0, quad = 0.0;
        if (r < 3) {
            const double w0 = xc[6], w1 = xc[7], w2 = xc[8];
            rate = dt * (r == 0 ? (cz * w0 - sz * w1) : (r == 1 ? (sz * w0 + cz * w1) : w2));
        } else if (r < 6) {
            rate = dt * xc[6 + r];
            if (r == 5) quad = 0.5 * dt * dt * xc[12];
        } else if (r == 11) rate = dt * xc[12];
        fr[r] = base; fr[NS + r] = rate; fr[2 * NS + r] = quad;     // [3][13] base | rate | quad (the solve scratch is dead)
        if (x_current && write) gx[r] = base;
    }
    __syncwarp();
    for (int e = NS + lane; e < (write ? (last + 1) * NS : 0); e += nl) {
        const int i = e / NS, r = e % NS;
        const double di = (double)i;
        double val = fma(di * (di - 1.0), fr[2 * NS + r], fma(di, fr[NS + r], fr[r]));
        if (r < 12) val += Xs[12 * (i - 1) + r];
        gx[e] = val;
    }
}

// Start point of the interior point.  Cold (gw == nullptr): the weight shared by the stance contacts of each step,
// clamped into the interior.  Warm (gw = a previous solution u[N,NC,3] of a neighbouring problem, `shift` = how many
// stages the horizon has moved since; the reference asks its solver for a warm start -- setWarmStart(true),
// mpc.cpp:386 / warm_start=True, mpc.py:342 -- but builds a fresh solver per call, so it never gets one): stage k takes
// the force of stage min(k + shift, N-1), pulled `warm_margin` newtons inside every row of the friction pyramid and of the
// fz box, so that the slacks s = b - C u are consistent with u AND bounded away from zero; a contact that was in swing in
// the shifted solution gets the cold value.  lambda = mu0 / s is set at the first pass (2).  tests/ipm_model.py:warm_point
// is the numpy twin.  `active` = this half-warp's problem takes part (control flow stays warp-uniform).
template <class LAY, int HL>
__device__ __forceinline__ void start_point(const LAY &L, double *smc, const DevParams &P, int lane, bool active, const double *gw, int shift)
{
    const int N = L.N, NC = L.NC, NCN = L.NCN;
    double *cfl = smc + L.oCf, *u = smc + L.oU, *sl = smc + L.oSl, *lam = smc + L.oLam;
    const double mu = P.mu;
    for (int c = lane; c < (active ? NCN : 0); c += HL) {
        const int k = c / NC;
        int ns = 0;
        for (int j = 0; j < NC; j++) { const double cj = cfl[k * NC + j]; ns += (P.fz_max * cj > P.fz_min * cj) ? 1 : 0; }
        const double cf = cfl[c], lo = P.fz_min * cf, hi = P.fz_max * cf;
        const bool st = hi > lo;
        double f0 = 0.0, f1 = 0.0, fz = 0.0;
        double *s6 = sl + 6 * c, *l6 = lam + 6 * c;
        if (st) {
            const double span = hi - lo;
            fz = P.weight / (double)ns;
            fz = fmin(fmax(fz, lo + 0.1 * span), hi - 0.1 * span);
            if (gw) {
                const int kp = k + shift < N ? k + shift : N - 1;
                const double *f = gw + (size_t)(kp * NC + (c - k * NC)) * 3;
                const double w0 = f[0], w1 = f[1], w2 = f[2];
                if (w2 > 0.0) {                                 // (false for NaN and for a contact that was in swing)
                    const double m = fmin(P.warm_margin, 0.1 * span);
                    fz = fmin(fmax(w2, lo + m), hi - m);
                    const double lim = fmax(mu * fz - m, 0.0);
                    f0 = fmin(fmax(w0, -lim), lim); f1 = fmin(fmax(w1, -lim), lim);
                }
            }
            s6[0] = mu * fz - f0; s6[1] = mu * fz + f0; s6[2] = mu * fz - f1; s6[3] = mu * fz + f1; s6[4] = hi - fz; s6[5] = fz - lo;
        } else {
#pragma unroll
            for (int i = 0; i < 6; i++) s6[i] = 1.0;
        }
        u[3 * c] = f0; u[3 * c + 1] = f1; u[3 * c + 2] = fz;
#pragma unroll
        for (int i = 0; i < 6; i++) l6[i] = 0.0;
    }
}

// ------------------------------------------------------------------------------------------------
// fused solve kernel: blockDim.x = 32 * warps_per_block.  HL = 32: warp -> problem.  HL = 16: two problems per warp, one
// per half-warp (the 6x6 inverse and LDL' of a stage, identical in every lane, are then issued once for two problems, and
// at the bench batch size no scheduler holds two competing warps).  `lane` is the lane WITHIN the problem throughout; all
// control flow stays warp-uniform (a half that has finished keeps executing with its state frozen), so every
// __syncwarp / shuffle is reached by all 32 lanes.  Both layouts give the same bits per problem.
// ------------------------------------------------------------------------------------------------
// One problem (per HL lanes), `sm` = its workspace of L.oEnd doubles -- shared memory, or (long horizons) global memory that the
// L1/L2 caches keep close: the code is the same, `__syncwarp()` orders both.
// REFILL (two problems per warp, persistent grid): a half-warp whose problem has stopped writes its results and takes the next
// problem from the launch's counter `next` while the other half goes on -- the outputs / prologue of the one run at an iteration
// boundary of the other, whose per-iteration quantities are then recomputed from its (untouched) iterate: same bits as ever.
template <class LAY, int HL, bool LAT = false, bool REFILL = false>
__device__ __forceinline__ void solve_problem(const DevParams &P, const BatchPtrs &io, long long B, const LAY &L, long long b, double *sm,
                                              double *smc, int lane, int half, double *lat = nullptr, unsigned long long *next = nullptr,
                                              unsigned long long next_base = 0)
{
    const int N = L.N, NC = L.NC, NU = 3 * NC, NV = NU * N, NCN = NC * N;
#ifdef G1MPC_PHASE_CLOCKS
    const long long ph_k0 = clock64();
#endif
    if (!REFILL && b - half >= B) return;          // whole warp out of range
    bool valid = b < B;                            // odd batch: the last half-warp repeats the last problem and writes nothing
    bool have = valid;                             // (REFILL) this half-warp holds a problem
    if (!valid) b = B - 1;
    // (smc = base of the per-contact region, offsets below L.oSt; sm = base of the stance list, the per-stage region and the scratch.  They are the
    //  same pointer except in the hybrid workspace variant of the kernel.)
    double *rr = smc + L.oR, *cfl = smc + L.oCf, *u = smc + L.oU, *sl = smc + L.oSl, *lam = smc + L.oLam;
    double *gf = smc + L.oGf, *Dst = smc + L.oDst, *prod = smc + L.oProd, *tv = (L.kTvInScratch ? sm : smc) + L.oTv, *E = sm + L.oE;
    double *dsv = sm + L.oRic, *dlv = sm + L.oRic + 6 * NCN;     // alias: valid only after the corrector solve
    int *stIdx = reinterpret_cast<int *>(sm + L.oSt);      // (the list of stance contacts belongs to the on-chip part: every per-contact pass starts from it)
    const double sr = P.sr, mu = P.mu;

    // per-problem state of this half-warp (set by `prologue`)
    StageC C;
    int fcount = 0;
    double cz = 0.0, sz = 0.0, m_ineq = 6.0;
    int bad_bounds = 0, n_stance = 0;
    bool warm = false;
    int status = G1MPC_MAX_ITER, it = 0;
    double gap = 0.0, rdmax = 0.0, gscale = 1.0, gap_start = 0.0;
    bool lam_init = false;
    // grad f is recomputed from u (two O(N) sweeps) only near the end; while the gap is large it is carried along with
    // the step, grad f += alpha (s r du + W^T y), because the multiplier wrench of the corrector solve IS M W du.  The
    // carried gradient drifts by ~1e-7 relative (the residual of the stage-wise solve), which the exact evaluations of
    // the last iterations remove; the iteration that stops always runs on an exact one (tests/test_ipm_model_cpu.py).
    bool exact = true;
    // `frozen`: this problem has stopped.  With one problem per warp that ends the loop; with two, the warp goes on until
    // both have stopped, the per-contact passes of the frozen one run over zero contacts (n_act) and nothing of its
    // state changes any more.
    bool frozen = false;

    // ---- linearise (mpc.cpp:153-221), stage the inputs, start point: for the half-warps with `act` (all of them unless REFILL).
    //      What a half-warp WITHOUT act writes here it writes again with the values its own problem gave before (inputs and lists
    //      only), its iterate (u, s, lambda, grad f) is not touched.
    auto prologue = [&](const bool act) {
        const double *gx0 = io.x0 + b * NS, *gxr = io.x_ref + b * (N * NS), *gct = io.contact + b * NCN;
        const double *gch = io.c_hor + b * (NCN * 3), *gpc = io.p_com + b * (N * 3);
        Lin lin;
        linearise(P, gxr, N, lin);
        if (act) {
#pragma unroll
            for (int i = 0; i < 9; i++) C.J[i] = lin.Iwi[i] * P.dt;
            C.bv = P.bv;
            C.F.dc = P.dt * lin.c; C.F.dsn = P.dt * lin.s; C.F.dt = P.dt;
            cz = lin.c; sz = lin.s;
        }
        int bb = 0, ns = 0;
        for (int c0 = 0; c0 < NCN; c0 += HL) {
            const int c = c0 + lane;
            bool st = false;
            if (c < NCN) {
            const int k = c / NC;
            rr[3 * c] = gch[3 * c] - gpc[3 * k]; rr[3 * c + 1] = gch[3 * c + 1] - gpc[3 * k + 1]; rr[3 * c + 2] = gch[3 * c + 2] - gpc[3 * k + 2];
            const double cf = gct[c];
            const double lo = P.fz_min * cf, hi = P.fz_max * cf;         // mpc.cpp:344,350
            if (hi < lo) bb = 1;
            cfl[c] = cf;
            st = hi > lo;
            }
            // compacted list of stance contacts (order preserved)
            const unsigned mw = __ballot_sync(kFullMask, st);
            const unsigned m = HL == 32 ? mw : ((mw >> (half * HL)) & ((1u << (HL & 31)) - 1u));      // the bits of this problem's lanes
            if (st) stIdx[ns + __popc(m & ((1u << lane) - 1u))] = c;
            ns += __popc(m);
        }
        if (act) { bad_bounds = bb; n_stance = ns; }
        // free-response error e_k = A^(k+1) x0 - xref_k on the 12 weighted states (mpc.cpp:256-270,369-375)
        {
            double x0r[13];
#pragma unroll
            for (int i = 0; i < 13; i++) x0r[i] = gx0[i];
            const double g0 = x0r[12], dt = P.dt;
            for (int k = lane; k < N; k += HL) {
                const double kk = (double)(k + 1);
                const double *xr = gxr + k * NS;
                double *Ek = E + 12 * k;
                Ek[0] = x0r[0] + kk * dt * (cz * x0r[6] - sz * x0r[7]) - xr[0];
                Ek[1] = x0r[1] + kk * dt * (sz * x0r[6] + cz * x0r[7]) - xr[1];
                Ek[2] = x0r[2] + kk * dt * x0r[8] - xr[2];
                Ek[3] = x0r[3] + kk * dt * x0r[9] - xr[3];
                Ek[4] = x0r[4] + kk * dt * x0r[10] - xr[4];
                Ek[5] = x0r[5] + kk * dt * x0r[11] + (0.5 * kk * (kk - 1.0)) * dt * dt * g0 - xr[5];
                Ek[6] = x0r[6] - xr[6]; Ek[7] = x0r[7] - xr[7]; Ek[8] = x0r[8] - xr[8];
                Ek[9] = x0r[9] - xr[9]; Ek[10] = x0r[10] - xr[10]; Ek[11] = x0r[11] + kk * dt * g0 - xr[11];
            }
        }
        __syncwarp();
        // start point (cold, or from a previous solution)
        if (act) warm = io.u_warm != nullptr;
        start_point<LAY, HL>(L, smc, P, lane, act, (act && warm) ? io.u_warm + b * NV : nullptr, io.warm_shift);
        if (act) {
            m_ineq = 6.0 * (double)(n_stance > 0 ? n_stance : 1);
            status = G1MPC_MAX_ITER; it = 0;
            gap = 0.0; rdmax = 0.0; gscale = 1.0; gap_start = 0.0;
            lam_init = false; exact = true; frozen = false;
        }
        __syncwarp();
    };
    if (REFILL) { prologue(have); if (!have) frozen = true; }
    else prologue(true);

    // ---- outputs: for the half-warps with `fin` (all of them unless REFILL) ------------------------------------
    auto epilogue = [&](const bool fin) {
    // A failed factorisation leaves the scratch in solve-phase state, and with two problems per warp the one that stopped
    // first has had its stage storage reused by the other's later iterations: evaluate w, X, V again from the returned u.
    if (HL == 16 || REFILL || status == G1MPC_NUMERICAL) {
        __syncwarp();
        gradient_sweeps<LAY, HL>(L, sm, smc, P, C, lane);
    }
    __syncwarp();
    // here w = W u, X = state response, gf = grad f(u) are consistent with the returned u
    {
        const unsigned bw = __ballot_sync(kFullMask, bad_bounds != 0);
        bad_bounds = (HL == 32 ? bw : ((bw >> (half * HL)) & ((1u << (HL & 31)) - 1u))) != 0u;
    }
    if (fin && bad_bounds && status == G1MPC_SOLVED) status = G1MPC_BAD_BOUNDS;
    // The slacks are carried by the iteration (ds = -Cb du), not recomputed from u, so after ~25 iterations u can sit
    // ~1e-9 N outside an active row (measured at N = 40, nc = 8).  Clamp the returned point back into the pyramid:
    // the change is six orders below the 1e-3 N parity tolerance and makes primal feasibility exact.
    if (status != G1MPC_NUMERICAL) {
        for (int i = lane; i < (fin ? n_stance : 0); i += HL) {
            const int c = stIdx[i];
            const double cf = cfl[c], lo = P.fz_min * cf, hi = P.fz_max * cf;
            const double fz = fmin(fmax(u[3 * c + 2], lo), hi), lim = mu * fz;
            u[3 * c] = fmin(fmax(u[3 * c], -lim), lim);
            u[3 * c + 1] = fmin(fmax(u[3 * c + 1], -lim), lim);
            u[3 * c + 2] = fz;
        }
    }
    __syncwarp();                                   // (outside the branch: with two problems per warp `status` differs per half)
    double *gu = io.u_opt + b * NV;
    for (int i = lane; i < (fin && valid ? NV : 0); i += HL) gu[i] = u[i];
    if (io.status && lane == 0 && fin && valid) io.status[b] = status;
    if (io.iters && lane == 0 && fin && valid) io.iters[b] = it;
    const double *wv = sm + L.oW, *Xs = sm + L.oXs;
    if (io.resid) {
        double obj2[2] = {0.0, 0.0}, pv = 0.0;      // (two accumulators: see wsum2)
        for (int i = lane, odd = 0; i < NV; i += HL, odd ^= 1) {
            const double t = 0.5 * sr * u[i] * u[i];
            if (HL == 16 && odd) obj2[1] += t; else obj2[0] += t;
        }
        // 1/2 u'Hu + g'u = 1/2 sr |u|^2 + 1/2 sum sQ (e^2 - E^2),  e = E + X
        for (int e = lane, odd = 0; e < 12 * N; e += HL, odd ^= 1) {
            const double t = 0.5 * P.scale * P.q[e % 12] * Xs[e] * (2.0 * E[e] + Xs[e]);
            if (HL == 16 && odd) obj2[1] += t; else obj2[0] += t;
        }
        for (int c = lane; c < NCN; c += HL) {
            const double fx = u[3 * c], fy = u[3 * c + 1], fz = u[3 * c + 2];
            const double cf = cfl[c], lo = P.fz_min * cf, hi = fmax(P.fz_max * cf, lo);
            const double lim = __dmul_rn(mu, fz);                  // the rounded product the clamp above used (no FMA contraction)
            pv = fmax(pv, fmax(fabs(fx) - lim, fabs(fy) - lim));
            pv = fmax(pv, fmax(fz - hi, lo - fz));
        }
        const double obj = wsum2<HL>(obj2);
        pv = wmax<HL>(pv);
        if (lane == 0 && fin && valid) {
            double *r = io.resid + b * 4;
            r[0] = gap; r[1] = rdmax; r[2] = fmax(pv, 0.0); r[3] = obj;
        }
    }
    if (io.y_opt) {
        // multipliers of the reference's 5 rows per contact (OSQP sign: y<0 at a lower bound)
        double *gy = io.y_opt + b * (5 * NCN);
        for (int c = lane; c < (fin && valid ? NCN : 0); c += HL) {
            double y5[5];
            const double cf = cfl[c];
            if (P.fz_max * cf > P.fz_min * cf) {
                const double *l6 = lam + 6 * c;
                y5[0] = -l6[1]; y5[1] = l6[0]; y5[2] = -l6[3]; y5[3] = l6[2]; y5[4] = l6[4] - l6[5];
            } else {
                // swing contact: any sign-feasible multiplier set that balances the gradient
                const double *vk = sm + L.oV + 6 * (c / NC), *r = rr + 3 * c;      // u = 0 here: grad f = W_c^T v
                const double g0 = (vk[1] * r[2] - vk[2] * r[1]) + vk[3], g1 = (vk[2] * r[0] - vk[0] * r[2]) + vk[4];
                const double g2 = (vk[0] * r[1] - vk[1] * r[0]) + vk[5];
                y5[0] = y5[1] = y5[2] = y5[3] = 0.0;
                if (-g0 <= 0.0) y5[0] = -g0; else y5[1] = -g0;
                if (-g1 <= 0.0) y5[2] = -g1; else y5[3] = -g1;
                y5[4] = -g2 - mu * (y5[0] - y5[1] + y5[2] - y5[3]);
            }
#pragma unroll
            for (int i = 0; i < 5; i++) gy[5 * c + i] = y5[i];
        }
    }
    rollout_rows(io.x_opt + b * (N * NS), io.x_current ? io.x_current + b * NS : nullptr, io.one_rollout ? 1 : N - 1,
                 sm + L.oVa, sm + L.oXs, P.dt, cz, sz, lane, HL, fin && valid);
    };

    PH_DECL
#ifdef G1MPC_PHASE_CLOCKS
    ph_t[13] = ph_last - ph_k0;                     // prologue: linearise, stage the inputs, start point
#endif
    const double exact_gap = 1e7 * P.gap_tol;
    int n_act = n_stance;
#pragma unroll 1
    for (;;) {
        PH(7)
    again:
        n_act = frozen ? 0 : n_stance;
        double p_gf = 0.0;
        if (HL == 32 ? exact : __any_sync(kFullMask, exact && !frozen)) {
            // (1) gradient: V <- Bm^T (M W u + h)
            gradient_sweeps<LAY, HL>(L, sm, smc, P, C, lane);
            PH(0)
            const double *Vt = sm + L.oV;
            for (int i = lane; i < (exact ? n_act : 0); i += HL) {
                const int c = stIdx[i], k = c / NC;
                const double *vk = Vt + 6 * k, *r = rr + 3 * c, *uc = u + 3 * c;
                const double g0 = sr * uc[0] + (vk[1] * r[2] - vk[2] * r[1]) + vk[3];
                const double g1 = sr * uc[1] + (vk[2] * r[0] - vk[0] * r[2]) + vk[4];
                const double g2 = sr * uc[2] + (vk[0] * r[1] - vk[1] * r[0]) + vk[5];
                gf[3 * c] = g0; gf[3 * c + 1] = g1; gf[3 * c + 2] = g2;
                p_gf = fmax(p_gf, fmax(fabs(g0), fmax(fabs(g1), fabs(g2))));
            }
        }
        // (2) per stance contact: dual residual, gap, reciprocals and D-state of the 3x3 block,
        //     predictor t = -Dblk^-1 grad f
        {
            // lambda_0 = mu0 / s with mu0 tied to the gradient scale at the start point (cold) or the fixed warm value
            const double gmax = wmax<HL>(p_gf);                 // (outside the branch: the halves of a warp may differ)
            if (!lam_init) {
                const double mu0 = warm ? P.warm_mu0 : fmax(1e-2, 0.1 * gmax);
                gscale = 1.0 + gmax;
                gap_start = mu0;
                for (int i = lane; i < n_act; i += HL) {
                    const int c = stIdx[i];
#pragma unroll
                    for (int j = 0; j < 6; j++) lam[6 * c + j] = mu0 / sl[6 * c + j];
                }
                lam_init = true;
            }
        }
        double p_gap2[2] = {0.0, 0.0}, p_rd = 0.0;
        for (int i = lane, odd = 0; i < n_act; i += HL, odd ^= 1) {
            const int c = stIdx[i];
            const double *s6 = sl + 6 * c, *l6 = lam + 6 * c;
            const double g0 = gf[3 * c], g1 = gf[3 * c + 1], g2 = gf[3 * c + 2];
            double p_gap = (HL == 16 && odd) ? p_gap2[1] : p_gap2[0];        // (see wsum2)
            const double r0 = g0 + l6[0] - l6[1], r1 = g1 + l6[2] - l6[3];
            const double r2 = g2 - mu * (l6[0] + l6[1] + l6[2] + l6[3]) + l6[4] - l6[5];
            p_rd = fmax(p_rd, fmax(fabs(r0), fmax(fabs(r1), fabs(r2))));
            double D[6];
            double *Ds = Dst + 3 * c;
#pragma unroll
            for (int j = 0; j < 6; j++) {
                const double si = frcp(s6[j]);
                p_gap += s6[j] * l6[j];
                D[j] = l6[j] * si;
            }
            // Dblk = sr I + Cb' D Cb = [[a,0,e],[0,b,f],[e,f,cc]] ; LDL' with a cancellation-free Schur term
            const double ia = frcp(sr + D[0] + D[1]), ib = frcp(sr + D[2] + D[3]);
            const double d2 = sr + D[4] + D[5] + mu * mu * ((sr * (D[0] + D[1]) + 4.0 * D[0] * D[1]) * ia +
                                                            (sr * (D[2] + D[3]) + 4.0 * D[2] * D[3]) * ib);
            const double id2 = frcp(d2);
            Ds[0] = ia; Ds[1] = ib; Ds[2] = id2;
            double Di[6];
            contact_dinv(D[0], D[1], D[2], D[3], ia, ib, id2, mu, Di);
            st6(prod + 6 * c, Di);                  // for the assembly of S_k (prod is dead from the step to the predictor pass)
            tv[3 * c] = -(Di[0] * g0 + Di[1] * g1 + Di[2] * g2);
            tv[3 * c + 1] = -(Di[1] * g0 + Di[3] * g1 + Di[4] * g2);
            tv[3 * c + 2] = -(Di[2] * g0 + Di[4] * g1 + Di[5] * g2);
            if (HL == 16 && odd) p_gap2[1] = p_gap; else p_gap2[0] = p_gap;
        }
        {
            const double gap_now = wsum2<HL>(p_gap2) / m_ineq, rd_now = wmax<HL>(p_rd);
            if (!frozen) {
                gap = gap_now; rdmax = rd_now;
                if (!(gap == gap) || !(rdmax == rdmax) || isinf(gap) || isinf(rdmax)) { status = G1MPC_NUMERICAL; frozen = true; }
            }
        }
        {
            // Safeguard of the warm start: a start point whose multipliers are far off stalls close to the boundary (steps of
            // a few per cent for tens of iterations -- seen on walkers that are falling over, tests/ipm_model.py).  If the gap
            // has not halved after kWarmCheckIt iterations the problem restarts from the cold point; the iterations spent count.
            const bool fall = warm && !frozen && it == kWarmCheckIt && gap > 0.5 * gap_start;
            if (HL == 32 ? fall : __any_sync(kFullMask, fall)) {
                __syncwarp();
                start_point<LAY, HL>(L, smc, P, lane, fall, nullptr, 0);
                if (fall) { warm = false; lam_init = false; exact = true; }
                __syncwarp();
                goto again;
            }
        }
        {
            // never stop on a carried gradient: evaluate it and look again
            const bool redo = !frozen && !exact && (gap <= P.gap_tol || it >= P.max_iter);
            if (HL == 32 ? redo : __any_sync(kFullMask, redo)) { if (redo) exact = true; __syncwarp(); goto again; }
        }
        // solved = complementarity gap AND dual residual |grad f + C' lambda|inf <= dual_tol (1 + |grad f(u_start)|inf); the
        // primal residual is zero by construction (slacks start consistent and move with ds = -Cb du; the returned point is
        // clamped).  A gap that converged on a dual residual that did not (long horizons sit on the accuracy floor of the
        // stage-wise solves) is reported, not hidden.
        if (!frozen && (gap <= P.gap_tol || n_stance == 0)) {
            // (no further iterations for the dual residual's sake: below gap_tol the barrier weights lambda/s span more than
            // twenty decades and further Newton steps LOSE accuracy -- measured: three more iterations moved the forces by
            // 3e-4 N and blew the residual of ill-conditioned problems up; the point is reported as it is)
            status = (rdmax <= P.dual_tol * gscale || n_stance == 0) ? G1MPC_SOLVED : G1MPC_DUAL_RESIDUAL;
            frozen = true;
        }
        if (!frozen && it >= P.max_iter) { status = G1MPC_MAX_ITER; frozen = true; }
        if (REFILL) {
            // a half-warp whose problem has stopped: results out, next problem in, while the other half waits at this iteration
            // boundary; then the top of the iteration runs again for both (for the half in mid-solve it recomputes what it had)
            const bool fin = frozen && have;
            if (__any_sync(kFullMask, fin)) {
                __syncwarp();
                epilogue(fin);
                unsigned long long nb = 0;
                if (fin && lane == 0) nb = next_base + atomicAdd(next, 1ULL);
                nb = __shfl_sync(kFullMask, nb, half * HL);
                if (fin) { b = (long long)nb; have = valid = b < B; if (!have) b = B - 1; }
                __syncwarp();
                prologue(fin && have);
                if (__all_sync(kFullMask, !have)) break;
                goto again;
            }
        } else if (HL == 32 ? frozen : __all_sync(kFullMask, frozen)) break;
        n_act = frozen ? 0 : n_stance;
        __syncwarp();

        PH(1)
        // (3) S_k, t_k, then the backward Riccati sweep (shared by predictor and corrector)
        assemble_stage<LAY, HL, true>(L, sm, smc, P, C, lane);
        __syncwarp();
        PH(2)
        if (LAT) { if (lane == 0) { lat[36] = C.F.dc; lat[37] = C.F.dsn; lat[38] = C.F.dt; } }      // (published with the first stage barrier)
        const bool fac_ok = factor_stages<LAY, HL, LAT>(L, sm, P, C, lane PH_ARGS, lat, fcount);
        if (LAT) {      // the partner warp is at most one stage behind: wait for its last step (also when the factorisation failed)
            fcount++;
            bar_sync64(kLatDoneBar);
        }
        if (!fac_ok && !frozen) { status = G1MPC_NUMERICAL; frozen = true; }
        if (!REFILL && (HL == 32 ? frozen : __all_sync(kFullMask, frozen))) break;      // (REFILL: handled at the next iteration boundary)
        n_act = frozen ? 0 : n_stance;
        PH(3)
        // (4) predictor.  With x = ds/s:  dl = -lam (1 + x), and the largest feasible step is
        //     1 / max_i max(-x_i, 1 + x_i)  -- no division per constraint.
        solve_stages<LAY, HL, LAT>(L, sm, C, lane PH_ARGS);
        PH(4)
        const double *zt = sm + L.oVa;
        double p_rmax = 0.0, p_sum2[2] = {0.0, 0.0};
        for (int i = lane, odd = 0; i < n_act; i += HL, odd ^= 1) {
            const int c = stIdx[i], k = c / NC;
            const double *z = zt + 6 * k, *r = rr + 3 * c, *Ds = Dst + 3 * c, *l6 = lam + 6 * c;
            double rt[3], X[18], si[6];
            double p_sum = (HL == 16 && odd) ? p_sum2[1] : p_sum2[0];          // (see wsum2)
#pragma unroll
            for (int j = 0; j < 6; j++) si[j] = frcp(sl[6 * c + j]);
            rt[0] = -gf[3 * c] - ((z[1] * r[2] - z[2] * r[1]) + z[3]);
            rt[1] = -gf[3 * c + 1] - ((z[2] * r[0] - z[0] * r[2]) + z[4]);
            rt[2] = -gf[3 * c + 2] - ((z[0] * r[1] - z[1] * r[0]) + z[5]);
            contact_X(l6[0] * si[0], l6[1] * si[1], l6[2] * si[2], l6[3] * si[3], Ds[0], Ds[1], Ds[2], sr, mu, X);
#pragma unroll
            for (int j = 0; j < 6; j++) {
                const double dsi = -(X[3 * j] * rt[0] + X[3 * j + 1] * rt[1] + X[3 * j + 2] * rt[2]);
                const double x = dsi * si[j];
                const double dli = -l6[j] * (1.0 + x);
                p_rmax = fmax(p_rmax, fmax(-x, 1.0 + x));
                prod[6 * c + j] = dsi * dli;
                p_sum += dsi * dli;
            }
            if (HL == 16 && odd) p_sum2[1] = p_sum; else p_sum2[0] = p_sum;
        }
        const double a_aff = fmin(1.0, frcp(wmax<HL>(p_rmax)));
        // sum (s + a ds)(lam + a dl) = (1 - a) s'lam + a^2 ds'dl for the affine direction
        const double g_aff = ((1.0 - a_aff) * gap * m_ineq + a_aff * a_aff * wsum2<HL>(p_sum2)) / m_ineq;
        double sig = g_aff * frcp(gap);
        sig = sig * sig * sig;
        const double sigma_gap = sig * gap;
        // Safeguard: a predictor blocked almost at once (a_aff < 0.3) carries a second-order term that is far off
        // (the affine step wants to go 1/a_aff times beyond the boundary); taken in full it can lock the iteration into a
        // two-cycle (2 of 131,072 random problems of config 5 ran into max_iter that way).  Scaled by a_aff / 0.3 there;
        // no effect on iterations with a_aff >= 0.3 (tests/test_ipm_model_cpu.py keeps the two problems).
        const double w2 = fmin(1.0, a_aff * (1.0 / 0.3));
        // (5) corrector right-hand side: t = Dblk^-1 (-grad f + Cb'q),  q = (ds dl - sigma mu)/s
        for (int i = lane; i < n_act; i += HL) {
            const int c = stIdx[i];
            const double *Ds = Dst + 3 * c, *l6 = lam + 6 * c;
            double q6[6], Di[6], si[6];
#pragma unroll
            for (int j = 0; j < 6; j++) { si[j] = frcp(sl[6 * c + j]); q6[j] = fma(w2, prod[6 * c + j], -sigma_gap) * si[j]; }
            const double r0 = -gf[3 * c] + (q6[0] - q6[1]);
            const double r1 = -gf[3 * c + 1] + (q6[2] - q6[3]);
            const double r2 = -gf[3 * c + 2] + (-mu * (q6[0] + q6[1] + q6[2] + q6[3]) + q6[4] - q6[5]);
            contact_dinv(l6[0] * si[0], l6[1] * si[1], l6[2] * si[2], l6[3] * si[3], Ds[0], Ds[1], Ds[2], mu, Di);
            tv[3 * c] = Di[0] * r0 + Di[1] * r1 + Di[2] * r2;
            tv[3 * c + 1] = Di[1] * r0 + Di[3] * r1 + Di[4] * r2;
            tv[3 * c + 2] = Di[2] * r0 + Di[4] * r1 + Di[5] * r2;
#pragma unroll
            for (int j = 0; j < 6; j++) prod[6 * c + j] = q6[j];
        }
        __syncwarp();
        assemble_stage<LAY, HL, false>(L, sm, smc, P, C, lane);
        __syncwarp();
        PH(5)
        solve_stages<LAY, HL>(L, sm, C, lane PH_ARGS);
        PH(4)
        // (6) combined direction, step length (the stage storage is dead now: ds | dl live there).
        //     dl = -lam (1 + x + q/lam)
        p_rmax = 0.0;
        for (int i = lane; i < n_act; i += HL) {
            const int c = stIdx[i], k = c / NC;
            const double *z = zt + 6 * k, *r = rr + 3 * c, *Ds = Dst + 3 * c, *l6 = lam + 6 * c, *q6 = prod + 6 * c;
            double rt[3], X[18], Di[6], si[6];
#pragma unroll
            for (int j = 0; j < 6; j++) si[j] = frcp(sl[6 * c + j]);
            const double wy0 = (z[1] * r[2] - z[2] * r[1]) + z[3], wy1 = (z[2] * r[0] - z[0] * r[2]) + z[4];     // W_c^T y
            const double wy2 = (z[0] * r[1] - z[1] * r[0]) + z[5];
            rt[0] = -gf[3 * c] + (q6[0] - q6[1]) - wy0;
            rt[1] = -gf[3 * c + 1] + (q6[2] - q6[3]) - wy1;
            rt[2] = -gf[3 * c + 2] + (-mu * (q6[0] + q6[1] + q6[2] + q6[3]) + q6[4] - q6[5]) - wy2;
            const double D0 = l6[0] * si[0], D1 = l6[1] * si[1], D2 = l6[2] * si[2], D3 = l6[3] * si[3];
            contact_dinv(D0, D1, D2, D3, Ds[0], Ds[1], Ds[2], mu, Di);
            contact_X(D0, D1, D2, D3, Ds[0], Ds[1], Ds[2], sr, mu, X);
            tv[3 * c] = Di[0] * rt[0] + Di[1] * rt[1] + Di[2] * rt[2];
            tv[3 * c + 1] = Di[1] * rt[0] + Di[3] * rt[1] + Di[4] * rt[2];
            tv[3 * c + 2] = Di[2] * rt[0] + Di[4] * rt[1] + Di[5] * rt[2];
#pragma unroll
            for (int j = 0; j < 6; j++) {
                const double dsi = -(X[3 * j] * rt[0] + X[3 * j + 1] * rt[1] + X[3 * j + 2] * rt[2]);
                const double x = dsi * si[j];
                const double rl = 1.0 + x + q6[j] * frcp(l6[j]);
                p_rmax = fmax(p_rmax, fmax(-x, rl));
                dsv[6 * c + j] = dsi; dlv[6 * c + j] = -l6[j] * rl;
            }
            prod[6 * c] = wy0; prod[6 * c + 1] = wy1; prod[6 * c + 2] = wy2;        // q is dead: keep W_c^T y for the gradient
        }
        // fraction to the boundary: 0.995 while the predictor is blocked, -> 1 quadratically once it is not (a_aff -> 1),
        // so that the last iterations are not held at a factor 200 per step (one iteration fewer on every preset)
        const double eta = fmin(0.99999, fmax(0.995, 1.0 - (1.0 - a_aff) * (1.0 - a_aff)));
        const double alpha = fmin(1.0, eta * frcp(wmax<HL>(p_rmax)));
        // (7) step (same lane -> same contacts as in (6): no barrier needed)
        for (int i = lane; i < n_act; i += HL) {
            const int c = stIdx[i];
#pragma unroll
            for (int p = 0; p < 3; p++) {
                u[3 * c + p] += alpha * tv[3 * c + p];
                gf[3 * c + p] += alpha * fma(sr, tv[3 * c + p], prod[6 * c + p]);
            }
#pragma unroll
            for (int j = 0; j < 6; j++) {
                lam[6 * c + j] += alpha * dlv[6 * c + j];
                sl[6 * c + j] += alpha * dsv[6 * c + j];
            }
        }
        if (!frozen) {
            exact = gap * (1.0 - alpha * (1.0 - sig)) < exact_gap;   // predicted gap of the next iterate
            it++;
        }
        __syncwarp();
        PH(6)
    }

    if (!REFILL) epilogue(true);
#ifdef G1MPC_PHASE_CLOCKS
    __syncwarp();
    PH(14)                                          // epilogue: outputs + rollout
    if (io.y_opt && lane == 0 && valid) { for (int i = 0; i < 16; i++) io.y_opt[b * (5 * NCN) + i] = (double)ph_t[i]; }
#endif
}

// ------------------------------------------------------------------------------------------------
// The launchable kernel.  WSG = false: warp -> problem(s), workspace in dynamic shared memory (the N = 10 path: everything
// of a problem on chip).  WSG = true: the workspace of a problem lives in GLOBAL memory (io.workspace, one slot per resident
// warp) and the grid is persistent -- warp w takes problems w, w + W, w + 2W, ... .  For the long-horizon shapes (n = 480-960,
// 80-160 KB per problem) shared memory held one or two warps per SM; this way the register file is the only limit
// (8 warps per SM) and the caches carry the working set.  WSG = 2 (hybrid): see the branch below.
// ------------------------------------------------------------------------------------------------
template <class LAY, int HL, int WSG>
__global__ void __launch_bounds__(128) srbd_mpc_solve_warp_kernel(const DevParams P, const BatchPtrs io, long long B, const LAY L)
{
    constexpr int PPW = 32 / HL;
    extern __shared__ __align__(16) double sm_all[];
    const int lane = threadIdx.x & (HL - 1), half = (threadIdx.x & 31) / HL, warp = threadIdx.x >> 5;
    if (WSG == 0) {
        double *sm = sm_all + (size_t)(warp * PPW + half) * L.oEnd;
        const long long b = ((long long)blockIdx.x * (blockDim.x >> 5) + warp) * PPW + half;
        solve_problem<LAY, HL>(P, io, B, L, b, sm, sm, lane, half);
    } else if (WSG == 1) {
        const long long w = (long long)blockIdx.x * (blockDim.x >> 5) + warp, nW = (long long)gridDim.x * (blockDim.x >> 5);
        double *sm = io.workspace + (size_t)(w * PPW + half) * L.oEnd;
        for (long long b0 = w * PPW; b0 < B; b0 += nW * PPW) {
            solve_problem<LAY, HL>(P, io, B, L, b0 + half, sm, sm, lane, half);
            __syncwarp();
        }
    } else {
        // hybrid: the stance list, the per-stage region and the scratch (offsets >= oSt: the dependent chains of the factorisation and
        // of the sweeps) in shared memory, the per-contact region (offsets < oSt: streamed by independent per-contact passes) in a
        // global-memory slot per resident warp, persistent grid
        const long long w = (long long)blockIdx.x * (blockDim.x >> 5) + warp, nW = (long long)gridDim.x * (blockDim.x >> 5);
        double *sm = sm_all + (size_t)(warp * PPW + half) * (L.oEnd - L.oSt) - L.oSt;
        double *smc = io.workspace + (size_t)(w * PPW + half) * L.oSt;
        for (long long b0 = w * PPW; b0 < B; b0 += nW * PPW) {
            solve_problem<LAY, HL>(P, io, B, L, b0 + half, sm, smc, lane, half);
            __syncwarp();
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Two problems per warp with REFILL (see solve_problem): persistent grid of one-warp CTAs, workspace in shared memory.  The warp no
// longer lasts as long as the slower of a fixed pair (E[max of two iteration counts] = 14.5 against a mean of 13.6 on config 3), and
// the launch has no tail of half-empty warps.  `next` counts the problems handed out after the first gridDim.x * 2.
// ------------------------------------------------------------------------------------------------
template <class LAY, int HL>
__global__ void __launch_bounds__(32) srbd_mpc_solve_refill_kernel(const DevParams P, const BatchPtrs io, long long B, const LAY L,
                                                                    unsigned long long *next)
{
    constexpr int PPW = 32 / HL;
    extern __shared__ __align__(16) double sm_all[];
    const int lane = threadIdx.x & (HL - 1), half = (threadIdx.x & 31) / HL;
    double *sm = sm_all + (size_t)half * L.oEnd;
    solve_problem<LAY, HL, false, true>(P, io, B, L, (long long)blockIdx.x * PPW + half, sm, sm, lane, half, nullptr, next,
                                        (unsigned long long)gridDim.x * PPW);
}

// ------------------------------------------------------------------------------------------------
// Latency mode: ONE problem per CTA of two warps (small batches -- the reference's own use is one updateMPC per control tick,
// mpc_to_wbid_node.cpp:494).  Warp 0 runs solve_problem as above; warp 1 runs the predictor's pre-pass and backward sweep one
// stage behind warp 0's factor loop (lat_helper), so that ~3 k of the ~36 k cycles of an interior-point iteration leave the
// critical path; ~1.4 k come back as the wait for the partner's last stage and as contention in the factor loop.  The
// corrector's right-hand side depends on the predictor's result, so only this one of the four sweeps of an iteration can hide.
// Measured (profiles/r2_f_latency_mode.txt): 213.0 -> 203.7 us p50 over the configs[0] replay, 342.9 -> 328.6 us at 148
// problems, bit-identical to the one-warp kernel.  The warps meet on named barriers (see factor_stages).
// ------------------------------------------------------------------------------------------------
template <class LAY, int HL>
__global__ void __launch_bounds__(64) srbd_mpc_solve_lat_kernel(const DevParams P, const BatchPtrs io, long long B, const LAY L)
{
    // HL = 32: one problem per CTA.  HL = 16: two (one per half-warp of BOTH warps: the partner warp sweeps the two problems of
    // its main warp in the same instructions).
    constexpr int PPW = 32 / HL;
    extern __shared__ __align__(16) double sm_all[];
    __shared__ int cta_done;                            // end-of-CTA flag for the partner warp
    __shared__ unsigned slot[2];                        // hardware warp slots of the two warps
    const int lane = threadIdx.x & (HL - 1), half = (threadIdx.x & 31) / HL, warp = threadIdx.x >> 5;
    double *sm = sm_all + (size_t)half * L.oEnd, *lat = sm + L.oProd;
    if (threadIdx.x == 0) cta_done = 0;
    if ((threadIdx.x & 31) == 0) { unsigned w; asm volatile("mov.u32 %0, %%warpid;" : "=r"(w)); slot[warp] = w; }
    __syncthreads();
    // Which of the two warps is the main one: a scheduler serves the warp slots congruent to it mod 4, and the two warps of a
    // CTA sit on neighbouring slots.  If warp 0 were always the main warp, the four CTAs of an SM would put their main warps on
    // schedulers 0, 2, 0, 2 and leave 1 and 3 to partners that mostly wait (measured: 2.91 against 3.59 M/s at 65,536 problems).
    // Alternating by slot group (slots 0-3, 4-7, ...) puts one main warp on every scheduler.  Both warps evaluate the same rule.
    const unsigned s_lo = slot[0] < slot[1] ? slot[0] : slot[1];
    const int main_warp = (((s_lo >> 2) & 1) == 0) == (slot[0] <= slot[1]) ? 0 : 1;
    if (warp == main_warp) {
        solve_problem<LAY, HL, true>(P, io, B, L, (long long)blockIdx.x * PPW + half, sm, sm, lane, half, lat);
        // release the partner, which waits for a stage that will not come
        __syncwarp();
        if ((threadIdx.x & 31) == 0) cta_done = 1;
        __syncwarp();
        bar_arrive64(1 + (L.N - 1));
    } else {
        lat_helper<LAY>(L, sm, lat, &cta_done, lane);
    }
}

} // namespace v2
