// node_adaptors.cuh -- the two callers either side of the hot path, batched on the device.
// Included by g1mpc_kernels.cu inside its anonymous namespace.
//
//   walk_inputs_kernel : what MPCNode::update builds before updateMPC (g1_mpc/src/mpc_to_wbid_node.cpp:385-470):
//                        reference trajectory (:403-435), p_com extraction (:461-468), contact points tiled over the
//                        horizon (:50-60), gaitPlanner's table window (:82-127,171-187) and the Raibert footstep
//                        overwrite (:205-309,507-524).  HBM-write bound: 8*(7+4+12+3)*N bytes out per problem.
//   wbid_refs_kernel   : what WBID.setReference derives from one solve (g1_mujoco_sim/src/wbid.py:244-297, called at
//                        ros_run_simulation.py:409 with x_opt[1], u_opt0 and the current foot points).
//
// All arithmetic is written with explicit round-to-nearest intrinsics in the reference's evaluation order, so the
// results are bit-identical to the CPU restatement (no FMA contraction); only sin/cos may differ in the last ulp.
#pragma once

namespace adaptors {

// mpc_to_wbid_node.cpp:84-127: 20 rows x [L-lower, L-upper, R-lower, R-upper]; row r, contact j as bit (3-j)
//   table A (gait_phase < 10): 5x1111 5x0011 5x1100 5x0011 ; table B: 5x1100 5x0011 5x1100 5x0011
__device__ __forceinline__ int table_row(bool late, int r)
{
    const int blk = r / 5;                        // 0..3
    if (blk == 0) return late ? 0xC : 0xF;
    return (blk & 1) ? 0x3 : 0xC;
}

constexpr double kVcmdX = 0.35;                   // :405
constexpr double kRefY = 7.44339342e-05, kRefZ = 5.97983255e-01;   // :412-413
constexpr double kFootY = 0.118508, kFootLowX = 0.0293, kFootUpX = 0.1406;   // :217-219
constexpr double kStance = 0.25, kVelGain = 0.02; // :27, :521

struct WalkPtrs {
    const double *x0, *contacts_now, *sim_time, *x_opt_prev;
    double *x_ref, *contact, *c_hor, *p_com, *landing;
};

// one warp per problem; lanes stride over the output elements
__global__ void __launch_bounds__(128) walk_inputs_kernel(const WalkPtrs io, long long B, int N, double dt, double g)
{
    const int lane = threadIdx.x & 31;
    const long long b = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (b >= B) return;
    const double x0x = io.x0[b * NS + 3];
    const int phase = (int)floor(__ddiv_rn(io.sim_time[b], dt));          // :105
    const bool late = phase >= 10;                                         // :106
    const int k = phase % N, sub = k % 5;                                  // :171, :213
    const int row0 = table_row(late, k);
    const bool single = __popc(row0) == 2;                                 // :222-227
    // Raibert landing points from the previous solve's x_opt rows 5-sub and 10-sub (:238-243, :276-281)
    double land[2][2] = {{0, 0}, {0, 0}};
    const int i1 = 5 - sub, i2 = 10 - sub;             // first loop rows [i1, i2), second [i2, 10)  (:245, :284)
    const int hi1 = i2 < N ? i2 : N;                   // N < 10: the reference would index past its buffers; clamp
    if (single) {
#pragma unroll
        for (int w = 0; w < 2; w++) {
            const int ir = w == 0 ? i1 : i2;
            if (ir >= N) continue;                 // the reference reads past the end here, but then never uses the value
            const double *xo = io.x_opt_prev + (b * N + ir) * NS;
#pragma unroll
            for (int a = 0; a < 2; a++) {
                const double p = xo[3 + a], v = xo[9 + a];
                const double vd = a == 0 ? kVcmdX : io.x_ref[(b * N + ir) * NS + 10];   // x_ref(i,9) = v_cmd(3); column 10 is never written (:424-425)
                const double t1 = __dmul_rn(__dmul_rn(__dmul_rn(v, kStance), 0.5), dt);
                land[w][a] = __dadd_rn(__dadd_rn(p, t1), __dmul_rn(kVelGain, __dsub_rn(v, vd)));   // :523-524
            }
        }
    }
    // x_ref columns written by the node (:410-433); the other columns keep the caller's values (x_ref is in/out)
    for (int i = lane; i < N; i += 32) {
        double *xr = io.x_ref + (b * N + i) * NS;
        const double px = __dadd_rn(x0x, __dmul_rn(__dmul_rn(kVcmdX, (double)i), dt));
        xr[2] = 0.0; xr[3] = px; xr[4] = kRefY; xr[5] = kRefZ; xr[8] = 0.0; xr[9] = kVcmdX; xr[12] = g;
        double *pc = io.p_com + (b * N + i) * 3;                           // :461-468
        pc[0] = px; pc[1] = kRefY; pc[2] = kRefZ;
    }
    for (int e = lane; e < 4 * N; e += 32) {                               // :175-187
        const int i = e >> 2, j = e & 3;
        io.contact[b * (4 * N) + e] = (double)((table_row(late, i + k) >> (3 - j)) & 1);
    }
    for (int e = lane; e < 12 * N; e += 32) {
        const int i = e / 12, c = e % 12;
        double v = io.contacts_now[b * 12 + c];                            // :50-60
        if (single && i >= i1 && i < 10) {
            const int w = i < i2 ? 0 : 1;                                  // :245, :284
            const bool left_swing = ((table_row(late, i + k) >> 3) & 1) == 0;   // contact_horizon[i][0] == 0
            const int foot = c / 6, pt = (c % 6) / 3, ax = c % 3;          // foot 0 = left; pt 0 = lower (heel), 1 = upper (toe)
            if (ax < 2 && foot == (left_swing ? 0 : 1)) {
                if (ax == 0) v = pt == 0 ? __dsub_rn(land[w][0], kFootLowX) : __dadd_rn(land[w][0], kFootUpX);   // :250,254 / :262,266
                else v = left_swing ? __dadd_rn(land[w][1], kFootY) : __dsub_rn(land[w][1], kFootY);            // :251,255 / :263,267
            }
        }
        io.c_hor[b * (12 * N) + e] = v;
    }
    if (io.landing && lane == 0) {
        // placeholder the node publishes (:257-258, :269-270): set by the first loop only, opposite-sign y as coded
        double lx = 0.0, ly = 0.0;
        if (single && i1 < hi1) {
            const int il = hi1 - 1;                                        // last row of the first loop decides
            const bool left_swing = ((table_row(late, il + k) >> 3) & 1) == 0;
            lx = land[0][0];
            ly = left_swing ? __dsub_rn(land[0][1], kFootY) : __dadd_rn(land[0][1], kFootY);
        }
        io.landing[b * 2] = lx; io.landing[b * 2 + 1] = ly;
    }
}

// refs[b] = R(9) | velocity(6) | acceleration_reference(6) | com_pos(3) | com_vel(3) | com_acc(3) | wrench refs(12)
constexpr int kWbidRefs = 42;

__global__ void __launch_bounds__(128) wbid_refs_kernel(const double *u_opt, const double *x_opt, const double *feet, double *refs,
                                                         long long B, int N, double wbid_mass)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const double *x1 = x_opt + (b * N + 1) * NS;                           // self.x_opt[1, :]  (ros_run_simulation.py:409)
    const double *u0 = u_opt + b * (N * 12);                               // u_opt0 = row 0
    const double *ft = feet + b * 12;
    double *o = refs + b * kWbidRefs;
    // R = euler_matrix(roll, pitch, yaw)[0:3,0:3], static xyz axes (wbid.py:249-250)
    double si, ci, sj, cj, sk, ck;
    sincos(x1[0], &si, &ci); sincos(x1[1], &sj, &cj); sincos(x1[2], &sk, &ck);
    const double cc = __dmul_rn(ci, ck), cs = __dmul_rn(ci, sk), sc = __dmul_rn(si, ck), ss = __dmul_rn(si, sk);
    o[0] = __dmul_rn(cj, ck); o[1] = __dsub_rn(__dmul_rn(sj, sc), cs); o[2] = __dadd_rn(__dmul_rn(sj, cc), ss);
    o[3] = __dmul_rn(cj, sk); o[4] = __dadd_rn(__dmul_rn(sj, ss), cc); o[5] = __dsub_rn(__dmul_rn(sj, cs), sc);
    o[6] = -sj;               o[7] = __dmul_rn(cj, si);                o[8] = __dmul_rn(cj, ci);
    // velocity = (linear, angular)  (wbid.py:259-261)
    o[9] = x1[9]; o[10] = x1[10]; o[11] = x1[11]; o[12] = x1[6]; o[13] = x1[7]; o[14] = x1[8];
    // angular acceleration reference = I_torso^-1 sum_i (r_i x omega),  r_i = foot_i - p   (wbid.py:264-283)
    const double w0 = x1[6], w1 = x1[7], w2 = x1[8];
    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const double r0 = __dsub_rn(ft[3 * i], x1[3]), r1 = __dsub_rn(ft[3 * i + 1], x1[4]), r2 = __dsub_rn(ft[3 * i + 2], x1[5]);
        s0 = __dadd_rn(s0, __dsub_rn(__dmul_rn(r1, w2), __dmul_rn(r2, w1)));
        s1 = __dadd_rn(s1, __dsub_rn(__dmul_rn(r2, w0), __dmul_rn(r0, w2)));
        s2 = __dadd_rn(s2, __dsub_rn(__dmul_rn(r0, w1), __dmul_rn(r1, w0)));
    }
    o[15] = 0.0; o[16] = 0.0; o[17] = 0.0;
    o[18] = __dmul_rn(__ddiv_rn(1.0, 8.20564e-2), s0); o[19] = __dmul_rn(__ddiv_rn(1.0, 8.05015e-2), s1); o[20] = __dmul_rn(__ddiv_rn(1.0, 0.32353e-2), s2);
    // CoM task (wbid.py:287-294): sum_forces = np.sum(np.reshape(u_opt0, (3, 4)), axis=1) -- as coded this sums
    // consecutive groups of four entries of u_opt0, not the per-axis totals; reproduced verbatim.
    o[21] = x1[3]; o[22] = x1[4]; o[23] = x1[5];
    o[24] = x1[9]; o[25] = x1[10]; o[26] = x1[11];
    const double grav[3] = {0.0, 0.0, -9.80665};
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const double sf = __dadd_rn(__dadd_rn(__dadd_rn(u0[4 * a], u0[4 * a + 1]), u0[4 * a + 2]), u0[4 * a + 3]);
        o[27 + a] = __dadd_rn(__ddiv_rn(sf, wbid_mass), grav[a]);
    }
    // wrench task references (wbid.py:296-297)
#pragma unroll
    for (int i = 0; i < 12; i++) o[30 + i] = u0[i];
}

// Closed loop of many walkers: the SRBD model itself is the plant.  After a solve, record the tick, then
//   x0 <- x_opt[1] (the first rollout step, mpc.cpp:436-441 = A x0 + B_0 u_0), contact points <- row 1 of the planner's
//   c_horizon (stance feet stay, a foot due to land sits at its Raibert point), sim_time += dt.
// The previous x_opt stays where it is: the next walk_inputs call reads it for the Raibert rule (:238-243).
__global__ void __launch_bounds__(128) advance_kernel(double *x0, double *contacts_now, double *sim_time, const double *x_opt,
                                                       const double *u_opt, const double *c_hor, const int32_t *status,
                                                       const int32_t *iters, double *traj_x, double *traj_u, int32_t *status_any,
                                                       int32_t *iters_sum, long long B, int N, int tick, int ticks, double dt)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const double *x1 = x_opt + (b * N + 1) * NS, *u0 = u_opt + b * (N * 12), *c1 = c_hor + (b * N + 1) * 12;
    double *x = x0 + b * NS;
    if (traj_x) for (int i = 0; i < NS; i++) traj_x[(b * ticks + tick) * NS + i] = x[i];
    if (traj_u) for (int i = 0; i < 12; i++) traj_u[(b * ticks + tick) * 12 + i] = u0[i];
    for (int i = 0; i < NS; i++) x[i] = x1[i];
    for (int i = 0; i < 12; i++) contacts_now[b * 12 + i] = c1[i];
    sim_time[b] = __dadd_rn(sim_time[b], dt);
    if (status_any) status_any[b] = (tick == 0 ? 0 : status_any[b]) | status[b];
    if (iters_sum) iters_sum[b] = (tick == 0 ? 0 : iters_sum[b]) + iters[b];
}

// ---------------------------------------------------------------------------------------------------------------------
// SURVEY 8f-1, Python half: GaitPlanner.plan_gait (gait_planner.py:27-62) + its mapping onto contact / c_horizon
// (main.py:259-285).  One thread per (problem, horizon step): it walks the phase list exactly as plan_gait builds it --
// the same floating-point additions in the same order (start times and foot positions are running sums) -- and stops at the
// phase that contains t_i = current_time + i dt; past the end the last phase applies (main.py:270-271).
//   phase 0: [0, ds) "both"; phase 1 + i: single support, support leg left for even i (gait_planner.py:40-51); after phase
//   1 + i the swing foot moves ahead by step_length (i == 0) or 2 step_length (:55-62).  Left foot starts at (0, -w), right
//   at (0, +w) (:19-20, as coded).
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) plan_gait_kernel(const double *planner, const double *current_time, double foot_offset,
                                                         double dt, double *contact, double *c_hor, long long B, int N)
{
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= B * N) return;
    const long long b = e / N;
    const int i = (int)(e % N);
    const double *pl = planner + b * 5;
    const double ds = pl[0], ss = pl[1], sl = pl[2], sw = pl[3];
    const int nsteps = (int)pl[4];
    const double t = __dadd_rn(current_time[b], __dmul_rn((double)i, dt));      // main.py:263
    double lx = 0.0, ly = -sw, rx = 0.0, ry = sw;                               // gait_planner.py:19-20
    double start = 0.0, end = __dadd_rn(0.0, ds);                               // :31-32
    int support = 0;                                                            // 0 both, 1 left, 2 right
    bool found = start <= t && t < end;                                         // main.py:266
    for (int k = 0; k < nsteps && !found; k++) {
        start = end;                                                            // current_time after the previous phase (:37,52)
        end = __dadd_rn(start, ss);                                             // :46
        support = (k % 2 == 0) ? 1 : 2;                                         // :41
        found = start <= t && t < end;
        if (!found && k + 1 < nsteps) {                                         // the feet of the NEXT phase (:55-62); the last phase keeps its own
            const double inc = k == 0 ? sl : __dmul_rn(2.0, sl);
            if (support == 1) rx = __dadd_rn(rx, inc); else lx = __dadd_rn(lx, inc);
        }
    }
    // (not found at all: the loop has left the last phase's support leg and feet in place, main.py:270-271)
    const bool lc = support == 0 || support == 1, rc = support == 0 || support == 2;   // main.py:282-283
    double *ct = contact + e * 4, *ch = c_hor + e * 12;
    ct[0] = lc ? 1.0 : 0.0; ct[1] = ct[0]; ct[2] = rc ? 1.0 : 0.0; ct[3] = ct[2];
    ch[0] = __dsub_rn(lx, foot_offset); ch[1] = ly; ch[2] = 0.0;                // left heel   (main.py:276)
    ch[3] = __dadd_rn(lx, foot_offset); ch[4] = ly; ch[5] = 0.0;                // left toe    (:277)
    ch[6] = __dsub_rn(rx, foot_offset); ch[7] = ry; ch[8] = 0.0;                // right heel  (:278)
    ch[9] = __dadd_rn(rx, foot_offset); ch[10] = ry; ch[11] = 0.0;              // right toe   (:279)
}

// testing_double_support (main.py:177-197): the CoM reference of horizon step i by the quarter of the run t_i falls in.
__global__ void __launch_bounds__(128) double_support_ref_kernel(const double *current_time, double total, double dt, double *x_ref,
                                                                  long long B, int N)
{
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= B * N) return;
    const long long b = e / N;
    const double t = __dadd_rn(current_time[b], __dmul_rn((double)(e % N), dt));
    double rx = 0.0, ry = 0.0, rz = 1.0;
    if (t < __ddiv_rn(total, 4.0)) ry = 0.1;
    else if (t < __ddiv_rn(total, 2.0)) { }
    else if (t < __ddiv_rn(__dmul_rn(3.0, total), 4.0)) rz = 0.7;
    else rx = 0.01;
    double *xr = x_ref + e * NS;
    xr[3] = rx; xr[4] = ry; xr[5] = rz;
}

// main.py:222-228,306-310: the state the NEXT solve starts from is x_opt[2] of this tick's rollout (x0 = x_ref[0] at the
// first tick), the state the next rollout starts from is x_opt[1].  `x0` / `carry` are [B,13] work arrays.
__global__ void __launch_bounds__(128) main_pre_kernel(double *x0, double *carry, const double *x_ref, const double *x_opt,
                                                        long long B, int N, int tick)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    for (int i = 0; i < NS; i++) {
        const double v = tick == 0 ? x_ref[b * N * NS + i] : x_opt[(b * N + 2) * NS + i];     // main.py:226,228
        x0[b * NS + i] = v;
        carry[b * NS + i] = tick == 0 ? v : x_opt[(b * N + 1) * NS + i];                     // main.py:306-310
    }
}
__global__ void __launch_bounds__(128) main_post_kernel(double *current_time, const double *x_opt, const double *u_opt,
                                                         const int32_t *status, const int32_t *iters, double *traj_x, double *traj_u,
                                                         int32_t *status_any, int32_t *iters_sum, long long B, int N, int tick,
                                                         int ticks, double dt)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    if (traj_x) for (int i = 0; i < NS; i++) traj_x[(b * ticks + tick) * NS + i] = x_opt[b * N * NS + i];       // main.py:314-321
    if (traj_u) for (int i = 0; i < 12; i++) traj_u[(b * ticks + tick) * 12 + i] = u_opt[b * N * 12 + i];       // main.py:322
    current_time[b] = __dadd_rn(current_time[b], dt);                                                          // main.py:326
    if (status_any) status_any[b] = (tick == 0 ? 0 : status_any[b]) | status[b];
    if (iters_sum) iters_sum[b] = (tick == 0 ? 0 : iters_sum[b]) + iters[b];
}

} // namespace adaptors
