/*
 * g1mpc_b200.h -- C ABI of the B200-native batched SRBD-MPC hot path.
 *
 * Drop-in boundary for g1_mpc::MPC::updateMPC (reference g1_mpc/include/mpc.hpp:77-82,
 * g1_mpc/src/mpc.cpp:448-473) and the Python twin MPC.update (g1_mpc/scripts/mpc.py:360-377), extended
 * from one problem to a batch of B independent problems.  Plain pointers and sizes only; no torch,
 * Eigen or CUDA types in the signatures (`stream` is a cudaStream_t passed as void*, NULL = default).
 *
 * Layouts (all double, row-major, contiguous, problem-major):
 *   x0        [B,13]        state  [roll,pitch,yaw, px,py,pz, wx,wy,wz, vx,vy,vz, g]   (mpc.py:30; setX0 mpc.hpp:91)
 *   x_ref     [B,N,13]      reference trajectory                                       (setXRefHor mpc.hpp:95)
 *   contact   [B,N,nc]      real-valued contact flags, multiply the fz bounds          (mpc.cpp:344,350)
 *   c_horizon [B,N,3*nc]    world contact-point positions                              (mpc.cpp:229)
 *   p_com     [B,N,3]       CoM used for the lever arms (independent input)            (mpc.cpp:230)
 *   x_current [B,13]|NULL   row 0 of x_opt if given                                    (mpc.cpp:430-433)
 *   u_opt     [B,N,3*nc]    ground-reaction forces, step-major decision vector         (mpc.cpp:417-420)
 *   x_opt     [B,N,13]      in/out: rows that the reference leaves untouched are left untouched
 *                           (row 0 without x_current; rows 2.. with one_rollout)       (mpc.cpp:427-446)
 *
 * Every entry point returns 0 on success or a negative g1mpc_error; numerical trouble inside one
 * problem is reported per problem in `status`, never as a call failure (the reference ignores the
 * solver status, mpc.cpp:405,414).  There is NO CPU fallback: without a CUDA device g1mpc_create fails.
 */
#ifndef G1MPC_B200_H
#define G1MPC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define G1MPC_NUM_STATES 13
#define G1MPC_ABI_VERSION 7

typedef enum g1mpc_error {
    G1MPC_OK = 0,
    G1MPC_ERR_INVALID_ARGUMENT = -1,
    G1MPC_ERR_UNSUPPORTED_SHAPE = -2, /* (horizon, num_contacts) exceeds shared memory, or an adaptor's fixed shape   */
    G1MPC_ERR_CUDA = -3,              /* see g1mpc_last_error() */
    G1MPC_ERR_NO_DEVICE = -4
} g1mpc_error;

/* per-problem status written by g1mpc_solve_batch */
typedef enum g1mpc_status {
    G1MPC_SOLVED = 0,          /* complementarity gap <= gap_tol AND dual residual <= dual_tol (1 + |grad f(u_start)|inf) */
    G1MPC_MAX_ITER = 1,        /* iteration limit hit; best iterate returned                            */
    G1MPC_BAD_BOUNDS = 2,      /* some contact had fz_max*c < fz_min*c (negative flag); treated as swing */
    G1MPC_NUMERICAL = 3,       /* NaN/Inf in the inputs or a breakdown of the factorisation             */
    G1MPC_DUAL_RESIDUAL = 4    /* gap converged, dual residual |H u + g + A'y|inf above dual_tol: result returned,
                                  accurate to what resid[1] says (long horizons: accuracy floor of the stage-wise solves) */
} g1mpc_status;

/* Constants the reference hard-codes in its constructors (mpc.hpp:22,136; mpc.cpp:25-27,90-98;
 * mpc.py:12-51) plus the interior-point settings of this implementation. */
typedef struct g1mpc_params {
    int32_t horizon;          /* N   (horizon_length_, mpc.hpp:19)            */
    int32_t num_contacts;     /* nc  (num_contacts_, mpc.cpp:12)              */
    double mu, fz_min, fz_max, dt;
    double mass;              /* robot_mass_ (mpc.hpp:136)                    */
    double gravity;           /* g_ (mpc.cpp:7); only used for the start point */
    double inertia_body[9];   /* row-major (mpc.cpp:25-27 / mpc.py:24-28)     */
    double q_weights[13];     /* diag(Q_weights_) (mpc.cpp:90-96)             */
    double r_weight;          /* diag(R_weights_) entry (mpc.cpp:98)          */
    double cost_scale;        /* 2: H=2(B'QB+R) (mpc.cpp:361,375); 1: Python (mpc.py:301,317) */
    double gap_tol;           /* stop when mean complementarity s'lambda/m <= gap_tol  (<=0: default 1e-10) */
    int32_t max_iter;         /* interior-point iteration limit               (<=0: default 40)     */
    int32_t layout;           /* lane layout of the solve kernel: 0 = chosen per launch (g1mpc_layout_info: latency mode -- two
                                 warps per problem -- up to 2 problems per SM, one problem per warp up to 4 per SM, two per
                                 warp above), 1 = one problem per warp, 2 = two per warp, 3 = latency mode, 4 = latency mode with
                                 two problems per main warp (measured slower than 2, never chosen automatically; horizon 10, 4
                                 contacts only; 2-3 where compiled: horizon <= 16 resp. <= 15, 4 contacts), 5 = two per warp with refill (persistent grid: a
                                 half-warp that has finished takes the next problem while its partner goes on; bench shape only;
                                 measured 5 % slower than 2 on large batches, never chosen automatically).  Results do not depend
                                 on it (bit-identical). */
    double dual_tol;          /* see G1MPC_SOLVED                             (<=0: default 1e-6)   */
    double warm_margin;       /* warm start: newtons a previous solution is pulled inside every constraint row
                                                                              (<=0: default 0.01 mass |g|) */
    double warm_mu0;          /* warm start: initial complementarity s*lambda (<=0: default 0.1)    */
    int32_t workspace;        /* where a problem's working set (stage factors, per-contact state) lives during the solve:
                                 1 = shared memory (everything on chip; needs it to fit 227 KB), 2 = global memory behind
                                 L1/L2 with a persistent grid (any horizon), 3 = hybrid (stage factors and scratch in shared
                                 memory, per-contact state in global memory, persistent grid), 0 = shared where at least four
                                 problems fit an SM, else chosen per launch among the three by batch size (g1mpc_layout_info).
                                 Same arithmetic, same bits.                                                           */
    int32_t reserved;
} g1mpc_params;

typedef struct g1mpc_handle g1mpc_handle;

/* Creates a solver context on CUDA device `device` (replaces MPC::MPC + initMatrices, mpc.cpp:10-151). */
int g1mpc_create(const g1mpc_params *params, int device, g1mpc_handle **out);
void g1mpc_destroy(g1mpc_handle *h);

/*
 * MPC::updateMPC over a batch, all pointers DEVICE memory (mpc.cpp:448-473): linearise + discretise,
 * condensed build (in factored form), friction-cone/contact-schedule constraints, QP solve, rollout --
 * one fused kernel launch, asynchronous on `stream`.
 * Optional outputs (NULL to skip):
 *   status [B] int32, iters [B] int32,
 *   resid  [B,4]  {complementarity gap, ||H u + g + A'y||inf, max primal violation, objective 1/2u'Hu+g'u},
 *   y_opt  [B,5*nc*N] multipliers of the reference's rows l <= Ac u <= u (OSQP sign convention).
 */
int g1mpc_solve_batch(g1mpc_handle *h, int64_t B, const double *x0, const double *x_ref,
                      const double *contact, const double *c_horizon, const double *p_com,
                      const double *x_current, int one_rollout, double *u_opt, double *x_opt,
                      int32_t *status, int32_t *iters, double *resid, double *y_opt, void *stream);

/*
 * g1mpc_solve_batch started from a previous solution -- the warm start the reference asks its solver for
 * (setWarmStart(true), mpc.cpp:386; warm_start=True, mpc.py:342) but never gets, because it constructs a fresh solver on
 * every call (mpc.cpp:380).  u_warm [B,N,3nc] (device) is the u_opt of a neighbouring problem, warm_shift the
 * number of stages the horizon has moved since (1 for consecutive control ticks, 0 for a re-solve); u_warm may alias
 * u_opt.  The minimiser is unique, so the result is the cold-start result to solver accuracy (not bit for bit); a start
 * that stalls is detected after 5 iterations and the problem restarts cold.  u_warm == NULL: identical to g1mpc_solve_batch.
 */
int g1mpc_solve_batch_warm(g1mpc_handle *h, int64_t B, const double *x0, const double *x_ref,
                           const double *contact, const double *c_horizon, const double *p_com,
                           const double *x_current, int one_rollout, const double *u_warm, int warm_shift,
                           double *u_opt, double *x_opt, int32_t *status, int32_t *iters, double *resid,
                           double *y_opt, void *stream);

/*
 * g1mpc_solve_batch that ALLOCATES NOTHING.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, same results as g1mpc_solve_batch (cold start).  It runs the global-memory-workspace variant of the solve
 * kernel (params.workspace = 2) with the workspace -- one working-set slot per resident warp -- in a file-scope
 * `static __device__` array of G1MPC_STATIC_WS_BYTES that every GPU holding the library keeps until the process exits; the
 * call itself makes no cudaMalloc, cudaFree or allocating library call (the array's device address and the kernel's occupancy
 * are looked up once per device in g1mpc_create).  Limit: one slot per SM must fit, i.e. the working set of one problem may
 * be at most G1MPC_STATIC_WS_BYTES / #SMs (226,719 B on a 148-SM B200: N = 10 / nc = 4 needs 28,096 B and runs at full
 * residency, N = 40 / nc = 8 needs 160,096 B and runs on 209 warps; N = 60 / nc = 8, 240,096 B, is rejected) -- beyond it the call returns
 * G1MPC_ERR_UNSUPPORTED_SHAPE with the sizes in g1mpc_last_error().  Any B runs (persistent grid).  All calls on one GPU share
 * the one array, from whatever handle or stream: calls that OVERLAP on the GPU race -- order them on one stream or with
 * events.  Nothing in it needs clearing between calls (a problem writes each word of its slot before reading it).
 */
#define G1MPC_STATIC_WS_BYTES (32u << 20)
int g1mpc_solve_batch_static_ws(g1mpc_handle *h, int64_t B, const double *x0, const double *x_ref,
                                const double *contact, const double *c_horizon, const double *p_com,
                                const double *x_current, int one_rollout, double *u_opt, double *x_opt,
                                int32_t *status, int32_t *iters, double *resid, double *y_opt, void *stream);

/*
 * MPC::computeRollout (mpc.cpp:427-446 / mpc.py:379-391) of a GIVEN u_opt, device pointers: x_opt[0] = x_current if given,
 * x_opt[i] = A x_opt[i-1] + B_{i-1} u_opt[i-1] for i = 1..N-1 (only i = 1 with one_rollout); A, B_i from x_ref (mean yaw),
 * c_horizon and p_com as in the build (mpc.cpp:153-254).  x_opt [B,N,13] is in/out like the member x_opt_.
 */
int g1mpc_rollout_batch(g1mpc_handle *h, int64_t B, const double *x_ref, const double *c_horizon,
                        const double *p_com, const double *u_opt, const double *x_current, int one_rollout,
                        double *x_opt, void *stream);

/* The same with HOST pointers (copies in, rolls out, copies x_opt back, synchronises): what a host-side computeRollout()
 * wrapper calls (include/g1mpc_mpc.hpp, g1_locomotion_b200/mpc.py). */
int g1mpc_rollout_host(g1mpc_handle *h, int64_t B, const double *x_ref, const double *c_horizon,
                       const double *p_com, const double *u_opt, const double *x_current, int one_rollout,
                       double *x_opt);

/*
 * Materialises the reference's intermediate matrices for a batch (device pointers, any may be NULL):
 *   A [B,13,13] (mpc.cpp:208-211)        Bd  [B,13N,3nc] (mpc.cpp:246-254)   Aqp [B,13N,13]  (mpc.cpp:256-270)
 *   Bqp [B,13N,3ncN] (mpc.cpp:272-291)   H   [B,3ncN,3ncN] (mpc.cpp:359-364) g [B,3ncN]      (mpc.cpp:366-376)
 *   Acon [B,5ncN,3ncN] dense (mpc.cpp:293-326)   l,u [B,5ncN] with IEEE +-inf (mpc.cpp:328-357)
 */
int g1mpc_emit_qp_batch(g1mpc_handle *h, int64_t B, const double *x0, const double *x_ref,
                        const double *contact, const double *c_horizon, const double *p_com, double *A,
                        double *Bd, double *Aqp, double *Bqp, double *H, double *g, double *Acon,
                        double *l, double *u, void *stream);

/*
 * Same as g1mpc_solve_batch but every pointer is HOST memory: copies host->device, solves, copies device->host and
 * synchronises.  x_opt is in/out.  This is the call the reference-facing MPC.update()/updateMPC wrapper makes
 * (g1_mpc/src/mpc.cpp:448-473 is called with Eigen matrices in host memory, mpc_to_wbid_node.cpp:494).
 * Caller buffers that are page-locked (cudaMallocHost / cudaHostRegister) are copied straight from / to where they
 * lie; pageable ones are staged through pinned buffers owned by the handle.
 */
int g1mpc_update_host(g1mpc_handle *h, int64_t B, const double *x0, const double *x_ref,
                      const double *contact, const double *c_horizon, const double *p_com,
                      const double *x_current, int one_rollout, double *u_opt, double *x_opt,
                      int32_t *status, int32_t *iters, double *resid);

/*
 * Host-API warm start (off by default = the reference's behaviour, a cold solve per call): when enabled, every
 * g1mpc_update_host / _begin call on a slot starts from the solution the previous call on that slot left on the device,
 * shifted by `shift` stages, provided the batch size is unchanged (otherwise that call is a cold start).
 */
int g1mpc_set_host_warm_start(g1mpc_handle *h, int enable, int shift);

/*
 * The same call in two halves, for a host that feeds one batch after the other (a fleet of simulated walkers, or
 * several controllers behind one GPU): _begin enqueues copy-in, solve and copy-out of a batch on the slot's streams
 * and returns; _end waits for that slot and (for pageable output buffers) unpacks the results.  Slots
 * 0 .. G1MPC_HOST_SLOTS-1 are independent, so batch i+1 can be on its way to the device while batch i is being solved.
 * All buffers of a slot must stay valid and untouched between _begin and _end.  g1mpc_update_host == _begin + _end on
 * slot 0.
 */
#define G1MPC_HOST_SLOTS 3
int g1mpc_update_host_begin(g1mpc_handle *h, int slot, int64_t B, const double *x0, const double *x_ref,
                            const double *contact, const double *c_horizon, const double *p_com,
                            const double *x_current, int one_rollout, double *u_opt, double *x_opt,
                            int32_t *status, int32_t *iters, double *resid);
int g1mpc_update_host_end(g1mpc_handle *h, int slot);

/*
 * SURVEY 8f-1 -- what the demo node builds before every updateMPC call, for a batch (device pointers):
 * MPCNode::update's reference trajectory (g1_mpc/src/mpc_to_wbid_node.cpp:403-435) and p_com extraction (:461-468),
 * callbackSrbdCurrent's tiling of the current contact points over the horizon (:50-60), and gaitPlanner (:71-320):
 * table window by gait_phase = floor(sim_time/dt) (:82-127,171-187) plus the Raibert footstep overwrite of the
 * swing-foot rows (:205-309, RaibertHeuristic :507-524) read from the previous solve's x_opt.
 *   x0 [B,13], contacts_now [B,12] (4 points x xyz), sim_time [B], x_opt_prev [B,N,13]  ->
 *   x_ref [B,N,13] IN/OUT (the node only writes columns 2,3,4,5,8,9,12; the rest keeps the caller's values),
 *   contact [B,N,4], c_horizon [B,N,12], p_com [B,N,3], landing [B,2] or NULL (landing_position x,y as published, :257-270).
 * Needs num_contacts == 4 and horizon <= 10 (the node's table has 20 rows).
 */
int g1mpc_walk_inputs_batch(g1mpc_handle *h, int64_t B, const double *x0, const double *contacts_now,
                            const double *sim_time, const double *x_opt_prev, double *x_ref, double *contact,
                            double *c_horizon, double *p_com, double *landing, void *stream);

/*
 * SURVEY 8f-2 -- what the WBID layer derives from one solve, WBID.setReference(t, x_opt[1], u_opt0, feet)
 * (g1_mujoco_sim/src/wbid.py:231-297, call site ros_run_simulation.py:409), for a batch (device pointers):
 *   refs [B,42] = R(9, euler_matrix of x_opt[1][0:3]) | base velocity (v, omega) | base acceleration reference
 *   (0,0,0, I_torso^-1 sum r_i x omega) | CoM position, velocity, acceleration reference (sum_forces/mass + g, with
 *   sum_forces = rows of reshape(u_opt0,(3,4)) exactly as coded, wbid.py:290) | wrench-task references (u_opt0).
 *   foot_positions [B,4,3]; wbid_mass = model.getMass() of the whole-body model.
 */
#define G1MPC_WBID_REFS 42
int g1mpc_wbid_refs_batch(g1mpc_handle *h, int64_t B, const double *u_opt, const double *x_opt,
                          const double *foot_positions, double wbid_mass, double *refs, void *stream);

/*
 * SURVEY 8f-3 -- closed loop of B walkers for `ticks` control periods, all on the device (three launches per tick:
 * g1mpc_walk_inputs_batch -> g1mpc_solve_batch with x_current = x0, full rollout -> advance).  The SRBD model is the
 * plant, in the manner of the reference's own SRBD-only loop (g1_mpc/scripts/main.py:222-326, which feeds a rollout row
 * back as the next x0): x0 <- x_opt[1], contact points <- row 1 of the planner's c_horizon, sim_time += dt.
 * In/out: x0 [B,13], contacts_now [B,12], sim_time [B], x_ref [B,N,13], x_opt [B,N,13] (the previous solve; zeros at
 * start), u_opt [B,N,12] (last tick's forces on return).  Optional outputs: traj_x [B,ticks,13] (state at the start of
 * each tick), traj_u [B,ticks,12] (first-step forces applied), status_any [B] (OR of the per-tick solver statuses).
 */
int g1mpc_closed_loop_batch(g1mpc_handle *h, int64_t B, int32_t ticks, double *x0, double *contacts_now,
                            double *sim_time, double *x_ref, double *x_opt, double *u_opt, double *traj_x,
                            double *traj_u, int32_t *status_any, void *stream);

/*
 * SURVEY 8f-1, Python half -- GaitPlanner.plan_gait (g1_mpc/scripts/gait_planner.py:27-62) and its mapping onto the MPC
 * inputs in the SRBD-only loop (g1_mpc/scripts/main.py:259-285), for a batch of planners (device pointers):
 *   planner [B,5] = double_support_time, single_support_time, step_length, step_width, num_steps (a whole number);
 *   current_time [B]  ->  contact [B,N,4], c_horizon [B,N,12] (left heel, left toe, right heel, right toe; z = 0),
 *   horizon step i looks the phase up at t_i = current_time + i dt, the last phase if none contains it; heel / toe sit
 *   foot_offset (0.08 in main.py:261) behind / ahead of the foot centre.  Needs num_contacts == 4.  Bit-exact against the
 *   reference class (same additions in the same order).
 */
int g1mpc_plan_gait_batch(g1mpc_handle *h, int64_t B, const double *planner, const double *current_time,
                          double foot_offset, double *contact, double *c_horizon, void *stream);

/*
 * SURVEY 8f-3 as main.py runs it (g1_mpc/scripts/main.py:200-326), B walkers, `ticks` control periods on the device, four
 * launches per tick: reference schedule of testing_double_support (main.py:177-197; tick > 0) -> g1mpc_plan_gait_batch ->
 * solve -> advance.  State feedback as coded: the solve of tick t > 0 uses x0 = x_opt[2] of the previous rollout
 * (main.py:228) while the rollout restarts from x_opt[1] (main.py:306-310).  warm_start != 0: every tick after the first
 * starts from the previous tick's u_opt shifted by one stage (g1mpc_solve_batch_warm).
 * In/out: x_ref [B,N,13] (main.py:207-209 initialises it: zeros, column 12 = g, column 5 = 1), x_opt [B,N,13], u_opt
 * [B,N,12], current_time [B].  Optional outputs: traj_x [B,ticks,13] (x_opt[0] after each tick's rollout, what main.py
 * records), traj_u [B,ticks,12], status_any [B], iters_sum [B] (interior-point iterations summed over the ticks).
 */
int g1mpc_main_loop_batch(g1mpc_handle *h, int64_t B, int32_t ticks, const double *planner, double total_duration,
                          double foot_offset, int warm_start, double *current_time, double *x_ref, double *x_opt,
                          double *u_opt, double *traj_x, double *traj_u, int32_t *status_any, int32_t *iters_sum,
                          void *stream);

/*
 * g1mpc_closed_loop_batch with the options SURVEY 8f-3 names: warm_start != 0 starts every tick after the first from
 * the previous tick's u_opt shifted by one stage; iters_sum [B] (optional) receives the interior-point iterations summed
 * over the ticks, so that cold and warm runs can be compared.
 */
int g1mpc_closed_loop_batch_ex(g1mpc_handle *h, int64_t B, int32_t ticks, int warm_start, double *x0,
                               double *contacts_now, double *sim_time, double *x_ref, double *x_opt, double *u_opt,
                               double *traj_x, double *traj_u, int32_t *status_any, int32_t *iters_sum, void *stream);

/* Dynamic shared memory (bytes; 0 = the workspace is in global memory) and threads per CTA of the solve kernel for this
 * handle's shape, and how many such CTAs an SM holds. */
int g1mpc_kernel_info(const g1mpc_handle *h, int32_t *smem_bytes, int32_t *threads, int32_t *ctas_per_sm);

/*
 * Lane layout a launch of B problems uses: problems_per_warp = 1, or -- for the shapes it is compiled for and above 4
 * problems per SM -- 2 (a half-warp each: the per-stage 6x6 factorisations, identical in every lane, are then issued
 * once for two problems), or -1 for the latency mode (up to 2 problems per SM: one problem per CTA of TWO warps, the
 * second sweeping behind the first's factorisation), or -2 / -3 for the hybrid workspace with one / two problems per warp
 * (persistent grid; smem_bytes = its on-chip part).  smem_bytes = 0: the global-memory workspace.  Results do not depend on the layout
 * (bit-identical).  smem_bytes / ctas_per_sm are per CTA of that layout.
 */
int g1mpc_layout_info(const g1mpc_handle *h, int64_t B, int32_t *problems_per_warp, int32_t *smem_bytes, int32_t *ctas_per_sm);

/* FP64 FMA throughput microbenchmark (dependent-chain-free DFMA loop on every SM), TFLOP/s. */
int g1mpc_fp64_peak_tflops(int device, double *tflops);

const char *g1mpc_last_error(void);
int g1mpc_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* G1MPC_B200_H */
